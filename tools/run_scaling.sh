for n in 2 4 8; do
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/r2_final_bench_n$n.json 2> gpurun_out/r2_final_bench_n$n.err; echo rc=$?
done
python - <<PY
import json
for n in (2,4,8):
    d=json.loads([l for l in open(f"gpurun_out/r2_final_bench_n{n}.json").read().strip().splitlines() if l.startswith("{")][-1])
    print(n, round(d["value"],1), round(d["ms_per_step"],4), "e2e", round(d["e2e"]["value"],1), round(d["e2e"]["ms_per_step"],4), d["parity"]["ok"], d["parity"]["max_rel_l2_over_ranks"], d["train_step"]["fused"])
PY
