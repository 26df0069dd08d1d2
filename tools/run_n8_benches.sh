# 8-GPU bench lines of BASELINE.json's multi-GPU configs (run under `gpurun --gpus 8`); outputs under gpurun_out/
set -x
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511"
timeout 400 $TR bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/r2_final_bench_n8.json 2> gpurun_out/r2_final_bench_n8.err; echo rc=$?
timeout 300 $TR bench.py --gpus 8 --steps 20 --warmup 5 --config swin > gpurun_out/r2_final_bench_swin_n8.json 2> gpurun_out/r2_final_bench_swin_n8.err; echo rc=$?
timeout 400 $TR bench.py --gpus 8 --steps 20 --warmup 5 --config large6m --no-train-step-leg > gpurun_out/r2_final_bench_large6m_n8.json 2> gpurun_out/r2_final_bench_large6m_n8.err; echo rc=$?
timeout 400 $TR bench.py --gpus 8 --steps 20 --warmup 5 --config garden1297 --no-train-step-leg > gpurun_out/r2_final_bench_garden1297_n8.json 2> gpurun_out/r2_final_bench_garden1297_n8.err; echo rc=$?
TR2="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512"
timeout 300 $TR2 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2_final_bench_n2.json 2> gpurun_out/r2_final_bench_n2.err; echo rc=$?
TR4="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29513"
timeout 300 $TR4 bench.py --gpus 4 --steps 20 --warmup 5 > gpurun_out/r2_final_bench_n4.json 2> gpurun_out/r2_final_bench_n4.err; echo rc=$?
