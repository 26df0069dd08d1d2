set -x
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511"
timeout 400 $TR bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/r2c_bench_n8.json 2> gpurun_out/r2c_bench_n8.err; echo rc=$?
timeout 300 $TR bench.py --gpus 8 --steps 20 --warmup 5 --config swin > gpurun_out/r2c_bench_swin_n8.json 2> gpurun_out/r2c_bench_swin_n8.err; echo rc=$?
timeout 400 $TR bench.py --gpus 8 --steps 20 --warmup 5 --config large6m --no-train-step-leg > gpurun_out/r2c_bench_large6m_n8.json 2> gpurun_out/r2c_bench_large6m_n8.err; echo rc=$?
tail -c 400 gpurun_out/r2c_bench_n8.json; tail -3 gpurun_out/r2c_bench_swin_n8.err
