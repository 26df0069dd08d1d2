mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 400 $TR --master-port 29511 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/r2k_bench_n8.json 2> gpurun_out/r2k_bench_n8.err
timeout 400 $TR --master-port 29512 tools/dp_check.py --P 3000000 --width 1920 --height 1080 --transports symm:dense,symm:records,symm-p2p:records --breakdown --train-steps 6 > gpurun_out/r2k_dp_n8.json 2> gpurun_out/r2k_dp_n8.err
timeout 300 $TR --master-port 29513 bench.py --gpus 8 --config large6m --steps 10 --warmup 3 --no-train-step-leg > gpurun_out/r2k_bench_large6m_n8.json 2> gpurun_out/r2k_bench_large6m_n8.err
timeout 300 $TR --master-port 29514 bench.py --gpus 8 --config garden1297 --steps 10 --warmup 3 --no-train-step-leg > gpurun_out/r2k_bench_garden1297_n8.json 2> gpurun_out/r2k_bench_garden1297_n8.err
timeout 300 $TR --master-port 29515 tools/dp_check.py --mcmc-steps 20 --mcmc-interval 5 --mcmc-P 100003 --skip-exchange > gpurun_out/r2k_mcmc_n8.json 2> gpurun_out/r2k_mcmc_n8.err
head -c 400 gpurun_out/r2k_bench_n8.json; echo; tail -c 300 gpurun_out/r2k_mcmc_n8.json; tail -3 gpurun_out/r2k_bench_n8.err
