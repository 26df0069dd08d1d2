mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2o_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/r2o_smoke.log
(timeout 1200 python -m pytest tests -q -m gpu 2>&1 | tail -15) > gpurun_out/r2o_tests.log 2>&1
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r2o_bench.json 2> gpurun_out/r2o_bench.err
for cfg in lego300k garden1297 large6m; do
  timeout 600 python bench.py --config $cfg --steps 10 --warmup 3 --no-train-step-leg > gpurun_out/r2o_bench_$cfg.json 2> gpurun_out/r2o_bench_$cfg.err
done
tail -12 gpurun_out/r2o_smoke.log; tail -6 gpurun_out/r2o_tests.log; head -c 300 gpurun_out/r2o_bench.json
