#!/usr/bin/env bash
# One GPU-box pass over HEAD: full GPU suite, smoke with NVTX ranges on, config-1 and headline bench lines, and the ncu
# launch list of the headline bench (taken only after the same command exited 0 without ncu). Outputs under gpurun_out/.
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/r3_gpu_tests_c.log 2>&1; echo "suite rc=$?"; tail -2 $O/r3_gpu_tests_c.log
GS_NVTX=1 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
python bench.py --config cpu100k --steps 50 --warmup 10 > $O/r3_bench_cpu100k.json 2> $O/r3_bench_cpu100k.err; echo "cpu100k rc=$?"
python -c "import json;d=json.loads(open('$O/r3_bench_cpu100k.json').read().strip().splitlines()[-1]);print(d['ms_per_step'],d['e2e']['ms_per_step'],d.get('config1_step'),d.get('ref_cuda',{}).get('ms_per_step'))"
python bench.py > $O/r3_bench_n1.json 2> $O/r3_bench_n1.err; echo "headline rc=$?"
python -c "import json;d=json.loads(open('$O/r3_bench_n1.json').read().strip().splitlines()[-1]);print(d['ms_per_step'],d['e2e']['ms_per_step'],d['ref_cuda']['ms_per_step'],d['ref_cuda']['e2e']['ms_per_step'],{k:v['ms'] for k,v in d['stages'].items()}, d['clocks'])"
python bench.py --steps 2 --warmup 3 --skip-baselines --no-train-step-leg > /dev/null 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/r3_launches_bench_3M_1080p.csv python bench.py --steps 2 --warmup 3 --skip-baselines --no-train-step-leg > $O/r3_ncu_bench.log 2>&1; echo "ncu rc=$?"
