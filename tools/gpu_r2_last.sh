#!/bin/bash
# last pass of round 2: full GPU suite, smoke, bench (own + reference + config 4), ncu of the estimator kernels of the final tree
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q -s > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
grep -E "passed|failed|Error" gpurun_out/pytest_gpu.log | tail -3
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke.log; tail -2 gpurun_out/smoke.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
timeout 900 python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/bench_reference.json 2> gpurun_out/bench_ref.err
timeout 900 python bench.py --workload config4 --steps 2 --warmup 1 > gpurun_out/bench_config4.json 2> gpurun_out/bench_c4.err; echo "config4 exit $?"
CFM_REPS=1 timeout 300 python tools/cfm_once.py 812 1 > gpurun_out/cfm_once.log 2>&1 &&
CFM_REPS=1 timeout 900 ncu --set full --import-source on --clock-control none -k regex:"s3_gemm_kernel|s3_attn_tc_kernel|s3_ffn_kernel" --launch-skip 560 --launch-count 14 \
  -o gpurun_out/prof_cfm_r02 -f python tools/cfm_once.py 812 1 > gpurun_out/ncu_cfm.log 2>&1
ncu -i gpurun_out/prof_cfm_r02.ncu-rep --page raw --csv > gpurun_out/prof_cfm_r02_raw.csv 2>/dev/null
python - <<'PY'
import json
d=json.load(open("gpurun_out/bench.json")); print(d["value"], d["ms_per_step"], d["e2e"]["value"], d["stages_ms"], d["gpu_launches"], d["roofline"]["frac"], d["cpu_baseline"]["value"], d["cuda_eager_baseline"]["value"], d["clocks"])
r=json.load(open("gpurun_out/bench_reference.json")); print(r["value"], r["ms_per_step"])
c=json.load(open("gpurun_out/bench_config4.json")); print(c["value"], c["ms_per_step"], c["stages_ms"], c["utterances_per_s"])
PY
