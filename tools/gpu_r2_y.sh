#!/bin/bash
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_boundary.py tests/test_flow_frontend.py -m gpu -q -s 2>&1 | grep -E "flow_inference_batch|passed|failed|Error|error|assert" | head -30
timeout 900 python bench.py --workload config4 --steps 2 --warmup 1 > gpurun_out/bench_config4.json 2> gpurun_out/bench_c4.err; echo "config4 exit $?"; tail -3 gpurun_out/bench_c4.err
python - <<'PY'
import json
c=json.load(open("gpurun_out/bench_config4.json")); print(c["value"], c["ms_per_step"], c["stages_ms"], c["utterances_per_s"], c.get("gpu_launches"))
PY
