#!/bin/bash
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_cfm_gpu.py tests/test_boundary.py tests/test_flow_frontend.py -m gpu -q -s > gpurun_out/pytest_cfm.log 2>&1; echo "pytest exit $?"
grep -E "batched solve|passed|failed|rror" gpurun_out/pytest_cfm.log | tail -24
for nb in 1 2 4 8 16 32 64; do timeout 300 python tools/cfm_batch_time.py $nb 2>&1 | tail -1; done | tee gpurun_out/cfm_batch_sweep.log
for sb in 16 64; do timeout 900 python bench.py --workload config4 --slots 64 --s3-batch $sb --steps 2 > gpurun_out/bench_c4_sb$sb.log 2> gpurun_out/bench_c4.err; echo "bench c4 sb=$sb exit $?"; python -c "
import json,sys; d=json.loads(open('gpurun_out/bench_c4_sb$sb.log').read()); print({k:d[k] for k in ('value','ms_per_step','stages_ms','utterances_per_s','rtf')})"; tail -3 gpurun_out/bench_c4.err; done
