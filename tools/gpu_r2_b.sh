#!/bin/bash
# round 2, GPU pass B: batched engine tests + timing sweep + config4 bench
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_t3_batch_gpu.py -x -q -s > gpurun_out/pytest_batch.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_batch.log
grep -E "max-abs|equal|passed|failed|Error|error|assert" gpurun_out/pytest_batch.log | tail -40
for u in 1 2 8 16 32 64; do timeout 300 python tools/t3b_time.py $u 2>&1 | tail -1; done | tee gpurun_out/t3b_sweep.log
timeout 900 python bench.py --workload config4 --slots 64 --steps 2 > gpurun_out/bench_c4.log 2> gpurun_out/bench_c4.err; echo "bench c4 exit $?"; cat gpurun_out/bench_c4.log; tail -5 gpurun_out/bench_c4.err
