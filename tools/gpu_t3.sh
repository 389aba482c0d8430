#!/bin/bash
# T3 iteration pass: GPU parity tests, phase timing, short bench
cd "$GRAFT_REPO_ROOT"
timeout 300 python -m pytest tests/test_t3_gpu.py -x -q --timeout 60 > gpurun_out/pytest_t3.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_t3.log
tail -15 gpurun_out/pytest_t3.log
timeout 300 python tools/t3_phase_times.py > gpurun_out/t3_phases.log 2>&1; tail -20 gpurun_out/t3_phases.log
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_t3.log 2>gpurun_out/bench_t3.err; cat gpurun_out/bench_t3.log; tail -3 gpurun_out/bench_t3.err
