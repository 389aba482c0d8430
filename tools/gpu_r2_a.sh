#!/bin/bash
# round 2, GPU pass A: full GPU test suite (new bench-shape parity tests), T3 A/B variants, bench (both arms),
# ncu --set full of the CFM GEMM / attention kernels
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -s > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
grep -E "rel-L2|max-abs|agree|passed|failed|Error|error" gpurun_out/pytest_gpu.log | tail -60
for v in base early maxs2 maxs1 early_maxs2; do
  echo -n "variant $v: "; timeout 200 python tools/t3_time.py _variants/libcbx_$v.so 2>&1 | tail -1
done | tee gpurun_out/t3_variants.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench exit $?"; cat gpurun_out/bench.log; tail -5 gpurun_out/bench.err
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.log 2> gpurun_out/bench_ref.err; cat gpurun_out/bench_ref.log; tail -3 gpurun_out/bench_ref.err
timeout 300 python tools/cfm_once.py 812 1 > gpurun_out/cfm_once.log 2>&1 &&
timeout 1200 ncu --set full --import-source on --clock-control none -k regex:"s3_gemm_kernel|s3_attn_kernel" --launch-skip 700 --launch-count 24 \
  -o gpurun_out/prof_cfm -f python tools/cfm_once.py 812 1 > gpurun_out/ncu_cfm.log 2>&1
tail -3 gpurun_out/ncu_cfm.log
ls -la gpurun_out | head -40
