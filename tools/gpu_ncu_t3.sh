#!/bin/bash
# ncu --set full capture of the T3 decode launch (256 tokens, as in bench.py) + raw metric export
cd "$GRAFT_REPO_ROOT"
timeout 300 python tools/t3_once.py 256 > gpurun_out/t3_once.log 2>&1 || { echo "t3_once failed"; tail -5 gpurun_out/t3_once.log; exit 1; }
timeout 1500 ncu --set full --import-source on --clock-control none -k regex:t3_megakernel --launch-skip 0 --launch-count 1 \
  -o gpurun_out/prof_t3 -f python tools/t3_once.py 256 > gpurun_out/ncu_t3.log 2>&1
tail -3 gpurun_out/ncu_t3.log
ncu -i gpurun_out/prof_t3.ncu-rep --page raw --csv > gpurun_out/prof_t3_raw.csv 2>/dev/null
ls -la gpurun_out/prof_t3*
