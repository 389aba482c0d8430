#!/bin/bash
# timing-only ablations of the one-pass attention (results are garbage): which pipe bounds it
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
timeout 300 python tools/cfm_batch_time.py 16 2>&1 | tail -2 | head -1
for v in a_noex2 a_trunc a_ld1 a_all; do echo "-- $v"; CBX_LIBCBX=$PWD/_variants/libcbx_$v.so timeout 300 python tools/cfm_batch_time.py 16 2>&1 | tail -2 | head -1; done
