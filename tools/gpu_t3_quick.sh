#!/bin/bash
# quick T3 iteration: parity tests + decode time per token (optionally A/B against variant libraries)
cd "$GRAFT_REPO_ROOT"
timeout 300 python -m pytest tests/test_t3_gpu.py -x -q --timeout 90 2>&1 | tail -4
timeout 200 python tools/t3_time.py 2>&1 | tail -2
for v in _variants/*.so; do [ -f "$v" ] && { echo "== $v"; timeout 200 python tools/t3_time.py "$v" 2>&1 | tail -1; }; done
true
