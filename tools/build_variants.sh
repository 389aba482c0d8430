#!/bin/bash
# A/B builds of libcbx.so with extra -D switches: tools/build_variants.sh name1:"-DX -DY" name2:"..."  -> _variants/libcbx_<name>.so
cd "$(dirname "$0")/.."
mkdir -p _variants
for v in "$@"; do
  name=${v%%:*}; fl=${v#*:}
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -shared -Xcompiler -fPIC -Xptxas -v $fl \
    -o _variants/libcbx_$name.so chatterbox_b200/csrc/*.cu > _variants/build_$name.log 2>&1 || { echo "$name: BUILD FAILED"; tail -5 _variants/build_$name.log; continue; }
  echo "$name: $(grep -A1 't3_megakernelI13' _variants/build_$name.log | grep -oE 'Used [0-9]+ registers' | head -1) $(grep -B1 -A1 't3_megakernelI13' _variants/build_$name.log | grep -oE '[0-9]+ bytes spill stores' | head -1)"
done
