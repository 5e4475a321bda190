#!/bin/bash
# Kernel-tuning builds of the SAME engine for tools/ab.sh: csrc/libba_engine_<name>.so = default flags + the variant's -D switches.
# usage: tools/build_variants.sh name:"-DX=1 -DY=2" ...      (run here; the .so files travel to the GPU box with the snapshot)
cd "$(dirname "$0")/../3dreconstruction_b200/csrc" || exit 1
for spec in "$@"; do
  name=${spec%%:*}; flags=${spec#*:}
  ( /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC --cudart=shared -shared $flags \
      -o libba_engine_$name.so ba_engine.cu -ldl 2>&1 | grep -E "error|spill" ; echo "built $name ($flags)" ) &
done
wait
