#!/bin/bash
# A/B timing of engine builds on the GPU box: tools/ab.sh <shape> <variant> [<variant> ...]   ("" = default build)
SHAPE=$1; shift
for v in "$@"; do
  if [ "$v" = "base" ]; then unset BA_ENGINE_LIB; else export BA_ENGINE_LIB=$PWD/3dreconstruction_b200/csrc/libba_engine_$v.so; fi
  echo "== variant $v"
  python tools/quick_gpu.py $SHAPE max_num_iterations=3 2>&1 | grep -E "plain|K1_|K1c|K2c|K3p|K4|K5|PCG|K7|K8|warm"
done
