#!/bin/bash
# A/B timing of engine builds on the GPU box: tools/ab.sh <shape> <variant> [<variant> ...]   ("base" = default build)
# Every run is under its own timeout: a hung kernel costs one minute, not the whole GPU call.
SHAPE=$1; shift
for v in "$@"; do
  if [ "$v" = "base" ]; then unset BA_ENGINE_LIB; else export BA_ENGINE_LIB=$PWD/3dreconstruction_b200/csrc/libba_engine_$v.so; fi
  echo "== variant $v"
  timeout -s KILL ${AB_TIMEOUT:-90} python tools/quick_gpu.py $SHAPE max_num_iterations=${AB_LM:-3} 2>&1 | grep -E "plain|K1_|K1c|K2c|K3p|K4|K5|PCG|K7|K8|warm|rror|Traceback"
  echo "   rc=${PIPESTATUS[0]}"
done
