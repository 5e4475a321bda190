#!/bin/bash
# A/B timing of engine builds on the GPU box: tools/ab.sh <shape> <variant> [<variant> ...]
#   variant = build[@VAR=value[,VAR=value...]]   build "base" = default library, else csrc/libba_engine_<build>.so; VARs are exported for the run
# Every run is under its own timeout: a hung kernel costs one minute, not the whole GPU call.
SHAPE=$1; shift
for spec in "$@"; do
  v=${spec%%@*}; envs=""; [ "$spec" != "$v" ] && envs=${spec#*@}
  if [ "$v" = "base" ]; then lib=""; else lib=$PWD/3dreconstruction_b200/csrc/libba_engine_$v.so; fi
  echo "== variant $spec"
  ( [ -n "$lib" ] && export BA_ENGINE_LIB=$lib; for kv in ${envs//,/ }; do export "$kv"; done
    timeout -s KILL ${AB_TIMEOUT:-90} python tools/quick_gpu.py $SHAPE max_num_iterations=${AB_LM:-3} 2>&1 | grep -E "plain|K0_|K1_|K1c|K2c|K3p|K3c|K4|K5|PCG|K7|K8|misc|warm|rror|Traceback"; echo "   rc=${PIPESTATUS[0]}" )
done
