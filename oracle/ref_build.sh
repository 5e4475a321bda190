#!/bin/bash
# oracle/ref_build.sh — builds oracle/_ref/libref_cost.so from the reference's OWN residual functors (see ref_cost_wrapper.cpp).
# Runs only where /root/reference exists (this container; __graft_entry__.build() calls it); outputs go to oracle/_ref/ only, which is
# git-ignored but travels to the GPU box.  Nothing of the reference is copied into tracked files.
set -euo pipefail
cd "$(dirname "$0")"
REF=${REFERENCE_ROOT:-/root/reference}/src
[ -f "$REF/adjuster/BundleAdjuster.h" ] || { echo "reference not present: nothing to build"; exit 0; }
mkdir -p _ref
# the private nested class ReprojectCost of BundleAdjuster.h: from its opening line to the first closing "    };"
awk '/^    class ReprojectCost \{/ {on = 1} on {print} on && /^    \};/ {exit}' "$REF/adjuster/BundleAdjuster.h" \
  | sed 's/^    class ReprojectCost {/class ReprojectCost {/' > _ref/reproject_cost.inc
grep -q "ceres::AngleAxisRotatePoint" _ref/reproject_cost.inc || { echo "ReprojectCost not found in the reference header"; exit 1; }
/usr/bin/g++ -O2 -std=c++20 -fPIC -ffp-contract=off -shared -Ishim -I. -I"$REF" -o _ref/libref_cost.so ref_cost_wrapper.cpp
echo "built oracle/_ref/libref_cost.so from $REF/adjuster/{BundleAdjuster.h:ReprojectCost,SnavelyReprojectionError.h}"
