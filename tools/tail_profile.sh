#!/bin/bash
# Developer tool: where does a tail round of resolve_kernel spend its cycles?  Builds an instrumented copy of the library
# (-DPS_TAIL_PROFILE, never the product) and runs the settled gravity pile of the default scene with it.
set -e
cd "$(dirname "$0")/.."
export PS_LIB=/tmp/libps_prof.so PS_NVCC_EXTRA="-DPS_TAIL_PROFILE"
python -c "import particle_simulator_b200 as ps; ps.build.build_lib(force=True)"
for prec in f32 f64; do
  PS_RESOLVE_TRACE=1 python tools/profile_collide.py --default-scene --precision $prec --gravity --warm-frames 100 --frames 1 2>&1 | grep -E "tail profile|ms_per_substep" | tail -3
done
