#!/bin/bash
# Developer sweep of the wavefront knobs on the two shipped scenes (GPU box). Usage: tools/sweep.sh
run() { echo "== $*"; env "$@" python tools/gpu_diag_lanes.py cornell-box 256 2>&1 | grep spp; env "$@" python tools/gpu_diag_lanes.py veach-mis 100 2>&1 | grep spp; }
run X=0
for v in 8 16 20; do run MCPT_REFILL_MIN=$v; done
for v in 12 20 24 32; do run MCPT_TRACE_CTAS_PER_SM=$v; done
for v in 2 4; do run MCPT_WF_PIPES=$v; done
for v in 8 16; do run MCPT_LEAF_MIN=$v; done
for v in 2 8; do run MCPT_BLOCK_MIN=$v; done
for v in 2 8; do run MCPT_NODE_STEPS=$v; done
