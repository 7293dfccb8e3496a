#!/bin/bash
# Developer sweep of wavefront knobs on the two shipped scenes (GPU box). Usage: tools/sweep.sh "VAR=a VAR=b ..." (one run per word)
run() { echo "== $*"; env "$@" python tools/gpu_diag_lanes.py cornell-box 256 2>&1 | grep spp; env "$@" python tools/gpu_diag_lanes.py veach-mis 100 2>&1 | grep spp; }
for cfg in "$@"; do run $cfg; done
