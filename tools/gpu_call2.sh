#!/bin/bash
# knob re-tuning sweep on the round-2 code
mkdir -p gpurun_out
run() {
  echo "== $*"
  env "$@" python tools/gpu_diag_lanes.py cornell-box 256 2>&1 | grep spp
  env "$@" python tools/gpu_diag_lanes.py veach-mis 100 2>&1 | grep spp
  env "$@" python tools/gpu_diag_lanes.py synthetic-5000000 32 2>&1 | grep spp
  env "$@" python tools/gpu_diag_lanes.py cornell-box 256 3840 2160 2>&1 | grep spp
}
{
run MCPT_NOP=1
run MCPT_NOP=2
for v in 2 4; do run MCPT_WF_PIPES=$v; done
for v in 9 15 18 24; do run MCPT_TRACE_CTAS_PER_SM=$v; done
for v in 6 8 10 14 16 20; do run MCPT_REFILL_MIN=$v; done
for v in 1 3; do run MCPT_NODE_STEPS=$v; done
for v in 1500000 2000000 4500000 6000000 9000000; do run MCPT_WF_SLOTS=$v; done
run MCPT_WF_PIPES=4 MCPT_TRACE_CTAS_PER_SM=16
run MCPT_WF_PIPES=2 MCPT_TRACE_CTAS_PER_SM=10
} > gpurun_out/c2_sweep.log 2>&1
tail -30 gpurun_out/c2_sweep.log
