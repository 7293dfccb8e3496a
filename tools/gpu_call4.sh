#!/bin/bash
mkdir -p gpurun_out
run() {
  echo "== $*"
  env "$@" python tools/gpu_diag_lanes.py cornell-box 256 2>&1 | grep -v "^\[diag\] pipe"
  env "$@" python tools/gpu_diag_lanes.py veach-mis 100 2>&1 | grep -v "^\[diag\] pipe"
  env "$@" python tools/gpu_diag_lanes.py synthetic-5000000 32 2>&1 | grep -v "^\[diag\] pipe"
  env "$@" python tools/gpu_diag_lanes.py cornell-box 256 3840 2160 2>&1 | grep -v "^\[diag\] pipe"
}
LF=MCPT_LIB=$PWD/montecarlopathtracing_b200/libmcpt_lf.so
LF2=MCPT_LIB=$PWD/montecarlopathtracing_b200/libmcpt_lf2.so
{
run MCPT_WF_PIPES=4
run $LF MCPT_WF_PIPES=4
run $LF2 MCPT_WF_PIPES=4
run $LF2 MCPT_WF_PIPES=4 MCPT_DIAG=1
run $LF2 MCPT_WF_PIPES=4 MCPT_NODE_STEPS=3
run $LF2 MCPT_WF_PIPES=4 MCPT_NODE_STEPS=1
run $LF2 MCPT_WF_PIPES=4 MCPT_REFILL_MIN=10
run $LF2 MCPT_WF_PIPES=4 MCPT_TRACE_CTAS_PER_SM=16
run $LF2 MCPT_WF_PIPES=3
} > gpurun_out/c4_ab.log 2>&1
MCPT_LIB=$PWD/montecarlopathtracing_b200/libmcpt_lf2.so python -m pytest tests/test_gpu_render.py tests/test_gpu_trace.py -m gpu -q -x > gpurun_out/c4_pytest_lf2.log 2>&1; tail -3 gpurun_out/c4_pytest_lf2.log
