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
{
run MCPT_NOP=1
run $LF
run MCPT_NOP=1 MCPT_DIAG=1
run $LF MCPT_DIAG=1
for v in 4 5 6; do run MCPT_WF_PIPES=$v; run $LF MCPT_WF_PIPES=$v; done
run $LF MCPT_WF_PIPES=6 MCPT_TRACE_CTAS_PER_SM=18
run $LF MCPT_NODE_STEPS=3
run $LF MCPT_REFILL_MIN=10
run $LF MCPT_REFILL_MIN=14
} > gpurun_out/c3_ab.log 2>&1
MCPT_LIB=$PWD/montecarlopathtracing_b200/libmcpt_lf.so python -m pytest tests/test_gpu_render.py tests/test_gpu_trace.py -m gpu -q -x > gpurun_out/c3_pytest_lf.log 2>&1; tail -3 gpurun_out/c3_pytest_lf.log
python -m pytest tests/test_gpu_render.py -m gpu -q -x > gpurun_out/c3_pytest.log 2>&1; tail -3 gpurun_out/c3_pytest.log
