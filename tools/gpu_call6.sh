#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x > gpurun_out/c6_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c6_pytest.log
tail -6 gpurun_out/c6_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/c6_smoke.log 2>&1; tail -1 gpurun_out/c6_smoke.log
run() {
  echo "== $*"
  env "$@" python tools/gpu_diag_lanes.py cornell-box 256 2>&1 | grep -v "^\[diag\] pipe"
  env "$@" python tools/gpu_diag_lanes.py veach-mis 100 2>&1 | grep -v "^\[diag\] pipe"
  env "$@" python tools/gpu_diag_lanes.py synthetic-5000000 32 2>&1 | grep -v "^\[diag\] pipe"
  env "$@" python tools/gpu_diag_lanes.py cornell-box 256 3840 2160 2>&1 | grep -v "^\[diag\] pipe"
}
{ run MCPT_NOP=1; run MCPT_NOP=2; } > gpurun_out/c6_ab.log 2>&1
python tools/ab_table.py gpurun_out/c6_ab.log
python tools/chunk_scaling.py > gpurun_out/c6_chunks.log 2>&1; cat gpurun_out/c6_chunks.log
MCPT_WF_MAX_SLOTS=67000000 MCPT_WF_MIN_GROUPS=8 python tools/chunk_scaling.py > gpurun_out/c6_chunks8.log 2>&1; cat gpurun_out/c6_chunks8.log
