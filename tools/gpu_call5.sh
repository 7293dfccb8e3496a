#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x > gpurun_out/c5_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c5_pytest.log
tail -4 gpurun_out/c5_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/c5_smoke.log 2>&1; tail -1 gpurun_out/c5_smoke.log
run() {
  echo "== $*"
  env "$@" python tools/gpu_diag_lanes.py cornell-box 256 2>&1 | grep -v "^\[diag\] pipe"
  env "$@" python tools/gpu_diag_lanes.py veach-mis 100 2>&1 | grep -v "^\[diag\] pipe"
  env "$@" python tools/gpu_diag_lanes.py synthetic-5000000 32 2>&1 | grep -v "^\[diag\] pipe"
  env "$@" python tools/gpu_diag_lanes.py cornell-box 256 3840 2160 2>&1 | grep -v "^\[diag\] pipe"
}
{ run MCPT_NOP=1; run MCPT_DIAG=1; } > gpurun_out/c5_ab.log 2>&1
python tools/chunk_scaling.py > gpurun_out/c5_chunks.log 2>&1; cat gpurun_out/c5_chunks.log
