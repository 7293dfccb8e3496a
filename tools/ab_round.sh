#!/bin/bash
# Developer A/B on the GPU box. Usage: tools/ab_round.sh "ENV=a ENV2=b" "ENV=c" ...  (one configuration per argument; "-" = defaults)
# Output: gpurun_out/ab.log
mkdir -p gpurun_out
{
for cfg in "$@"; do
  [ "$cfg" = "-" ] && cfg="MCPT_NOP=1"
  echo "== $cfg"
  env $cfg MCPT_DIAG=1 python tools/gpu_diag_lanes.py cornell-box 256 2>&1 | grep -v "^\[diag\] pipe"
  env $cfg MCPT_DIAG=1 python tools/gpu_diag_lanes.py veach-mis 100 2>&1 | grep -v "^\[diag\] pipe"
  env $cfg MCPT_DIAG=1 python tools/gpu_diag_lanes.py synthetic-5000000 16 2>&1 | grep -v "^\[diag\] pipe"
done
} > gpurun_out/ab.log 2>&1
tail -5 gpurun_out/ab.log
