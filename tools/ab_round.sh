#!/bin/bash
# Developer A/B on the GPU box. Usage: tools/ab_round.sh "ENV=a ENV2=b" "ENV=c" ...  (one configuration per argument; "-" = defaults)
# AB_DIAG=1 adds the lane diagnostics, AB_SCENES picks the workloads.  Output: gpurun_out/ab.log
mkdir -p gpurun_out
SC=${AB_SCENES:-"cornell-box:256 veach-mis:100 synthetic-5000000:16"}
{
for cfg in "$@"; do
  [ "$cfg" = "-" ] && cfg="MCPT_NOP=1"
  echo "== $cfg"
  for s in $SC; do
    env $cfg MCPT_DIAG=$AB_DIAG python tools/gpu_diag_lanes.py ${s%%:*} ${s##*:} 2>&1 | grep -v "^\[diag\] pipe"
  done
done
} > gpurun_out/ab.log 2>&1
tail -5 gpurun_out/ab.log
