#!/bin/bash
# GPU run C: tests; A/B of the shadow early-out; ncu --set full (with source) of wf_trace and wf_shade mid-render.
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x > gpurun_out/c_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c_pytest.log
tail -6 gpurun_out/c_pytest.log
q() { python tools/profile_render.py --workload $1 --spp $2 --renders 3 --builds 3 | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$3', d['workload'], d['spp'], 'render_ms', round(d['render_ms'],2), 'Msamples/s', round(d['samples']/3/d['render_ms']*1e-3,1), 'build_ms', round(d['build_ms'],3))"; }
for spec in cornell-box-1024x1024:256 veach-mis-1200x900:100 synthetic-5M-1920x1080:64 cornell-box-3840x2160:256; do
  w=${spec%%:*}; s=${spec##*:}
  q $w $s early-out
  MCPT_NO_SHADOW_EARLY_OUT=1 q $w $s no-early-out
done
for k in wf_trace wf_shade; do
  timeout 600 ncu --set full --import-source on --clock-control none --kernel-name regex:$k --launch-skip 120 --launch-count 1 -f -o gpurun_out/c_$k python tools/profile_render.py --workload cornell-box-1024x1024 --spp 256 > gpurun_out/c_ncu_$k.log 2>&1
  echo "ncu $k rc=$?"
done
timeout 600 ncu --set full --import-source on --clock-control none --kernel-name regex:wf_shade --launch-skip 120 --launch-count 1 -f -o gpurun_out/c_wf_shade_4k python tools/profile_render.py --workload cornell-box-3840x2160 --spp 64 > gpurun_out/c_ncu_shade4k.log 2>&1
echo "ncu shade 4k rc=$?"
ls -la gpurun_out/*.ncu-rep
