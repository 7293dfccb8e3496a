#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/i_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/i_pytest.log
tail -6 gpurun_out/i_pytest.log
q() { python tools/profile_render.py --workload $1 --spp $2 --renders 3 --builds 5 | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$3', d['workload'], d['spp'], 'render_ms', round(d['render_ms'],2), 'Msamples/s', round(d['samples']/3/d['render_ms']*1e-3,1), 'build_ms', round(d['build_ms'],3), 'trace launches', d['trace_launches'])"; }
for spec in cornell-box-1024x1024:256 veach-mis-1200x900:100 synthetic-5M-1920x1080:64 cornell-box-3840x2160:256; do
  w=${spec%%:*}; s=${spec##*:}
  q $w $s now
done
