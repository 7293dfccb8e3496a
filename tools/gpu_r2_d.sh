#!/bin/bash
# GPU run D: tests after the build rework; build + render timing; ncu launch list of the build.
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x > gpurun_out/d_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/d_pytest.log
tail -6 gpurun_out/d_pytest.log
q() { python tools/profile_render.py --workload $1 --spp $2 --renders 3 --builds 5 | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$3', d['workload'], d['spp'], 'render_ms', round(d['render_ms'],2), 'Msamples/s', round(d['samples']/3/d['render_ms']*1e-3,1), 'build_ms', round(d['build_ms'],3), 'build launches', d['build_launches'])"; }
for spec in cornell-box-1024x1024:256 veach-mis-1200x900:100 synthetic-1M-1920x1080:16 synthetic-5M-1920x1080:64 cornell-box-3840x2160:256; do
  w=${spec%%:*}; s=${spec##*:}
  q $w $s now
done
for w in cornell-box-1024x1024 synthetic-5M-1920x1080; do
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name regex:'morton|Radix|radix|refit|emit|ploc|wide|block_sort' --csv --log-file gpurun_out/d_build_$w.csv python tools/profile_render.py --workload $w --spp 1 --builds 3 > /dev/null 2>&1
python tools/ncu_summary.py launches gpurun_out/d_build_$w.csv gpurun_out/d_build_$w.md; cat gpurun_out/d_build_$w.md | tail -16
done
