#!/bin/bash
# GPU run A of round 2: all GPU tests, smoke, then short bench lines for the four workloads.
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/a_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/a_pytest.log
tail -5 gpurun_out/a_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/a_smoke.log 2>&1; tail -2 gpurun_out/a_smoke.log
for w in cornell-box-1024x1024-spp256 veach-mis-1200x900-spp100 synthetic-5M-1920x1080-spp64; do
  python bench.py --steps 3 --warmup 3 --workload $w --no-cpu-baseline > gpurun_out/a_bench_$w.json 2> gpurun_out/a_bench_$w.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/a_bench_$w.json").read().strip().splitlines()[-1])
    print("$w", round(d["value"],1), "Msamples/s build", round(d["lbvh_build_ms"],3), "ms launches", d["gpu_launches"], "e2e", round(d["e2e"]["value"],1))
except Exception as e:
    print("$w FAILED", e); print(open("gpurun_out/a_bench_$w.err").read()[-1500:])
PY
done
