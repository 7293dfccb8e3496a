#!/bin/bash
# Round-2 re-entry: validation of HEAD + probes
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x > gpurun_out/c1_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c1_pytest.log
tail -4 gpurun_out/c1_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/c1_smoke.log 2>&1; tail -1 gpurun_out/c1_smoke.log
python tools/chunk_scaling.py > gpurun_out/c1_chunks.log 2>&1; cat gpurun_out/c1_chunks.log
for lib in "" _qs1 _qs2; do
  echo "== lib$lib"
  for s in cornell-box:256 veach-mis:100 synthetic-5000000:16; do
    MCPT_LIB=$PWD/montecarlopathtracing_b200/libmcpt$lib.so python tools/gpu_diag_lanes.py ${s%%:*} ${s##*:} 2>&1 | grep -v "^\[diag\] pipe"
  done
  MCPT_LIB=$PWD/montecarlopathtracing_b200/libmcpt$lib.so python tools/gpu_diag_lanes.py cornell-box 256 3840 2160 2>&1 | grep -v "^\[diag\] pipe"
done > gpurun_out/c1_ab.log 2>&1
cat gpurun_out/c1_ab.log
MCPT_DIAG_ALIVE=1 python tools/gpu_diag_lanes.py cornell-box 256 > gpurun_out/c1_alive_cb.log 2>&1
MCPT_DIAG_ALIVE=1 python tools/gpu_diag_lanes.py cornell-box 256 3840 2160 > gpurun_out/c1_alive_4k.log 2>&1
MCPT_DIAG=1 python tools/gpu_diag_lanes.py cornell-box 256 2>&1 | grep -v "^\[diag\] pipe" > gpurun_out/c1_lanes.log
cat gpurun_out/c1_lanes.log
python bench.py --steps 2 --warmup 3 > gpurun_out/c1_bench.json 2> gpurun_out/c1_bench.err; echo "bench rc=$?"; tail -2 gpurun_out/c1_bench.err
