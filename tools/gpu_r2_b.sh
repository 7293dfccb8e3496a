#!/bin/bash
# GPU run B of round 2: all GPU tests, the default bench line, launch list, per-kernel instruction counters for every workload.
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/b_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/b_pytest.log
tail -8 gpurun_out/b_pytest.log
python bench.py --steps 3 --warmup 3 > gpurun_out/b_bench.json 2> gpurun_out/b_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/b_bench.err
python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/b_bench_ref.json 2> gpurun_out/b_bench_ref.err; echo "ref rc=$?"
M="smsp__inst_executed.sum,smsp__thread_inst_executed.sum,dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum"
for spec in cornell-box-1024x1024:64 veach-mis-1200x900:25 synthetic-5M-1920x1080:16 cornell-box-3840x2160:16; do
  w=${spec%%:*}; s=${spec##*:}
  timeout 900 ncu --metrics $M --clock-control none --csv --log-file gpurun_out/prof_$w.csv python tools/profile_render.py --workload $w --spp $s > gpurun_out/prof_$w.json 2> gpurun_out/prof_$w.err
  echo "prof $w rc=$? $(wc -l < gpurun_out/prof_$w.csv) rows"
done
