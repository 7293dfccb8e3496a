#!/bin/bash
# Final single-GPU evidence run of round 2.
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/z_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/z_pytest.log
tail -4 gpurun_out/z_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/z_smoke.log 2>&1; tail -1 gpurun_out/z_smoke.log
# per-kernel instruction / DRAM counters of whole renders -> profiles/kernel_metrics.json
MET="smsp__inst_executed.sum,smsp__thread_inst_executed.sum,dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum"
for spec in cornell-box-1024x1024:64 veach-mis-1200x900:25 synthetic-5M-1920x1080:16 cornell-box-3840x2160:16; do
  w=${spec%%:*}; s=${spec##*:}
  timeout 900 ncu --metrics $MET --clock-control none --csv --log-file gpurun_out/prof_$w.csv python tools/profile_render.py --workload $w --spp $s > gpurun_out/prof_$w.json 2> gpurun_out/prof_$w.err
  echo "prof $w rc=$? $(wc -l < gpurun_out/prof_$w.csv) rows"
done
# --set full: every kernel once or twice (two scenes), then mid-render launches of the two wavefront kernels on three workloads, the megakernel
for sc in cornell-box synthetic-1M; do
  timeout 900 ncu --set full --import-source on --clock-control none -k regex:'^(?!.*(wf_trace|wf_shade)).*$' -f -o gpurun_out/r2_all_$sc python tools/profile_all_kernels.py --scene $sc > gpurun_out/r2_all_$sc.log 2>&1; echo "all $sc rc=$?"
done
for spec in cornell-box-1024x1024:256:120 veach-mis-1200x900:100:60 synthetic-5M-1920x1080:64:60; do
  IFS=: read w s skip <<< "$spec"
  for k in wf_trace wf_shade; do
    timeout 600 ncu --set full --import-source on --clock-control none --kernel-name regex:$k --launch-skip $skip --launch-count 1 -f -o gpurun_out/r2_${k}_$w python tools/profile_render.py --workload $w --spp $s > gpurun_out/r2_${k}_$w.log 2>&1; echo "full $k $w rc=$?"
  done
done
timeout 600 ncu --set full --import-source on --clock-control none --kernel-name regex:pt_megakernel --launch-count 1 -f -o gpurun_out/r2_megakernel python tools/profile_render.py --workload cornell-box-1024x1024 --spp 16 --mode megakernel > gpurun_out/r2_megakernel.log 2>&1; echo "full megakernel rc=$?"
# the two shared-memory variants of wf_trace (A/B builds): counters for why they lose
for lib in sstack stop; do
  MCPT_LIB=$PWD/montecarlopathtracing_b200/libmcpt_$lib.so timeout 600 ncu --set full --clock-control none --kernel-name regex:wf_trace --launch-skip 120 --launch-count 1 -f -o gpurun_out/r2_wf_trace_$lib python tools/profile_render.py --workload cornell-box-1024x1024 --spp 256 > gpurun_out/r2_wf_trace_$lib.log 2>&1; echo "full wf_trace $lib rc=$?"
done
ls -la gpurun_out/*.ncu-rep | awk '{print $5, $9}'
