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
# summarise on the box (the raw reports exceed what gpurun copies back), then drop them
for f in gpurun_out/r2_wf_trace_*.ncu-rep gpurun_out/r2_wf_shade_*.ncu-rep gpurun_out/r2_megakernel.ncu-rep; do
  b=${f%.ncu-rep}
  python tools/ncu_summary.py metrics $f ${b}_ncu.md
  python tools/ncu_summary.py lines $f ${b}_lines.md 2>/dev/null
done
python tools/ncu_summary.py source gpurun_out/r2_wf_trace_cornell-box-1024x1024.ncu-rep gpurun_out/r2_wf_trace_cornell-box-1024x1024_source.md
python tools/ncu_summary.py metrics gpurun_out/r2_all_cornell-box.ncu-rep gpurun_out/r2_all_cornell-box_ncu.md
python tools/ncu_summary.py metrics gpurun_out/r2_all_synthetic-1M.ncu-rep gpurun_out/r2_all_synthetic-1M_ncu.md
python tools/roofline_table.py gpurun_out/r2_roofline.md gpurun_out/r2_all_cornell-box.ncu-rep:cornell-box gpurun_out/r2_all_synthetic-1M.ncu-rep:synthetic-1M \
  gpurun_out/r2_wf_trace_cornell-box-1024x1024.ncu-rep:cornell-box-mid-render gpurun_out/r2_wf_shade_cornell-box-1024x1024.ncu-rep:cornell-box-mid-render \
  gpurun_out/r2_wf_trace_veach-mis-1200x900.ncu-rep:veach-mis-mid-render gpurun_out/r2_wf_shade_veach-mis-1200x900.ncu-rep:veach-mis-mid-render \
  gpurun_out/r2_wf_trace_synthetic-5M-1920x1080.ncu-rep:synthetic-5M-mid-render gpurun_out/r2_wf_shade_synthetic-5M-1920x1080.ncu-rep:synthetic-5M-mid-render \
  gpurun_out/r2_megakernel.ncu-rep:cornell-box-megakernel gpurun_out/r2_wf_trace_sstack.ncu-rep:variant-smem-stack gpurun_out/r2_wf_trace_stop.ncu-rep:variant-smem-top256
rm -f gpurun_out/*.ncu-rep
# the default bench line and the reference arm, as the driver runs them
python bench.py --steps 3 --warmup 3 > gpurun_out/z_bench.json 2> gpurun_out/z_bench.err; echo "bench rc=$?"; tail -2 gpurun_out/z_bench.err
python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/z_bench_ref.json 2> gpurun_out/z_bench_ref.err; echo "ref rc=$?"
du -sh gpurun_out
