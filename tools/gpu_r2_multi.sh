#!/bin/bash
# Multi-GPU validation (run with gpurun --gpus N): C-ABI multi-GPU tests, the strong-scaling bench line at 1..N GPUs with the image hash.
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi -L
python -m pytest tests/test_gpu_multi.py -m gpu -q > gpurun_out/m_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/m_pytest.log
tail -6 gpurun_out/m_pytest.log
python bench.py --gpus 1 --steps 2 --warmup 1 --no-sub-records --no-cpu-baseline > gpurun_out/m_bench_1.json 2> gpurun_out/m_bench_1.err; echo "bench 1 rc=$?"
for g in 2 4 8; do
  if [ $g -le $N ]; then
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $g --master-addr 127.0.0.1 --master-port $((29500+g)) bench.py --gpus $g --steps 2 --warmup 1 > gpurun_out/m_bench_$g.json 2> gpurun_out/m_bench_$g.err; echo "bench $g rc=$?"
    tail -3 gpurun_out/m_bench_$g.err
  fi
done
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/m_bench_*.json")):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f, "n", d["n_gpus"], "value", round(d["value"],1), "ms/step", round(d["ms_per_step"],1), "exchange_ms", round(d["exchange_ms"],3), "e2e", round(d["e2e"]["value"],1), "sha", d["image_sha256"][:16], "mrays", round(d["mrays_per_s"],1))
    except Exception as e:
        print(f, "FAILED", e)
PY
