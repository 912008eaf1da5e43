#!/bin/bash
# One-GPU verification: GPU test suite, headline bench (kept as profiles/bench/final1_cnn.json), MLP bench,
# the GEMM shared-memory cap sweep and the performance record.
mkdir -p gpurun_out
timeout 400 python -m pytest tests -m gpu -q --timeout 300 2>&1 | tail -3
python bench.py --steps 200 --warmup 10 > gpurun_out/final1_cnn.json 2> gpurun_out/final1_cnn.err
python bench.py --model mlp --steps 30 --warmup 5 > gpurun_out/final1_mlp.json 2> gpurun_out/final1_mlp.err
for kb in 110; do
  TFD_GEMM_SMEM_KB=$kb python bench.py --steps 200 --warmup 10 > gpurun_out/cap${kb}_cnn.json 2>/dev/null
  TFD_GEMM_SMEM_KB=$kb python bench.py --model mlp --steps 30 --warmup 5 > gpurun_out/cap${kb}_mlp.json 2>/dev/null
done
python scripts/perf_table.py cuda > gpurun_out/performance_b200.txt 2>&1
python - <<'PY'
import glob, json
for f in sorted(glob.glob("gpurun_out/final1_*.json") + glob.glob("gpurun_out/cap*_*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f, "ms/step %.4f  img/s %.0f  e2e %.0f  launches %s clocks %s" % (d["ms_per_step"], d["value"], d["e2e"]["value"], d["gpu_launches"], d["clocks"]))
    except Exception as e:
        print(f, "unreadable:", e)
PY
cat gpurun_out/performance_b200.txt
