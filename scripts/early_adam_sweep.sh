#!/bin/bash
# Sweep of the overlapped-Adam launch width (TFD_EARLY_ADAM_BLOCKS) on one GPU; one bench.py JSON line per setting.
mkdir -p gpurun_out
python bench.py --steps 200 --warmup 10 > gpurun_out/ea_off.json 2> gpurun_out/ea_off.err
for n in 37 74 148 222 296; do
  TFD_EARLY_ADAM=1 TFD_EARLY_ADAM_BLOCKS=$n python bench.py --steps 200 --warmup 10 > gpurun_out/ea_$n.json 2> gpurun_out/ea_$n.err
done
python - <<'PY'
import glob, json
for f in sorted(glob.glob("gpurun_out/ea_*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f, "ms/step %.4f  img/s %.0f  e2e %.0f  sm_mhz %s" % (d["ms_per_step"], d["value"], d["e2e"]["value"], d["clocks"]["sm_mhz"]))
    except Exception as e:
        print(f, "unreadable:", e)
PY
