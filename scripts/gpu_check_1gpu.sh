#!/bin/bash
# One-GPU check used during development: kernel + engine tests, in-graph kernel timelines of both precision arms,
# bench.py for both arms and the graph-captured / eager PyTorch baselines.  Run under gpurun from the repo root.
python -m pytest tests/test_gpu_kernels.py tests/test_gpu_engine.py -m gpu -q -x 2>&1 | tail -3
for p in bf16 tf32; do echo "== $p"; TFD_PRECISION=$p python scripts/step_timeline.py 2>&1 | grep -v Warning | tail -15; done
show() { python -c "
import json,sys
d=json.load(open(sys.argv[1])); print(sys.argv[2], d['dtype'], 'ms/step', round(d['ms_per_step'],4), 'hot', round(d.get('ms_per_step_hot_l2',0),4), 'e2e', round(d['e2e']['ms_per_step'],4))" "$1" "$2" || tail -3 "${1%.json}.err"; }
for p in tf32 bf16; do python bench.py --steps 20 --warmup 3 --precision $p 2>gpurun_out/k_b1_$p.err | tail -1 > gpurun_out/k_b1_$p.json; show gpurun_out/k_b1_$p.json ours; done
if [ "$1" = "ddp" ]; then
for p in tf32 bf16; do python bench.py --impl ddp --graphs --precision $p --steps 20 --warmup 3 > gpurun_out/ddpg1_$p.json 2> gpurun_out/ddpg1_$p.err; show gpurun_out/ddpg1_$p.json ddp-graph; done
python bench.py --impl ddp --precision tf32 --steps 20 --warmup 3 2>gpurun_out/ddpe1_tf32.err | tail -1 > gpurun_out/ddpe1_tf32.json; show gpurun_out/ddpe1_tf32.json ddp-eager
fi
