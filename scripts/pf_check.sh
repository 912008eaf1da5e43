#!/bin/bash
# 2-GPU check of the copy-engine prefetch against the copy-kernel prefetch: cluster tests, then bench at both settings.
N=${1:-2}
timeout 600 python -m pytest tests/test_gpu_cluster.py -x -q -m gpu 2>&1 | tail -3
for pf in ce kernel; do for p in tf32 bf16; do
  TFD_PREFETCH=$pf timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29800 + RANDOM % 100)) bench.py --gpus $N --steps 20 --warmup 3 --precision $p 2>gpurun_out/pf_${pf}_$p.err | tail -1 > gpurun_out/pf_${N}gpu_${pf}_$p.json
  python -c "
import json,sys
d=json.load(open(sys.argv[1])); print(sys.argv[1], d['dtype'], 'ms/step', round(d['ms_per_step'],4), 'e2e', round(d['e2e']['ms_per_step'],4))" gpurun_out/pf_${N}gpu_${pf}_$p.json || tail -5 gpurun_out/pf_${pf}_$p.err
done; done
TFD_PRECISION=tf32 timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29899 scripts/step_timeline.py 2>&1 | grep -v Warn | tail -17 | tee gpurun_out/pf_timeline_${N}gpu_tf32.txt
