#!/bin/bash
# Scaling run on one 8-GPU box: bench.py at N = 8, 4, 2, 1 for both precision arms (driver protocol: --steps 20 --warmup 3),
# the graph-captured DDP baseline at 8 and 1, BASELINE config 3 (1 ps + 7 workers; transport auto = NVSwitch multicast)
# and an in-graph timeline of rank 0 at 8 GPUs.  JSON lines land in gpurun_out/r2_*.json.
run() { # name nproc args...
  local name=$1 n=$2; shift 2
  if [ "$n" = "1" ]; then timeout 200 python bench.py --gpus 1 "$@" 2>gpurun_out/$name.err | tail -1 > gpurun_out/$name.json
  else timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29800 + RANDOM % 100)) bench.py --gpus $n "$@" 2>gpurun_out/$name.err | tail -1 > gpurun_out/$name.json; fi
  python -c "
import json,sys
d=json.load(open(sys.argv[1])); print(sys.argv[2], d.get('impl'), d['dtype'], 'ms/step', round(d['ms_per_step'],4), 'e2e', round(d['e2e']['ms_per_step'],4), d['config'].get('parallelism'))" gpurun_out/$name.json $name || tail -3 gpurun_out/$name.err
}
for n in 8 4 2 1; do for p in tf32 bf16; do run r2_${n}gpu_$p $n --steps 20 --warmup 3 --precision $p; done; done
run r2_8gpu_ddpgraph_tf32 8 --impl ddp --graphs --precision tf32 --steps 20 --warmup 3
run r2_1gpu_ddpgraph_tf32 1 --impl ddp --graphs --precision tf32 --steps 20 --warmup 3
for p in tf32 bf16; do run r2_8gpu_cfg3_$p 8 --steps 20 --warmup 3 --layout dedicated --ps 1 --precision $p; done
TFD_PRECISION=tf32 timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29899 scripts/step_timeline.py 2>&1 | grep -v Warn | tail -17 > gpurun_out/r2_timeline_8gpu_tf32.txt; cat gpurun_out/r2_timeline_8gpu_tf32.txt
