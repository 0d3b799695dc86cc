#!/bin/sh
# usage: exp/multi.sh N [extra bench args]  -> gpurun_out/multi_nN*.json (C4 tile-sharded over N GPUs)
cd "$(dirname "$0")/.."
N=$1; shift
for e2e in ${E2E_MODES:-host-frame copy}; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
    bench.py --gpus $N --steps 10 --warmup 3 --e2e $e2e "$@" > gpurun_out/multi_n${N}_$e2e.json 2> gpurun_out/multi_n${N}_$e2e.err
  echo "n$N $e2e rc=$?"
  python -c "
import json
d=json.loads([l for l in open('gpurun_out/multi_n${N}_$e2e.json') if l.startswith('{')][-1])
print('$e2e', 'frame', d['frame_ms'], 'value', d['value'], 'e2e', d['e2e']['value'], d['e2e']['ms_per_step'], 'verified', d.get('gather_equals_nccl_path'), d.get('host_frame_equals_device_frame'))" || tail -5 gpurun_out/multi_n${N}_$e2e.err
done
