#!/bin/sh
# usage: exp/multi2.sh N [extra bench args]  -> gpurun_out/r02_n$N.json (C4 tile-sharded over N GPUs), key figures on stdout
cd "$(dirname "$0")/.."
N=$1; shift
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
  bench.py --gpus $N --steps 10 --warmup 3 "$@" > gpurun_out/r02_n$N.json 2> gpurun_out/r02_n$N.err
echo "n$N rc=$?"
python -c "
import json
d=json.loads([l for l in open('gpurun_out/r02_n$N.json') if l.startswith('{')][-1])
print('N=$N frame', d['frame_ms'], 'value', d['value'], 'e2e', d['e2e']['value'], d['e2e']['ms_per_step'], 'verified', d.get('gather_equals_nccl_path'), d.get('host_frame_equals_device_frame'))
print(' single_gpu_same_workload', d.get('single_gpu_same_workload'), 'strong_eff', d.get('strong_efficiency_same_workload'))
print(' stage', d['stage_ms'], 'launches', d['gpu_launches'])" || tail -5 gpurun_out/r02_n$N.err
