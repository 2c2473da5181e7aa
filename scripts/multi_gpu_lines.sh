#!/bin/bash
# N-GPU evidence pass (torchrun, one rank per GPU): bench lines of c2, c3 and c5; every run under its own timeout
N=${1:-8}; TAG=${2:-r02p}
mkdir -p gpurun_out
for cfg in c2 c3 c5; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29520 + RANDOM % 100)) \
    bench.py --gpus $N --config $cfg --no-cpu-baseline > gpurun_out/${TAG}_${cfg}_${N}gpu.json 2> gpurun_out/${TAG}_${cfg}_${N}gpu.err
  echo "$cfg rc=$?"
  python -c "
import json; d=json.load(open('gpurun_out/${TAG}_${cfg}_${N}gpu.json')); print('$cfg', '%.5g'%d['value'], '%.4g ms'%d['ms_per_step'], 'e2e %.5g'%d['e2e']['value'], d.get('single_pass_job',{}).get('value'))"
done
