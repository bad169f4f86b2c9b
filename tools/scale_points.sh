#!/bin/bash
# bench.py at N GPUs for c2 and c3 (no driver figure): tools/scale_points.sh N  -> gpurun_out/scale_N/
N=$1; OUT=gpurun_out/scale_$N; mkdir -p $OUT
port=29530
for cfg in "c2 1000" "c3 50"; do
  set -- $cfg
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port \
    bench.py --gpus $N --config $1 --steps $2 --warmup 5 --cpu-seconds 1 --no-io > $OUT/bench_$1.json 2> $OUT/bench_$1.err
  port=$((port+1))
  python -c "
import json;d=json.loads([l for l in open('$OUT/bench_$1.json') if l.startswith('{')][-1]);e=d['e2e'];print('$1', d['n_gpus'], round(d['value']), [round(x,1) for x in d['regions_ms']], 'e2e', round(e['value']), round(e['frac_of_link'],3), 'bits', round(e['bitpacked']['value']), 'parity', d['parity_checked']['mismatches'])"
done
