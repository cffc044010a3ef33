#!/bin/bash
# Step time with the fc1 tcgen05 GEMMs switched on (threshold 128) vs off, per batch: where should they switch on?
for B in ${SWEEP_BATCHES:-128 192 256 384}; do
  for m in 100000 128; do
    B200MNIST_FC1_TC_MIN=$m timeout 120 python bench.py --batch $B --steps 1000 --warmup 50 --no-cpu-baseline --no-config5 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); k=d['kernel_us']; print('B=$B fc1_tc_min=$m', round(d['ms_per_step']*1e3,1), 'us/step;', {n[:22]:round(v,1) for n,v in k.items() if 'fc' in n})"
  done
done
