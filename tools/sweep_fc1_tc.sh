#!/bin/bash
# Step time with the fc1 tcgen05 GEMMs switched on from batch 256 vs the default threshold.
for B in 256 384 512 768; do
  for m in 100000 256; do
    B200MNIST_FC1_TC_MIN=$m timeout 120 python bench.py --batch $B --steps 1000 --warmup 50 --no-cpu-baseline 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); k=d['kernel_us']; print('B=$B fc1_tc_min=$m', round(d['ms_per_step']*1e3,1), 'us/step;', {n[:22]:round(v,1) for n,v in k.items() if 'fc' in n})"
  done
done
