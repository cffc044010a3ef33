#!/bin/bash
# Step time with the tcgen05 kernels forced on / off at mid-size batches (where should they switch on?).
for B in ${SWEEP_BATCHES:-128 256 512}; do
  for tc in 0 1; do
    B200MNIST_CONV2_TC=$tc B200MNIST_CONV2_DGRAD_TC=$tc B200MNIST_CONV2_WGRAD_TC=$tc timeout 120 python bench.py --batch $B --steps 1500 --warmup 50 --no-cpu-baseline 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('B=$B tc=$tc', round(d['ms_per_step']*1e3,1), 'us/step', round(d['value']), 'img/s')"
  done
done
