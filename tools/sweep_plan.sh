#!/bin/bash
# development aid: sweep the split plan at B=64 (fc1_split,dg_split,c2w_split,c2dg_gx)
for plan in "0,0,0,0" "0,1,0,0" "0,2,0,0" "0,8,0,0" "5,0,0,0" "10,0,0,0" "0,0,16,0" "0,0,32,40" "0,0,32,20" "0,1,32,27"; do
  B200MNIST_PLAN=$plan python bench.py --steps 2000 --warmup 50 --no-cpu-baseline --profile-iters 5 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); k=d['kernel_us']; print('$plan', round(d['ms_per_step']*1e3,2), {a.split(' (')[0][:14]:round(b,1) for a,b in k.items()})"
done
