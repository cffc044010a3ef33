export B200MNIST_CONV2_WGRAD_TC=1 B200MNIST_CONV2_DGRAD_TC=1 B200MNIST_CONV2_TC=1
timeout 100 python bench.py --steps 20 --warmup 5 --no-graph > /dev/null 2>&1 || timeout 100 python bench.py --steps 20 --warmup 5 > /dev/null 2>&1
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_tc_b64.csv python bench.py --steps 20 --warmup 5 > gpurun_out/ncu_tc.log 2>&1
tail -2 gpurun_out/ncu_tc.log
