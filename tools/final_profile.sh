#!/bin/bash
# Round-end evidence: GPU test suite log and the ncu launch list of the default bench command (run after the same
# command has exited 0 without ncu).
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -2 gpurun_out/pytest_gpu.log
timeout 100 python bench.py --steps 8 --warmup 4 --no-config5 --no-cpu-baseline --profile-iters 2 --settle 0 > /dev/null 2>&1 || exit 1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_final_b64.csv \
    python bench.py --steps 8 --warmup 4 --no-config5 --no-cpu-baseline --profile-iters 2 --settle 0 > gpurun_out/ncu_final.log 2>&1
tail -c 300 gpurun_out/ncu_final.log
