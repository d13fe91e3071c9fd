#!/bin/bash
# Round-2 launch list of the device loop alone (no e2e slices, no extras, no PPO): a plain run, then the same command under
# ncu's duration-only pass.
set -x
cd "$(dirname "$0")/.."
O=gpurun_out
BENCH="python bench.py --steps 2 --warmup 3 --no-extras --no-ppo --e2e-steps 0 --cpu-envs 2"
$BENCH > $O/r2b_plain_bench.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/r2b_launches_c3.csv $BENCH > $O/r2b_ncu_bench.log 2>&1
tail -c 600 $O/r2b_plain_bench.log
