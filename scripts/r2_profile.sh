#!/bin/bash
# Round-2 evidence run (one gpurun call): launch lists, ncu --set full of the decode kernel and of the backward, and the
# decode kernel's DRAM bytes at every BASELINE shape.  Each ncu run follows a plain run of the same command.
set -x
cd "$(dirname "$0")/.."
O=gpurun_out
BENCH="python bench.py --steps 2 --warmup 3 --no-extras --no-ppo --e2e-steps 1 --cpu-envs 2"
$BENCH > $O/r2_plain_bench.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r2_launches_c3.csv $BENCH > $O/r2_ncu_bench.log 2>&1
python scripts/decode_sweep.py 16384 50 > $O/r2_plain_dec.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:decode_tc_kernel -s 3 -c 1 -o $O/r2_decode python scripts/decode_sweep.py 16384 50 > $O/r2_ncu_dec.log 2>&1
for cfg in "1024 20 1 10" "8192 20" "4096 10" "4096 20" "4096 50" "4096 100" "4096 150" "4096 200"; do
  tag=$(echo $cfg | tr ' ' '_')
  python scripts/decode_sweep.py $cfg > $O/r2_plain_$tag.log 2>&1 && \
  ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:decode_tc_kernel -s 3 -c 1 --csv --log-file $O/r2_dram_$tag.csv python scripts/decode_sweep.py $cfg > /dev/null 2>&1
done
python scripts/ppo_time.py 12800 21 > $O/r2_plain_ppo.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -s 150 -c 160 --csv --log-file $O/r2_bwd_launches.csv python scripts/ppo_time.py 12800 21 > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:"bwd_dw_kernel|attn_bwd_env_kernel|bwd_mlp_tc_kernel" -s 20 -c 6 -o $O/r2_bwd python scripts/ppo_time.py 12800 21 > $O/r2_ncu_bwd.log 2>&1
ls -la $O | tail -30
