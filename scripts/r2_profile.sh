#!/bin/bash
# Round-2 evidence run (one gpurun call): launch list of the device loop, ncu --set full of the decode kernel, of the
# encoder kernels and of the backward, and the decode kernel's DRAM bytes at every BASELINE shape.  Each ncu run follows a
# plain run of the same command.
set -x
cd "$(dirname "$0")/.."
O=gpurun_out
P=${1:-r2f}
BENCH="python bench.py --steps 2 --warmup 3 --no-extras --no-ppo --e2e-steps 0 --cpu-envs 2"
$BENCH > $O/${P}_plain_bench.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/${P}_launches_c3.csv $BENCH > $O/${P}_ncu_bench.log 2>&1
python scripts/decode_sweep.py 16384 50 > $O/${P}_plain_dec.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:decode_tc_kernel -s 3 -c 1 -o $O/${P}_decode python scripts/decode_sweep.py 16384 50 > $O/${P}_ncu_dec.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"enc_block_tc|enc_embed_qkv_tc|attn_tc_kernel" -s 15 -c 5 -o $O/${P}_enc python scripts/decode_sweep.py 16384 50 > $O/${P}_ncu_enc.log 2>&1
for cfg in "1024 20 1 10" "8192 20" "4096 10" "4096 20" "4096 50" "4096 100" "4096 150" "4096 200"; do
  tag=$(echo $cfg | tr ' ' '_')
  python scripts/decode_sweep.py $cfg > $O/${P}_plain_$tag.log 2>&1 && \
  ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:decode_tc_kernel -s 3 -c 1 --csv --log-file $O/${P}_dram_$tag.csv python scripts/decode_sweep.py $cfg > /dev/null 2>&1
done
python scripts/ppo_time.py 12800 21 > $O/${P}_plain_ppo.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -s 150 -c 160 --csv --log-file $O/${P}_bwd_launches.csv python scripts/ppo_time.py 12800 21 > /dev/null 2>&1
ls -la $O | tail -30
