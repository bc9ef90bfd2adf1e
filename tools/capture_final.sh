#!/bin/bash
# One GPU call that produces everything profiles/r2_* quotes for the library as built: a plain default bench run first (its line is
# the number), then -- each only after the plain run exited 0 -- the serialized launch lists (1 block and 8 blocks) and one
# `--set full` capture of the hoisted-rotation kernels.  usage: gpurun --timeout 1500 -- bash tools/capture_final.sh <tag>
set -u
tag=${1:-r2_final}
out=gpurun_out
mkdir -p $out
python bench.py > $out/${tag}_bench1.log 2> $out/${tag}_bench1.err || { echo "plain bench failed"; tail -5 $out/${tag}_bench1.err; exit 1; }
python tools/bench_summary.py $out/${tag}_bench1.log
# the commands ncu will wrap, plain first: they must exit 0 without it
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-share --ncu-range > $out/${tag}_plain1.log 2>&1 || { echo "plain 1-block run failed"; exit 1; }
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-share --total-blocks 8 --ncu-range > $out/${tag}_plain8.log 2>&1 || { echo "plain 8-block run failed"; exit 1; }
NCU="ncu --clock-control none --profile-from-start off"
$NCU --metrics gpu__time_duration.sum --csv --log-file $out/${tag}_launches_1b.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-share --ncu-range > $out/${tag}_ncu1.log 2>&1
$NCU --metrics gpu__time_duration.sum --csv --log-file $out/${tag}_launches_8b.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-share --total-blocks 8 --ncu-range > $out/${tag}_ncu8.log 2>&1
$NCU --set full --import-source on -k regex:'k_moddown_w|k_ks_p_inner|k_ks_q_finish' -s 12 -c 8 -o $out/${tag}_ks -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-share --ncu-range > $out/${tag}_ncu_full.log 2>&1
$NCU --set full --import-source on -k regex:'k_diag_mac|k_modup_rows|k_rescale_rows|k_rescale_cols|k_tensor' -c 6 -o $out/${tag}_other -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-share --total-blocks 8 --ncu-range > $out/${tag}_ncu_other.log 2>&1
ls -la $out | grep ${tag}
