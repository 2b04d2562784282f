#!/bin/bash
# One GPU-box session (run under gpurun): parity tests, the bench line, an A/B of the sweep's exact
# path, and the ncu captures the profiles/ summaries are made from.  Outputs under gpurun_out/.
# usage: bash tools/gpu_session.sh [tag]
tag=${1:-s1}
out=gpurun_out
mkdir -p $out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $out/${tag}_smi.txt 2>&1
timeout 1200 python -m pytest tests -m gpu -q -p no:cacheprovider > $out/${tag}_tests.log 2>&1
echo "tests rc=$?" >> $out/${tag}_tests.log
timeout 900 python bench.py --steps 5 --warmup 3 > $out/${tag}_bench.json 2> $out/${tag}_bench.err
echo "bench rc=$?" >> $out/${tag}_bench.err
PTGS_SWEEP_EXACT=legacy timeout 600 python bench.py --steps 5 --warmup 3 --no-c5 --no-cpu-baseline --secondary '' \
    > $out/${tag}_bench_legacy_exact.json 2> $out/${tag}_bench_legacy_exact.err
timeout 300 python tools/vis_once.py c4 2 > $out/${tag}_vis_once.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:sweep_kernel -c 6 -f \
    -o $out/${tag}_sweep python tools/vis_once.py c4 2 > $out/${tag}_ncu_sweep.log 2>&1
timeout 300 python tools/partition_once.py c4 > $out/${tag}_partition_once.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv \
    --log-file $out/${tag}_launches_partition_c4.csv python tools/partition_once.py c4 > $out/${tag}_ncu_launches.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off \
    -k regex:'vox_|radius_count|gather_sorted|hist_kernel|partition_kernel|expand_|aos_' -c 60 -f \
    -o $out/${tag}_partition python tools/partition_once.py c4 > $out/${tag}_ncu_partition.log 2>&1
tail -5 $out/${tag}_tests.log
head -c 1500 $out/${tag}_bench.json
