#!/bin/bash
# One GPU-box session (run under gpurun): parity tests, the bench line, the stage times with
# and without the hull pre-filter, and the ncu captures the profiles/ summaries are made from.  The .ncu-rep files are
# summarised ON THE BOX (gpurun_out/ is limited to 64 MiB) and removed.  Outputs under gpurun_out/.
# usage: bash tools/gpu_session.sh [tag] [steps: tests bench stages ncu_sweep launches launches_bench ncu_partition]
tag=${1:-s1}; shift
steps=${@:-tests bench stages ncu_sweep launches ncu_partition}
out=gpurun_out
mkdir -p $out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $out/${tag}_smi.txt 2>&1
for s in $steps; do case $s in
tests)
  timeout 1200 python -m pytest tests -m gpu -q -p no:cacheprovider > $out/${tag}_tests.log 2>&1
  echo "tests rc=$?" >> $out/${tag}_tests.log ;;
bench)
  timeout 900 python bench.py --steps 5 --warmup 3 > $out/${tag}_bench.json 2> $out/${tag}_bench.err
  echo "bench rc=$?" >> $out/${tag}_bench.err ;;
stages)
  for m in prefilter noprefilter; do for nt in ${HULL_NT:-256}; do
    PTGS_HULL_THREADS=$nt timeout 300 python tools/vis_stages.py c4 5 $m > $out/${tag}_vis_stages_${m}_$nt.log 2>&1
  done; done ;;
ncu_sweep)
  timeout 300 python tools/vis_once.py c4 2 > $out/${tag}_vis_once.log 2>&1 && \
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:'sweep_kernel|hull_kernel|pf_|cand_runs|sample_pairs' -s 8 -c 8 -f \
      -o /tmp/${tag}_sweep python tools/vis_once.py c4 2 > $out/${tag}_ncu_sweep.log 2>&1
  if [ -f /tmp/${tag}_sweep.ncu-rep ]; then
    tests=$(python -c "import json;print(json.load(open('$out/${tag}_vis_once.log'))['tests'])" 2>/dev/null || echo 0)
    python tools/ncu_summary.py /tmp/${tag}_sweep.ncu-rep > $out/${tag}_ncu_sweep_summary.txt 2>&1
    python tools/ncu_lines.py /tmp/${tag}_sweep.ncu-rep sweep_kernel '<(int)1' > $out/${tag}_ncu_sweep1_lines.txt 2>&1
    python tools/ncu_lines.py /tmp/${tag}_sweep.ncu-rep sweep_kernel '<(int)2' > $out/${tag}_ncu_sweep2_lines.txt 2>&1
    python tools/ncu_lines.py /tmp/${tag}_sweep.ncu-rep hull_kernel > $out/${tag}_ncu_hull_lines.txt 2>&1
    python tools/ncu_lines.py /tmp/${tag}_sweep.ncu-rep cand_runs_kernel > $out/${tag}_ncu_cand_runs_lines.txt 2>&1
    python tools/ncu_json.py /tmp/${tag}_sweep.ncu-rep 'sweep_kernel<1' $tests c4 $out/${tag}_sweep1_ncu.json > /dev/null 2>> $out/${tag}_ncu_sweep.log
  fi ;;
launches)
  timeout 300 python tools/partition_once.py c4 > $out/${tag}_partition_once.log 2>&1 && \
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv \
      --log-file $out/${tag}_launches_partition_c4.csv python tools/partition_once.py c4 > $out/${tag}_ncu_launches.log 2>&1
  python tools/launch_summary.py $out/${tag}_launches_partition_c4.csv > $out/${tag}_launches_partition_c4_summary.txt 2>&1 ;;
launches_bench)
  # the launch list of the bench command itself (the timed steps are the "step" ranges of the summary)
  timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none --csv \
      --log-file $out/${tag}_launches_bench.csv python bench.py --steps 2 --warmup 3 --no-c5 --no-cpu-baseline --secondary '' \
      > $out/${tag}_launches_bench.log 2>&1
  python tools/launch_summary.py $out/${tag}_launches_bench.csv > $out/${tag}_launches_bench_summary.txt 2>&1
  gzip -f $out/${tag}_launches_bench.csv ;;
ncu_c5)
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:'sweep_kernel|cov_cull|order_' -c 12 -f \
      -o /tmp/${tag}_c5 python tools/configs_report.py c5 > $out/${tag}_ncu_c5.log 2>&1
  if [ -f /tmp/${tag}_c5.ncu-rep ]; then
    python tools/ncu_summary.py /tmp/${tag}_c5.ncu-rep > $out/${tag}_ncu_c5_summary.txt 2>&1
    python tools/ncu_lines.py /tmp/${tag}_c5.ncu-rep sweep_kernel '<(int)0' > $out/${tag}_ncu_c5_sweep0_lines.txt 2>&1
  fi ;;
ncu_partition)
  timeout 900 ncu --set full --clock-control none --profile-from-start off \
      -k regex:'vox_|radius_count|gather_sorted|hist_kernel|partition_kernel|expand_|aos_|cand_plan|keep_' -c 80 -f \
      -o /tmp/${tag}_partition python tools/partition_once.py c4 > $out/${tag}_ncu_partition.log 2>&1
  [ -f /tmp/${tag}_partition.ncu-rep ] && python tools/ncu_summary.py /tmp/${tag}_partition.ncu-rep > $out/${tag}_ncu_partition_summary.txt 2>&1 ;;
esac; done
du -sh $out
tail -5 $out/${tag}_tests.log 2>/dev/null
head -c 1200 $out/${tag}_bench.json 2>/dev/null
tail -n 3 $out/${tag}_vis_stages_*.log 2>/dev/null
