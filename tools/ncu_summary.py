"""Summarise an .ncu-rep (read on the CPU box): key raw metrics per kernel and
the hottest SASS instructions by warp-stall samples.
usage: python tools/ncu_summary.py gpurun_out/prof.ncu-rep [kernel-regex]"""
import csv, io, subprocess, sys, collections

rep = sys.argv[1]
kre = sys.argv[2] if len(sys.argv) > 2 else None
METRICS = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
           "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
           "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
           "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
           "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
           "sm__throughput.avg.pct_of_peak_sustained_elapsed",
           "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
           "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
           "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
           "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
           "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
           "dram__bytes_read.sum", "dram__bytes_write.sum",
           "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
           "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum"]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
idx = {h: i for i, h in enumerate(hdr)}
for r in rows[2:]:
    print("== %s" % r[idx["Kernel Name"]][:90])
    for m in METRICS:
        if m in idx:
            print("   %-66s %s %s" % (m, r[idx[m]], units[idx[m]]))
args = ["ncu", "-i", rep, "--page", "source", "--csv"]
if kre:
    args += ["--kernel-name", "regex:" + kre]
src = subprocess.run(args, capture_output=True, text=True).stdout
kern = None
tab = collections.OrderedDict()
for line in csv.reader(io.StringIO(src)):
    if not line:
        continue
    if line[0] == "Kernel Name":
        kern = line[1]
        tab[kern] = []
        continue
    if line[0] == "Address":
        cols = {h: i for i, h in enumerate(line)}
        continue
    if kern is None:
        continue
    try:
        tab[kern].append((int(line[cols["Warp Stall Sampling (All Samples)"]]), int(line[cols["Instructions Executed"]]),
                          line[cols["Source"]].strip(), len(tab[kern])))
    except (ValueError, KeyError):
        pass
for k, ins in tab.items():
    tot = sum(i[0] for i in ins) or 1
    tin = sum(i[1] for i in ins) or 1
    print("\n== hottest SASS of %s (samples %d, warp instructions %d)" % (k[:80], tot, tin))
    ops = collections.Counter()
    for s, n, t, _ in ins:
        ops[t.split()[0] if not t.startswith("@") else t.split()[1]] += n
    print("   instruction mix: " + ", ".join("%s %.1f%%" % (o, 100.0 * c / tin) for o, c in ops.most_common(14)))
    for s, n, t, pos in sorted(ins, key=lambda x: -x[0])[:18]:
        print("   %5.1f%% samples  %5.1f%% insts  #%04d  %s" % (100.0 * s / tot, 100.0 * n / tin, pos, t[:90]))
