"""One kernel launch of an .ncu-rep (ncu --set full) -> the JSON summary bench.py reads for its
`roofline` object (profiles/r02_sweep1_ncu.json): DRAM bytes of the launch, issue-slot and pipe
utilisation, executed lane operations per nominal point-camera test.
usage: python tools/ncu_json.py rep.ncu-rep kernel-name-substring n_tests workload-label out.json
       (the LONGEST launch whose demangled name contains the substring is taken: the full-size sweep,
       not the sample pass)"""
import csv
import io
import json
import subprocess
import sys

rep, kre, n_tests, label, out = sys.argv[1], sys.argv[2], float(sys.argv[3]), sys.argv[4], sys.argv[5]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
idx = {h: i for i, h in enumerate(hdr)}
rows = rows[:2] + [r for r in rows[2:] if kre in r[idx["Kernel Name"]]]
if len(rows) < 3:
    sys.exit("no launch whose name contains %r" % kre)


def val(r, name):
    try:
        return float(r[idx[name]].replace(",", ""))
    except (KeyError, ValueError):
        return None


def to_ms(r):
    v, u = val(r, "gpu__time_duration.sum"), units[idx["gpu__time_duration.sum"]]
    return v / 1e6 if u in ("ns", "nsecond") else v / 1e3 if u in ("us", "usecond") else v


def to_bytes(r, name):
    v, u = val(r, name), units[idx[name]].lower()
    scale = {"byte": 1.0, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(u, 1.0)
    return v * scale if v is not None else None


best = max(rows[2:], key=to_ms)
inst = val(best, "smsp__inst_executed.sum")
lanes = val(best, "smsp__thread_inst_executed_per_inst_executed.ratio")
rd, wr = to_bytes(best, "dram__bytes_read.sum"), to_bytes(best, "dram__bytes_write.sum")
d = {
    "kernel": best[idx["Kernel Name"]][:80], "workload": label, "gpu_time_ms_under_ncu": to_ms(best),
    "dram_bytes_read": rd, "dram_bytes_write": wr, "dram_bytes_per_launch": (rd or 0) + (wr or 0),
    "issue_active_pct": val(best, "smsp__issue_active.avg.pct_of_peak_sustained_active"),
    "fma_pipe_pct": val(best, "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active"),
    "alu_pipe_pct": val(best, "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active"),
    "warp_instructions": inst, "active_lanes_per_instruction": lanes,
    "lane_ops_per_test": (inst * lanes / n_tests) if inst and lanes and n_tests else None,
    "nominal_tests": n_tests,
    "l2_hit_pct": val(best, "lts__t_sector_hit_rate.pct"),
    "dram_throughput_pct": val(best, "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
    "occupancy_pct": val(best, "sm__warps_active.avg.pct_of_peak_sustained_active"),
    "registers": val(best, "launch__registers_per_thread"),
    "source": "ncu --set full --clock-control none, %s (longest launch matching %s)" % (rep.split("/")[-1], kre),
}
with open(out, "w") as f:
    json.dump(d, f, indent=1)
print(json.dumps(d, indent=1))
