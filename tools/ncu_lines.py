"""Per-source-line totals (instructions executed, stall samples) of one kernel launch in an .ncu-rep.
usage: python tools/ncu_lines.py rep.ncu-rep kernel-base-name-regex [substring of the demangled name, e.g. "<(int)1"]
(ncu's --kernel-name matches the base name without template arguments; the longest matching launch is reported)"""
import csv, io, subprocess, sys, collections
rep, kre = sys.argv[1], sys.argv[2]
sub = sys.argv[3] if len(sys.argv) > 3 else None
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass",
                      "--kernel-name", "regex:" + kre], capture_output=True, text=True).stdout
sections, cur, cols, fpath = [], None, None, ""
for row in csv.reader(io.StringIO(out)):
    if not row:
        continue
    if row[0] == "File Path":
        fpath = row[1].split("/")[-1]
        continue
    if row[0] == "Function Name":
        if cur is None or cur["name"] != row[1] or fpath == "visibility.cu" and any(k[0] == fpath for k in cur["lines"]):
            cur = {"name": row[1], "lines": collections.OrderedDict()}
            sections.append(cur)
        continue
    if row[0] == "Line No":
        cols = {h: i for i, h in enumerate(row)}
        first_src = row.index("Source")
        continue
    if cols is None or cur is None or row[0] == "":
        continue
    try:
        ln = int(row[0])
        ins = int(row[cols["Instructions Executed"]]); smp = int(row[cols["Warp Stall Sampling (All Samples)"]])
        thr = int(row[cols["Thread Instructions Executed"]])
    except (ValueError, KeyError, IndexError):
        continue
    e = cur["lines"].setdefault((fpath, ln), [0, 0, 0, row[first_src].strip()])
    e[0] += ins; e[1] += smp; e[2] += thr
if sub:
    sections = [x for x in sections if sub in x["name"]]
if not sections:
    sys.exit("no sections parsed")
sel = max(sections, key=lambda s: sum(v[0] for v in s["lines"].values()))
tot_i = sum(v[0] for v in sel["lines"].values()) or 1
tot_s = sum(v[1] for v in sel["lines"].values()) or 1
print("kernel %s: %d warp instructions, %d samples (per source line incl. inlined callees)" % (sel["name"][:70], tot_i, tot_s))
for ln, v in sorted(sel["lines"].items(), key=lambda kv: -kv[1][0])[:45]:
    print("%s:%-5d  %5.1f%% inst  %5.1f%% smp  thr/inst %4.1f  %s" % (ln[0][:12], ln[1], 100.0 * v[0] / tot_i, 100.0 * v[1] / tot_s,
                                                               v[2] / max(v[0], 1), v[3][:100]))
