"""Summarise `ncu -i X.ncu-rep --page source --csv` of ONE kernel: executed instructions by opcode, shared-memory
wavefronts by opcode, and a table of contiguous SASS regions (split where the execution count changes) with their share
of instructions, stall samples and shared-memory wavefronts.  Usage: python tools/ncu_source_summary.py src.csv [min_pct]"""
import csv
import sys
from collections import Counter

rows = list(csv.reader(open(sys.argv[1])))
min_pct = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hdr_i]
col = {h: i for i, h in enumerate(hdr)}
data = [r for r in rows[hdr_i + 1:] if len(r) == len(hdr)]


def num(r, name):
    try:
        return float(r[col[name]])
    except ValueError:
        return 0.0


tot_inst = sum(num(r, "Instructions Executed") for r in data)
tot_samp = sum(num(r, "# Samples") for r in data)
tot_wf = sum(num(r, "L1 Wavefronts Shared") for r in data)
print("SASS instructions %d, executed %.4g, stall samples %d, shared wavefronts %.4g (ideal %.4g)" % (
    len(data), tot_inst, tot_samp, tot_wf, sum(num(r, "L1 Wavefronts Shared Ideal") for r in data)))
ops, wfs = Counter(), Counter()
for r in data:
    src = r[col["Source"]].strip()
    if src.startswith("@"):
        src = src.split(None, 1)[1]
    op = src.split()[0].split(".")[0] if src else "?"
    full = src.split()[0] if src else "?"
    ops[op] += num(r, "Instructions Executed")
    wfs[full] += num(r, "L1 Wavefronts Shared")
print("executed by opcode (% of all):")
print("  " + "  ".join("%s %.1f" % (o, 100 * n / tot_inst) for o, n in ops.most_common(24)))
print("shared-memory wavefronts by opcode (% of all):")
print("  " + "  ".join("%s %.1f" % (o, 100 * n / tot_wf) for o, n in wfs.most_common(10) if n > 0))
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
# regions: maximal runs with the same execution count
regions, start = [], 0
for i in range(1, len(data) + 1):
    if i == len(data) or num(data[i], "Instructions Executed") != num(data[start], "Instructions Executed"):
        regions.append((start, i))
        start = i
print("regions (>= %.1f %% of instructions or samples): [first-last] n exec/instr inst%% samp%% wf%% top stalls" % min_pct)
for a, b in regions:
    seg = data[a:b]
    inst = sum(num(r, "Instructions Executed") for r in seg)
    samp = sum(num(r, "# Samples") for r in seg)
    wf = sum(num(r, "L1 Wavefronts Shared") for r in seg)
    if 100 * inst / tot_inst < min_pct and 100 * samp / tot_samp < min_pct:
        continue
    st = Counter({s[6:]: sum(num(r, s) for r in seg) for s in stalls})
    top = " ".join("%s=%d" % (k, round(100 * v / max(samp, 1))) for k, v in st.most_common(5))
    opc = Counter()
    for r in seg:
        src = r[col["Source"]].strip()
        if src.startswith("@"):
            src = src.split(None, 1)[1]
        opc[src.split()[0].split(".")[0] if src else "?"] += 1
    print("[%4d-%4d] n=%4d exec=%.3g inst%%=%5.1f samp%%=%5.1f wf%%=%5.1f  %s | %s" % (
        a, b - 1, b - a, num(seg[0], "Instructions Executed"), 100 * inst / tot_inst, 100 * samp / tot_samp,
        100 * wf / max(tot_wf, 1), top, " ".join("%s:%d" % kv for kv in opc.most_common(6))))
