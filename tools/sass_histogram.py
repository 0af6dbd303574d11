"""Per-kernel SASS opcode histogram of the shipped library (cuobjdump -sass): the Blackwell markers (UTC*MMA = tcgen05.mma,
LDTM / STTM = tcgen05.ld / st, UTMALDG / UBLKCP = TMA, HMMA = legacy mma.sync, FFMA2 = packed fp32) and the top opcodes.
    python tools/sass_histogram.py [lib.so] > profiles/rN_sass_histogram.txt"""
import collections
import os
import re
import subprocess
import sys

lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                                         "gpflow_b200", "lib", "libgpflow_b200.so")
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
kern, hist = None, collections.OrderedDict()
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        kern = m.group(1)
        hist[kern] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)", line)
    if m and kern:
        hist[kern][m.group(1)] += 1
demangle = subprocess.run(["c++filt"] + list(hist.keys()), capture_output=True, text=True).stdout.splitlines()
MARK = ["UTCHMMA", "UTCQMMA", "UTCIMMA", "UTCOMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "UTCBAR", "SYNCS", "HMMA", "FFMA2", "FFMA",
        "DFMA", "MUFU", "RED", "REDG", "ATOMG"]
tot = collections.Counter()
print("SASS opcode histogram of %s (sm_100a)\n" % os.path.basename(lib))
for (k, h), name in zip(hist.items(), demangle):
    n = sum(h.values())
    marks = "  ".join("%s=%d" % (m, h[m]) for m in MARK if h[m])
    top = ", ".join("%s %d" % kv for kv in h.most_common(8))
    print("%s\n   %d instructions | %s\n   top: %s\n" % (re.sub(r"\(.*", "", name)[:150], n, marks or "-", top))
    tot.update(h)
print("whole library: " + "  ".join("%s=%d" % (m, tot[m]) for m in MARK if tot[m]))
