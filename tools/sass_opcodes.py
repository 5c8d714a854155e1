"""Opcode histogram per kernel of libq4k_b200.so (cuobjdump -sass): which SASS instructions each product kernel is made of —
the evidence for "tcgen05 / TMA / IMMA really are in there" (UTCIMMA, UTCHMMA, LDTM, STTM, UTMALDG, UBLKCP, IMMA.16832, IDP)."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "ggml-neon-opt_b200", "lib", "libq4k_b200.so")
out = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
demangle = lambda n: subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip()
KEY = ["UTCIMMA", "UTCHMMA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UBLKCP", "IMMA", "IDP", "SYNCS", "UCGABAR", "LDS", "LDG", "STG", "IMAD", "LOP3", "FFMA"]
kern, hist = None, collections.OrderedDict()
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        kern = demangle(m.group(1)).replace("(anonymous namespace)::", "").replace("q4k::", "")
        kern = re.sub(r"^void ", "", re.sub(r"\((q4k::)?\w*(Params|const|float|int|long|unsigned|signed|char|\*|,| )[^)]*\)$|\(.*\)$", "", kern))
        hist[kern] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
    if m and kern:
        hist[kern][m.group(1)] += 1
print(f"{os.path.basename(so)}: {len(hist)} kernels; columns = instruction counts in the SASS listing (static, not executed)")
print(f"{'kernel':78s} {'total':>6s} " + " ".join(f"{k:>7s}" for k in KEY))
for k, h in hist.items():
    print(f"{k[:78]:78s} {sum(h.values()):6d} " + " ".join(f"{h.get(x, 0):7d}" for x in KEY))
