"""Per-kernel counts of the SASS mnemonics that show what a kernel is built on:
tcgen05 MMA (UTCHMMA / UTCQMMA ...), tensor-memory traffic (LDTM / STTM), TMA (UTMALDG / UTMASTG / UTMAREDG), mbarrier
(SYNCS), packed FP32 (FFMA2), plain FFMA, HMMA (none expected).  Reads the built library with cuobjdump.
    python tools/sass_summary.py > profiles/sass_summary.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "stereoformer_b200", "libsfb200.so")
KEYS = ["UTCHMMA", "UTCQMMA", "UTCIMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAREDG", "SYNCS", "FFMA2", "FFMA", "HMMA",
        "LDS", "STS", "LDG", "STG", "ATOM", "RED", "MEMBAR", "MUFU"]
out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
counts, name = collections.OrderedDict(), None
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip() or m.group(1)
        name = re.sub(r"\(.*", "", name)
        counts[name] = collections.Counter()
        continue
    if name is None:
        continue
    m = re.search(r"^\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
    if m:
        op = m.group(1)
        counts[name]["total"] += 1
        for k in KEYS:
            if op == k or (k in ("UTCHMMA", "UTCQMMA", "UTCIMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAREDG") and op.startswith(k)):
                counts[name][k] += 1
                break
print("# %s  (cuobjdump -sass, sm_100a)" % os.path.relpath(LIB, ROOT))
print("%-72s %6s " % ("kernel", "instrs") + " ".join("%8s" % k for k in KEYS))
for n, c in counts.items():
    print("%-72s %6d " % (n[:72], c["total"]) + " ".join("%8d" % c[k] for k in KEYS))
