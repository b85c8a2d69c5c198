"""Blackwell-native evidence from the built library: per kernel, how many tcgen05 / TMEM / TMA / mbarrier
SASS instructions it holds (B200_PROFILING.md "What proves a Blackwell-native kernel").
usage: python scripts/sass_summary.py [k8s_b200/csrc/libtfgraft.so] > profiles/sass_summary.txt"""
import collections
import re
import subprocess
import sys

so = sys.argv[1] if len(sys.argv) > 1 else "k8s_b200/csrc/libtfgraft.so"
OPS = ["UTCHMMA", "UTCQMMA", "UTCMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAREDG", "UBLKCP", "UTCBAR", "SYNCS", "UTCATOM", "HMMA", "LDGSTS"]
out = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
counts = collections.OrderedDict()
cur = None
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1)
        counts[cur] = collections.Counter()
        continue
    if cur is None:
        continue
    m = re.search(r"^\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d\s+)?([A-Z0-9_.]+)", line)
    if m:
        op = m.group(1)
        for o in OPS:
            if op.startswith(o):
                counts[cur][o + (op[len(o):len(o) + 3] if o == "UTMALDG" else "")] += 1
demangled = subprocess.run(["c++filt"] + list(counts), capture_output=True, text=True).stdout.splitlines()
total = collections.Counter()
print("SASS opcode counts per kernel of %s (kernels without any of the listed opcodes are omitted)" % so)
print("tcgen05.mma = UTC*MMA, tcgen05.ld = LDTM, TMA loads/stores = UTMALDG/UTMASTG/UTMAREDG, bulk copy = UBLKCP,")
print("tcgen05.commit = UTCBAR, mbarrier = SYNCS; HMMA (legacy mma.sync) must be absent.\n")
for (name, c), dm in zip(counts.items(), demangled):
    if not c:
        continue
    total.update(c)
    short = re.sub(r"\(.*", "", dm).replace("void ", "").replace("tfg::", "")[:90]
    print("%-92s %s" % (short, "  ".join("%s=%d" % kv for kv in sorted(c.items()))))
print("\nTOTAL  " + "  ".join("%s=%d" % kv for kv in sorted(total.items())))
