"""Writes profiles/sass_summary.txt: per kernel, the counts of the SASS mnemonics that prove the Blackwell-native paths
(UTC*MMA = tcgen05.mma, LDTM/STTM = tcgen05.ld/st, UTMALDG/UTMASTG = TMA, SYNCS = mbarrier, multimem / RED / ATOM)."""
import collections
import re
import subprocess
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
lib = ROOT / "hyperion_b200/lib/libhyperion_b200.so"
out = subprocess.run(["cuobjdump", "-sass", str(lib)], capture_output=True, text=True).stdout
pat = re.compile(r"\b(UTCHMMA|UTCQMMA|UTCMMA|UTCBAR|UTCATOMSWS|LDTM|STTM|UTMALDG|UTMASTG|UTMAPF|UBLKCP|SYNCS|HMMA|LDGSTS|RED|ATOMG|ATOM|MULTIMEM|LDGMC|REDG|ST\.E\.MULTIMEM|MEMBAR|MUFU)\S*")
kern = None
counts = collections.OrderedDict()
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        kern = m.group(1)
        counts[kern] = collections.Counter()
        continue
    if kern:
        for mm in pat.finditer(line):
            counts[kern][mm.group(1)] += 1
demangle = subprocess.run(["c++filt"], input="\n".join(counts), capture_output=True, text=True).stdout.splitlines()
lines = ["# SASS mnemonic counts per kernel of libhyperion_b200.so (cuobjdump -sass, sm_100a)",
         "# UTCHMMA = tcgen05.mma kind::f16/tf32, LDTM = tcgen05.ld, UTMALDG = cp.async.bulk.tensor (TMA), SYNCS = mbarrier", ""]
for (k, c), d in zip(counts.items(), demangle):
    if not c:
        continue
    name = re.sub(r"\(.*", "", d)[:110]
    lines.append(f"{name}: " + ", ".join(f"{m}={n}" for m, n in sorted(c.items())))
(ROOT / "profiles").mkdir(exist_ok=True)
(ROOT / "profiles/sass_summary.txt").write_text("\n".join(lines) + "\n")
print("\n".join(lines[:40]))
