"""SASS census of libgsplat_b200.so per kernel: counts of the mnemonics that prove which hardware path a kernel uses
(B200_PROFILING.md): UTCHMMA = tcgen05.mma, LDTM = tcgen05.ld (TMEM -> registers), UTCBAR = tcgen05.commit, LDGMC = multimem.ld_reduce (NVLS; multimem.st is
an STG.E.128.STRONG.SYS on the multicast address), MATCH.ANY = warp match (stable radix ranking), REDG / ATOMG = global reductions / atomics.

    python profiles/sass_census.py [lib.so] > profiles/r2/r2_sass_census.json      (runs on the build container: cuobjdump only)"""
import collections
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "3dgs-avatar-release_b200", "lib", "libgsplat_b200.so")
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
PATS = {"UTCHMMA": r"\bUTCHMMA", "UTCBAR": r"\bUTCBAR", "LDTM": r"\bLDTM", "UTCATOMSWS": r"UTCATOMSWS", "LDGMC": r"\bLDGMC",
        "STG.128.SYS": r"STG\.E\.128\.STRONG\.SYS", "MATCH.ANY": r"MATCH\.ANY", "REDG": r"\bREDG\.", "ATOMG": r"\bATOMG\.", "ATOMS": r"\bATOMS\.",
        "MUFU.EX2": r"MUFU\.EX2", "MUFU.RCP": r"MUFU\.RCP", "FFMA": r"\bFFMA", "SHFL": r"\bSHFL", "BAR.SYNC": r"\bBAR\.", "ACQBULK/UBLKCP": r"UBLKCP|UTMALDG",
        "instructions": r"^\s+/\*[0-9a-f]{4,5}\*/"}
cnt, cur = collections.defaultdict(collections.Counter), None
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1)
        continue
    if cur:
        for k, p in PATS.items():
            if re.search(p, line):
                cnt[cur][k] += 1
names = subprocess.run(["c++filt"], input="\n".join(cnt), capture_output=True, text=True).stdout.splitlines()
out = {}
for mangled, name in zip(cnt, names):
    name = name.replace("(anonymous namespace)::", "")
    name = re.sub(r"^void ", "", name)
    name = re.sub(r"\(.*$", "", name)
    out[name] = {k: v for k, v in sorted(cnt[mangled].items())}
json.dump(dict(sorted(out.items())), sys.stdout, indent=1)
