#!/usr/bin/env python3
"""Per-kernel SASS mnemonic counts of the built library: which kernels use the Blackwell tensor core (UTCQMMA /
UTCHMMA = tcgen05.mma), TMEM loads (LDTM = tcgen05.ld), TMA tensor loads (UTMALDG), bulk copies (UBLKCP),
legacy mma.sync (HMMA / QMMA), cluster / DSMEM traffic and PDL (griddepcontrol).  Usage:
    python scripts/sass_summary.py [vllm_b200/lib/libvllm_b200.so] > profiles/r02_sass_summary.txt
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "vllm_b200", "lib", "libvllm_b200.so")
PAT = [("UTCQMMA", r"\bUTCQMMA"), ("UTCHMMA", r"\bUTCHMMA"), ("UTCQMMA.2CTA", r"\bUTCQMMA\.2CTA"), ("LDTM", r"\bLDTM"),
       ("UTMALDG", r"\bUTMALDG"), ("UTMALDG.2CTA", r"\bUTMALDG\S*\.2CTA"), ("UBLKCP", r"\bUBLKCP"),
       ("HMMA", r"\bHMMA"), ("SYNCS", r"\bSYNCS"), ("UCGABAR", r"\bUCGABAR"), ("ST.cluster", r"\bST\S*\.CLUSTER|STAS"),
       ("ACQBULK/PDL", r"\bACQBULK|\bPREEXIT|griddep")]
txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True, check=True).stdout
demangle = {}
fn, counts, instrs = None, collections.OrderedDict(), {}
for ln in txt.splitlines():
    m = re.match(r"\s*Function : (\S+)", ln)
    if m:
        fn = m.group(1)
        counts[fn] = collections.Counter()
        instrs[fn] = 0
        continue
    if fn is None or "/*" not in ln:
        continue
    body = ln.split("/*")[1] if ln.strip().startswith("/*") else ln
    if not re.search(r"/\*[0-9a-f]{4}\*/", ln):
        continue
    instrs[fn] += 1
    for name, pat in PAT:
        if re.search(pat, ln):
            counts[fn][name] += 1
names = list(counts)
dm = subprocess.run(["cu++filt"] + names, capture_output=True, text=True).stdout.splitlines() if names else []
print(f"# SASS summary of {os.path.relpath(so, ROOT)} (cuobjdump -sass, sm_100a); columns = instruction counts per kernel")
print("# " + "  ".join(n for n, _ in PAT) + "  | total instructions | kernel")
for i, fn in enumerate(names):
    c = counts[fn]
    short = (dm[i] if i < len(dm) else fn)
    short = re.sub(r"\((?:[^()]|\([^()]*\))*\)$", "", short)
    print("  ".join(f"{c[n]:5d}" for n, _ in PAT) + f"  | {instrs[fn]:6d} | {short}")
