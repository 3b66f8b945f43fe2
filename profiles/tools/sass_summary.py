#!/usr/bin/env python
"""Per-kernel SASS summary of the shipped library (cuobjdump -sass): instruction counts and the mnemonics that show how
each kernel moves data and synchronises.  usage: python profiles/tools/sass_summary.py > profiles/r02/sass_excerpt.txt"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
SO = os.path.join(ROOT, "nano-gpu-scheduler_b200", "libngs_dealer.so")
CLASSES = [
    ("LDG.E.*256 (256-bit global loads)", r"\bLDG\.E\.\S*256"),
    ("UBLKCP (TMA bulk copy global->shared)", r"\bUBLKCP"),
    ("SYNCS (mbarrier)", r"\bSYNCS"),
    ("BAR.SYNC/ARV", r"\bBAR\.(SYNC|ARV)"),
    ("ATOMG.*MAX.64 (grid arg-max word)", r"\bATOMG\S*MAX\S*64"),
    ("REDG (fire-and-forget add)", r"\bREDG"),
    ("REDUX (warp reduce)", r"\bREDUX"),
    ("SHFL", r"\bSHFL"),
    ("VOTE", r"\bVOTE"),
    ("ST.E.*SYS / LD.E.*SYS (peer / host words)", r"\b(ST|LD)\.E\S*SYS"),
    ("DADD/DMUL/MUFU.RCP64 (fp64 Rate)", r"\b(DADD|DMUL|DFMA|MUFU\.RCP64H)"),
    ("CALL (out-of-line routines: exact search, wide rows)", r"\bCALL"),
    ("STL/LDL (stack)", r"\b(STL|LDL)"),
    ("HMMA/UTCMMA/tcgen05 (tensor core)", r"\b(HMMA|UTC\S*MMA|IMMA)"),
]
out = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True).stdout
arch = sorted(set(re.findall(r"arch = (sm_\w+)", out)))
demangle = lambda n: subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip()  # noqa: E731
kern = collections.OrderedDict()
cur = None
for ln in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", ln)
    if m:
        name = demangle(m.group(1))
        name = re.sub(r"\(anonymous namespace\)::", "", name)
        name = re.sub(r"^void ", "", name).split("(")[0]
        cur = kern.setdefault(name, collections.Counter())
        continue
    m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(.*?);", ln)
    if m and cur is not None:
        cur["_n"] += 1
        for label, rx in CLASSES:
            if re.search(rx, m.group(1)):
                cur[label] += 1
print(f"SASS summary of nano-gpu-scheduler_b200/libngs_dealer.so (cuobjdump -sass; round 2), cubin architectures: {', '.join(arch)}")
print("template arguments: k_assume<PLAN, WIDE>, k_scan_tma<WIDE>, k_bind<WIDE>, k_schedule<ALL_WORK, WIDE>, k_schedule_one[_random]<WIDE>;")
print("WIDE = the snapshot holds nodes of 9..12 GPUs (only then launched)\n")
for name in sorted(kern):
    c = kern[name]
    print(f"{name}: {c['_n']} instructions")
    for label, _ in CLASSES:
        if c[label]:
            print(f"    {label:<48} {c[label]}")
