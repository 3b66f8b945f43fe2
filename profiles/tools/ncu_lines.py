#!/usr/bin/env python
"""Join an `ncu --page source --csv` dump (per-SASS-instruction samples) with `nvdisasm -g`
line info of the same cubin, and print the hottest source lines of one kernel.

usage: ncu_lines.py <report.ncu-rep> <library.so> <kernel-substring> [top]
"""
import csv
import io
import os
import re
import subprocess
import sys
import tempfile
from collections import defaultdict


def sass_lines(so, kernel):
    tmp = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=tmp, check=True, capture_output=True)
    out = []
    for f in os.listdir(tmp):
        if not f.endswith(".cubin"):
            continue
        dis = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, f)], capture_output=True, text=True).stdout
        cur_fn, cur_line, active = None, None, False
        for ln in dis.splitlines():
            m = re.match(r"\s*\.section\s+\.text\.(\S+?),", ln)
            if m:
                # the mangled name carries the identifier's length: "10k_schedule" must not match "14k_schedule_one";
                # "k_schedule<0, 0>" selects the instantiation with those bool arguments ("10k_scheduleILb0ELb0EE")
                base, _, targ = kernel.partition("<")
                args = [a.strip() for a in targ.rstrip(">").split(",")] if targ else []  # bool template arguments: "0, 1"
                want = "%d%s" % (len(base), base) + ("I" + "".join("Lb%sE" % a for a in args) + "E" if args else "")
                active = want in m.group(1) or kernel == m.group(1)
                cur_line = None
                continue
            if not active:
                continue
            m = re.match(r'\s*//## File "([^"]+)", line (\d+)', ln)
            if m:
                cur_line = (os.path.basename(m.group(1)), int(m.group(2)))
                continue
            m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
            if m:
                out.append((int(m.group(1), 16), m.group(2).strip(), cur_line))
        if out:
            break
    return out


def main():
    rep, so, kernel = sys.argv[1], sys.argv[2], sys.argv[3]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "-k", "regex:" + kernel.partition("<")[0] + "$"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[hi]
    ci = {h: i for i, h in enumerate(hdr)}
    body = [r for r in rows[hi + 1:] if len(r) == len(hdr)]
    sass = sass_lines(so, kernel)
    if len(sass) != len(body):
        print(f"warning: {len(sass)} SASS instructions in the cubin vs {len(body)} in the report", file=sys.stderr)
    stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    by_line = defaultdict(lambda: [0, 0, defaultdict(int)])
    tot_s = tot_i = 0
    for k, r in enumerate(body):
        line = sass[k][2] if k < len(sass) else None
        s = int(r[ci["# Samples"]] or 0)
        ie = int(r[ci["Instructions Executed"]] or 0)
        a = by_line[line]
        a[0] += s
        a[1] += ie
        for c in stall_cols:
            v = int(r[ci[c]] or 0)
            if v:
                a[2][c] += v
        tot_s += s
        tot_i += ie
    print(f"kernel {kernel}: {tot_s} samples, {tot_i} warp-instructions executed, {len(body)} SASS instructions")
    for line, (s, ie, st) in sorted(by_line.items(), key=lambda kv: -kv[1][0])[:top]:
        reasons = ", ".join(f"{k[6:]}={v}" for k, v in sorted(st.items(), key=lambda kv: -kv[1])[:3])
        where = f"{line[0]}:{line[1]}" if line else "?"
        print(f"{100 * s / max(tot_s, 1):5.1f}% samples {100 * ie / max(tot_i, 1):5.1f}% inst  {where:28s} {reasons}")


if __name__ == "__main__":
    main()
