#!/usr/bin/env python3
"""Per-source-line view of an ncu capture: joins the SASS page of a .ncu-rep (instructions
executed, active threads, stall samples per instruction) with nvdisasm's line table of the
same kernel in libcavept.so (-lineinfo), and prints the hottest source lines.

  python profiles/hotspots.py gpurun_out/prof_trace.ncu-rep 'k_traceILb1ELb0' [launch-id] [top-n]

The library must be the build the capture was taken from (same SASS order)."""
import collections
import csv
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "caveengine_b200", "libcavept.so")


def line_table(mangled_substr):
    tmp = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", LIB], cwd=tmp, check=True, capture_output=True)
    table = []
    for cub in sorted(os.listdir(tmp)):
        out = subprocess.run(["nvdisasm", "--print-line-info", os.path.join(tmp, cub)], capture_output=True, text=True).stdout
        inside, cur = False, ("?", 0)
        for ln in out.splitlines():
            if ln.startswith(".text."):
                inside = mangled_substr in ln
                continue
            if not inside:
                continue
            m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
            if m:
                cur = (os.path.basename(m.group(1)), int(m.group(2)))
                continue
            m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
            if m:
                table.append((int(m.group(1), 16), cur, m.group(2).strip()))
        if table:
            break
    return table


def main():
    rep, sub = sys.argv[1], sys.argv[2]
    launch = sys.argv[3] if len(sys.argv) > 3 else "1"
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 45
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    heads = [i for i, r in enumerate(rows) if "Instructions Executed" in r]
    h = heads[int(launch) - 1]
    end = heads[int(launch)] if int(launch) < len(heads) else len(rows)
    hdr = rows[h]
    ie, it, isamp = hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed"), hdr.index("# Samples")
    inst = [(r[1], int(r[ie] or 0), int(r[it] or 0), int(r[isamp] or 0)) for r in rows[h + 1:end] if len(r) == len(hdr)]
    table = line_table(sub)
    if len(table) != len(inst):
        print(f"# warning: {len(table)} SASS instructions in the library vs {len(inst)} in the capture", file=sys.stderr)
    per = collections.defaultdict(lambda: [0, 0, 0, 0])
    tot_e = sum(x[1] for x in inst) or 1
    tot_t = sum(x[2] for x in inst)
    tot_s = sum(x[3] for x in inst) or 1
    for (off, loc, sass), (_, e, t, s) in zip(table, inst):
        p = per[loc]
        p[0] += e
        p[1] += t
        p[2] += s
        p[3] += 1
    print(f"# {sub} launch {launch}: {tot_e} warp-instructions, {tot_t / tot_e:.2f} active threads/instruction, {len(inst)} SASS instructions")
    print("#  %inst  %stall-samples  active  n_sass  file:line")
    src_cache = {}
    for loc, (e, t, s, n) in sorted(per.items(), key=lambda kv: -kv[1][0])[:top]:
        f, l = loc
        text = ""
        for d in ("caveengine_b200/csrc", "include"):
            pth = os.path.join(ROOT, d, f)
            if os.path.exists(pth):
                src_cache.setdefault(pth, open(pth).read().splitlines())
                if 0 < l <= len(src_cache[pth]):
                    text = src_cache[pth][l - 1].strip()[:90]
        print(f"{100 * e / tot_e:7.2f} {100 * s / tot_s:7.2f} {t / max(e, 1):7.1f} {n:6d}  {f}:{l}  {text}")


if __name__ == "__main__":
    main()
