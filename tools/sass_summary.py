#!/usr/bin/env python
"""Per-kernel SASS evidence for libhfops.so (runs without a GPU).

    python tools/sass_summary.py [--out profiles/r2_sass_summary.txt]

Disassembles the shipped library with `cuobjdump -sass` and counts, per kernel, the mnemonics that
show which Blackwell mechanisms the kernel really uses (B200_PROFILING.md: names that prove TMA /
DSMEM / mbarrier use):
    UBLKCP      1-D bulk copy global -> shared (cp.async.bulk; the TMA engine, no tensor map)
    UTMALDG/STG tensor-map TMA (none expected: tiles here are 12-byte AoS records)
    STAS        st.async into a peer CTA's shared memory (DSMEM push)
    SYNCS       mbarrier operations
    REDUX       redux.sync warp reductions
    LDG.E.128 / STG.E.128   128-bit global accesses (vectorised gathers / streaming stores)
    LDS.128     128-bit shared-memory reads
    UTC*MMA / LDTM / STTM   tcgen05 tensor-core / TMEM (none expected: K = 3, index bit-exactness)
"""
import argparse
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "heterofusionrcnn_b200", "libhfops.so")
COLS = ["UBLKCP", "UTMALDG", "UTMASTG", "STAS", "SYNCS", "REDUX", "LDG.E.128", "STG.E.128", "LDS.128", "FFMA2", "UTCMMA", "LDTM"]


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), stdout=subprocess.PIPE, text=True, check=True).stdout
    return out.splitlines()


def short(name):
    name = re.sub(r"\(anonymous namespace\)::", "", name)
    name = re.sub(r"^void ", "", name)
    name = re.sub(r"^hf::", "", name)
    m = re.match(r"([A-Za-z0-9_]+(?:<[^(]*>)?)\(", name)
    return m.group(1) if m else name


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    sass = subprocess.run(["cuobjdump", "-sass", LIB], stdout=subprocess.PIPE, text=True, check=True).stdout
    kernels = collections.OrderedDict()
    cur = None
    arch = set()
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = kernels.setdefault(m.group(1), collections.Counter())
            continue
        m = re.search(r"arch = (sm_\w+)", line)
        if m:
            arch.add(m.group(1))
        if cur is None:
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if not m:
            continue
        op = m.group(1)
        cur["_total"] += 1
        for c in COLS:
            if c == "UTCMMA":
                hit = op.startswith("UTC") and "MMA" in op
            elif c == "REDUX":
                hit = "REDUX" in op  # REDUX.* and CREDUX.* (redux.sync min/max lower to CREDUX on sm_100a)
            elif c in ("LDG.E.128", "STG.E.128"):
                hit = op.startswith(c[:3] + ".") and ".128" in op  # incl. LDG.E.NA.128.CONSTANT, STG.E.EF.128
            else:
                hit = op == c or op.startswith(c + ".")
            if hit:
                cur[c] += 1
    names = demangle(list(kernels))
    rows = []
    for (mangled, cnt), nm in zip(kernels.items(), names):
        rows.append((short(nm), cnt))
    rows.sort(key=lambda r: r[0])
    w = max(len(r[0]) for r in rows) if rows else 10
    w = min(w, 70)
    lines = ["# tools/sass_summary.py: cuobjdump -sass heterofusionrcnn_b200/libhfops.so  (%d kernels, arch %s)" % (
        len(rows), ",".join(sorted(arch)) or "?"),
             "# counts of SASS mnemonics per kernel; see the tool's docstring for what each one proves", ""]
    lines.append("%-*s %7s " % (w, "kernel", "instrs") + " ".join("%9s" % c for c in COLS))
    tot = collections.Counter()
    for nm, cnt in rows:
        lines.append("%-*s %7d " % (w, nm[:w], cnt["_total"]) + " ".join("%9d" % cnt[c] for c in COLS))
        tot.update(cnt)
    lines.append("")
    lines.append("%-*s %7d " % (w, "TOTAL", tot["_total"]) + " ".join("%9d" % tot[c] for c in COLS))
    text = "\n".join(lines) + "\n"
    if args.out:
        open(os.path.join(ROOT, args.out) if not os.path.isabs(args.out) else args.out, "w").write(text)
    sys.stdout.write(text)


if __name__ == "__main__":
    main()
