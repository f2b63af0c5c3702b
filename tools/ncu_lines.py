#!/usr/bin/env python
"""Attribute ncu warp-stall samples to source lines (ncu's CLI source page is SASS only).

    python tools/ncu_lines.py REPORT.ncu-rep LIB.so KERNEL_SUBSTR [top_n]

Joins `ncu --page source --csv` (per-SASS-instruction samples) with `nvdisasm -g` line info of the
same kernel in LIB.so by instruction offset; lines in CUDA/system headers are charged to the last
line of this repo seen before them."""
import csv
import io
import os
import re
import subprocess
import sys
import tempfile
from collections import defaultdict


def main():
    rep, lib, kern = sys.argv[1:4]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    tmp = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=tmp, check=True, stdout=subprocess.DEVNULL)
    dis = []                                                   # (one cubin per translation unit)
    for cubin in sorted(os.path.join(tmp, f) for f in os.listdir(tmp) if f.endswith(".cubin")):
        dis += subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout.splitlines()
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    kname = rows[0][1]
    hdr = rows[1]
    ia, isrc, isamp = hdr.index("Address"), hdr.index("Source"), hdr.index("Warp Stall Sampling (All Samples)")
    stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
    insts = [r for r in rows[2:] if len(r) == len(hdr)]
    base = int(insts[0][ia], 16)
    # find the function in the disassembly: mangled name containing all tokens of kern
    norm = lambda t: re.sub(r"\((bool|int)\)", "", t).replace("false", "0").replace("true", "1").replace(" ", "")
    want = None
    for m in re.finditer(r"^\.text\.(\S+):$", "\n".join(dis), flags=re.M):
        dem = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        if kern in dem and norm(dem).split("(spa::")[0] == norm(kname).split("(spa::")[0]:
            want = m.group(1)
            break
    assert want, f"kernel {kname} not found in {lib}"
    line_of = {}
    cur = ("?", 0)
    inside = False
    for ln in dis:
        if ln.startswith(".text." + want + ":"):
            inside = True
            continue
        if inside and ln.startswith("//-----"):
            break
        if not inside:
            continue
        m = re.match(r'\s*//## File "([^"]+)", line (\d+)', ln)
        if m:
            if "/root/repo" in m.group(1) or "csrc" in m.group(1):
                cur = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        m = re.match(r"\s*/\*([0-9a-f]+)\*/", ln)
        if m:
            line_of[int(m.group(1), 16)] = cur
    per = defaultdict(lambda: [0, defaultdict(int)])
    total = 0
    for r in insts:
        off = int(r[ia], 16) - base
        s = int(r[isamp] or 0)
        total += s
        key = line_of.get(off, ("?", 0))
        per[key][0] += s
        for i in stall_cols:
            v = int(r[i] or 0)
            if v:
                per[key][1][hdr[i]] += v
    print(f"kernel: {kname}\ntotal samples {total}")
    # per source file: samples and the stall reasons behind them
    byfile = defaultdict(lambda: [0, defaultdict(int)])
    for (f, l), (sm, st) in per.items():
        byfile[f][0] += sm
        for k, v in st.items():
            byfile[f][1][k] += v
    for f, (sm, st) in sorted(byfile.items(), key=lambda kv: -kv[1][0]):
        tops = ", ".join(f"{k[6:]} {v}" for k, v in sorted(st.items(), key=lambda kv: -kv[1])[:8])
        print(f"  {f:18s} {sm:6d} samples ({100 * sm / max(total, 1):4.1f} %): {tops}")
    srcs = {}
    for (f, l), (s, st) in sorted(per.items(), key=lambda kv: -kv[1][0])[:top]:
        if f not in srcs:
            path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "cann-ops-adv-vllm_b200", "csrc", f)
            srcs[f] = open(path).read().splitlines() if os.path.exists(path) else []
        text = srcs[f][l - 1].strip()[:70] if 0 < l <= len(srcs[f]) else ""
        tops = ",".join(f"{k[6:]}:{v}" for k, v in sorted(st.items(), key=lambda kv: -kv[1])[:3])
        print(f"{s:7d} {100 * s / max(total, 1):5.1f}%  {f}:{l:<4d} {text:70s} [{tops}]")


if __name__ == "__main__":
    main()
