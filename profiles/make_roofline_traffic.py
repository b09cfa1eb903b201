#!/usr/bin/env python
"""Turns an ncu capture of DRAM bytes into profiles/roofline_traffic.json, stamped with a hash of the kernel sources.
    ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv --log-file gpurun_out/traffic.csv \
        -k regex:'k_interp_flat|k_interp_trunc|k_spgemm' python bench.py --steps 1 --warmup 0 --no-e2e --no-cpu-baseline --parity-n 0
    python profiles/make_roofline_traffic.py gpurun_out/traffic.csv
`dram_bytes_per_launch` of a kernel class = DRAM bytes of all its kernels in ONE step / the number of profiler scopes of that
class in one step (bench.py's `launches_timed` / steps), i.e. per launch in the sense of `roofline.achieved`.  bench.py only
reports the number when `source_sha` equals the hash of the sources it runs (source_sha() below, which bench.py imports).  The hash is
taken over the code with comments and white space removed (code_only): what the compiler sees."""
import csv
import glob
import hashlib
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CLASSES = {"interp_rows": ("k_interp_flat", "k_interp_trunc", "k_interp_ovf"), "spgemm_numeric": ("k_spgemm_smem<", "k_spgemm_gmem<"),
           "spgemm_symbolic": ("k_spgemm_count_try", "k_spgemm_bound")}
SCOPES_PER_STEP = {"interp_rows": 4, "spgemm_numeric": 8, "spgemm_symbolic": 8}   # scopes above bench.py's 32 MB threshold at lap7_256


def code_only(text):
    """the C++ / CUDA source without // and /* */ comments (string and character literals respected) and with every run of white
    space collapsed: what the compiler sees, so that a comment or layout edit does not orphan a traffic capture"""
    out, i, n = [], 0, len(text)
    while i < n:
        c = text[i]
        if c == "/" and i + 1 < n and text[i + 1] == "/":
            while i < n and text[i] != "\n":
                i += 1
        elif c == "/" and i + 1 < n and text[i + 1] == "*":
            j = text.find("*/", i + 2)
            i = n if j < 0 else j + 2
            out.append(" ")
        elif c in "\"'":
            j = i + 1
            while j < n and text[j] != c:
                j += 2 if text[j] == "\\" else 1
            out.append(text[i:j + 1])
            i = j + 1
        else:
            out.append(c)
            i += 1
    return " ".join("".join(out).split())


def source_sha():
    h = hashlib.sha256()
    for f in sorted(glob.glob(os.path.join(ROOT, "hypre-mini-app_b200", "csrc", "*"))):
        h.update(os.path.basename(f).encode())
        h.update(code_only(open(f, encoding="utf-8").read()).encode())
    return h.hexdigest()[:16]


def main():
    rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if l.startswith('"'))]
    hdr = rows[0]
    name, metric, value, unit = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    tot = {c: 0.0 for c in CLASSES}
    launches = {c: 0 for c in CLASSES}
    for r in rows[1:]:
        if not r[metric].startswith("dram__bytes"):
            continue
        for c, pats in CLASSES.items():
            if any(p in r[name] for p in pats):
                # symbolic / numeric instantiations of k_spgemm_smem / k_spgemm_gmem differ by their last template argument
                if c == "spgemm_numeric" and re.search(r"k_spgemm_(smem<\d+, |gmem<)0>", r[name]):
                    c = "spgemm_symbolic"
                tot[c] += float(r[value].replace(",", "")) * scale.get(r[unit], 1.0)
                launches[c] += 1 if r[metric].endswith("read.sum") else 0
                break
    out = {"source_sha": source_sha(), "source_sha_kind": "sha256 over csrc/* with comments and white space removed (code_only)", "how": "ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum over one step of bench.py (lap7_256); "
                                              "bytes of all kernels of the class / profiler scopes of the class per step"}
    for c in CLASSES:
        if launches[c]:
            out[c] = {"dram_bytes_per_launch": tot[c] / SCOPES_PER_STEP[c], "dram_bytes_per_step": tot[c], "kernel_launches_per_step": launches[c]}
    json.dump(out, open(os.path.join(ROOT, "profiles", "roofline_traffic.json"), "w"), indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
