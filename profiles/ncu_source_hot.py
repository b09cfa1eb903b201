#!/usr/bin/env python
"""Hot source lines of one launch in an .ncu-rep captured with --import-source on (-lineinfo builds).
usage: ncu_source_hot.py file.ncu-rep [launch-index (0-based)] [top N]"""
import csv
import io
import subprocess
import sys


def main():
    rep = sys.argv[1]
    idx = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
    # one block per launch, each starting with a "Kernel Name" line
    blocks, cur = [], []
    for line in out.splitlines():
        if line.startswith('"Function Name"'):
            if cur:
                blocks.append(cur)
            cur = [line]
        elif cur:
            cur.append(line)
    if cur:
        blocks.append(cur)
    b = blocks[idx]
    print(b[0][:160])
    rows = list(csv.reader(io.StringIO("\n".join(l for l in b[1:] if not l.startswith('"File Path"')))))
    hdr = rows[0]
    ci = {}
    for i, h in enumerate(hdr):
        ci.setdefault(h, i)
    src = ci.get("Source")
    rows = [rows[0]] + [r for r in rows[1:] if r and r[0] != ""]   # per-line aggregates only
    samp = ci.get("# Samples") or ci.get("Warp Stall Sampling (All Samples)")
    inst = ci.get("Instructions Executed")
    recs = []
    for r in rows[1:]:
        try:
            recs.append((int(r[samp] or 0), int(r[inst] or 0), r[0], r[src].strip()[:150]))
        except Exception:
            pass
    ts, ti = sum(r[0] for r in recs) or 1, sum(r[1] for r in recs) or 1
    print(f"total samples {ts}, warp instructions {ti}")
    for s, i, ln, text in sorted(recs, reverse=True)[:top]:
        print(f"{100.0 * s / ts:6.2f}% smp {100.0 * i / ti:6.2f}% inst  {ln:>6s}  {text}")


if __name__ == "__main__":
    main()
