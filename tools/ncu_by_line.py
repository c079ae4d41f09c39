"""Joins an `ncu --page source --csv` SASS table of one kernel with `nvdisasm --print-line-info-inline`
of the same cubin and sums executed instructions / stall samples per source line.

    cuobjdump -xelf all build/kernels_fast.o                    # -> rtw_kernels.sm_100a.cubin
    nvdisasm --print-line-info-inline rtw_kernels.sm_100a.cubin > all.txt
    ncu -i rep.ncu-rep --page source --csv --launch-count 1 > k.csv
    python tools/ncu_by_line.py k.csv all.txt wf_extend_kernel [outer|inner|inner@<kernel line>]

`inner` keys by the innermost (file:line) of every instruction, `outer` by the line of the kernel's
own file the instruction was inlined into (so a whole inlined call shows up as one row);
`inner@328` is `inner` restricted to what was inlined into line 328 of the kernel's file; `outer2` keys one call level
below the kernel's own lines (a kernel that is one call of an inlined body, like mega_kernel).
"""
import csv
import re
import sys
from collections import defaultdict


def load_listing(path, kernel):
    """offset -> (inner (file, line), outer (file, line)) for the kernel's .text section."""
    out = {}
    on = False
    inner = outer = None
    pending = []
    for ln in open(path, errors="replace"):
        if ln.startswith(".text."):
            on = kernel in ln and "variant_fast" in ln
            continue
        if not on:
            continue
        m = re.match(r'\s*//## File "([^"]+)", line (\d+)(?: inlined at "([^"]+)", line (\d+))?', ln)
        if m:
            pending.append(m)
            continue
        m2 = re.match(r"\s*/\*([0-9a-f]+)\*/\s+(.*?);", ln)
        if m2:
            if pending:
                first = pending[0]
                inner = (first.group(1).split("/")[-1], int(first.group(2)))
                last = pending[-1]
                outer = (last.group(1).split("/")[-1], int(last.group(2)))
                second = pending[-2] if len(pending) > 1 else last      # one call level below the kernel's own lines
                outer2 = (second.group(1).split("/")[-1], int(second.group(2)))
                pending = []
            out[int(m2.group(1), 16)] = (inner, outer, m2.group(2).strip(), outer2)
    return out


def main():
    csv_path, listing, kernel = sys.argv[1:4]
    mode = sys.argv[4] if len(sys.argv) > 4 else "outer"
    only = None
    if "@" in mode:
        mode, at = mode.split("@")
        only = int(at)
    lst = load_listing(listing, kernel)
    rows = list(csv.reader(open(csv_path)))
    hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[hi]
    ix = {h: i for i, h in enumerate(hdr)}
    body = rows[hi + 1:]
    base = int(body[0][0], 16)
    agg = defaultdict(lambda: [0, 0, 0, 0])
    tot = [0, 0, 0, 0]
    for r in body:
        off = int(r[0], 16) - base
        info = lst.get(off)
        if only is not None and not (info and info[1][1] == only):
            continue
        key = (info[0] if mode == "inner" else info[3] if mode == "outer2" else info[1]) if info else ("?", 0)
        vals = [int(r[ix["Instructions Executed"]] or 0), int(r[ix["Thread Instructions Executed"]] or 0),
                int(r[ix["# Samples"]] or 0), int(r[ix["Warp Stall Sampling (Not-issued Samples)"]] or 0)]
        for k in range(4):
            agg[key][k] += vals[k]
            tot[k] += vals[k]
    print(f"total: warp-instr {tot[0]}  thread-instr {tot[1]}  avg lanes {tot[1] / max(1, tot[0]):.1f}  samples {tot[2]}")
    print(f"{'where':34s} {'warp-instr':>11s} {'%':>6s} {'lanes':>6s} {'samples':>8s} {'%':>6s} {'not-issued%':>11s}")
    for key, v in sorted(agg.items(), key=lambda kv: -kv[1][0 if only is not None else 2])[:60]:
        print(f"{key[0] + ':' + str(key[1]):34s} {v[0]:11d} {100 * v[0] / tot[0]:6.1f} {v[1] / max(1, v[0]):6.1f} {v[2]:8d} "
              f"{100 * v[2] / max(1, tot[2]):6.1f} {100 * v[3] / max(1, v[2]):11.1f}")


if __name__ == "__main__":
    main()
