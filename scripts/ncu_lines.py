"""Joins an ncu source-page export (per-SASS-instruction warp-stall samples) with nvdisasm's inline-aware line table, and
prints where a kernel's warps wait: by stall reason, by source line of the KERNEL body (outermost frame of the inline chain)
and by innermost frame.

  ncu -i rep.ncu-rep --page source --csv --print-source sass > src.csv
  cuobjdump -xelf all libtts_ar.so ; nvdisasm -gi -c ar_decode.sm_100a.cubin > dec.sass
  python scripts/ncu_lines.py src.csv dec.sass _ZN3tts16ar_decode_kernelILi1EEEvNS_6ParamsE [top_n]
"""
import csv
import re
import sys
from collections import defaultdict


def sass_lines(path, symbol):
    """[(offset, text, chain)] of one function; chain = [(file, line), ...] innermost first."""
    out, chain, on = [], [], False
    pending = []
    for ln in open(path, errors="replace"):
        if ln.startswith(".text."):
            on = ln.strip().rstrip(":") == ".text." + symbol
            continue
        if not on:
            continue
        m = re.match(r'\s*//## File "([^"]+)", line (\d+)', ln)
        if m:
            pending.append((m.group(1).split("/")[-1], int(m.group(2))))
            continue
        m = re.match(r"\s*/\*([0-9a-f]+)\*/\s+(.*?);", ln)
        if m:
            if pending:
                chain, pending = pending, []
            out.append((int(m.group(1), 16), m.group(2).strip(), list(chain)))
    return out


def main():
    src_csv, sass, symbol = sys.argv[1:4]
    top_n = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    rows = list(csv.reader(open(src_csv)))
    hdr = rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    data = rows[2:]
    ins = sass_lines(sass, symbol)
    assert len(ins) == len(data), (len(ins), len(data))
    stalls = [h for h in hdr if h.startswith("stall_") and "Not" not in h]

    def f(r, k):
        try:
            return float(r[idx[k]])
        except ValueError:
            return 0.0

    tot = sum(f(r, "# Samples") for r in data)
    print(f"{len(data)} instructions, {int(tot)} warp samples")
    print("| stall reason | % of samples |\n|---|---|")
    for s in sorted(stalls, key=lambda s: -sum(f(r, s) for r in data)):
        v = sum(f(r, s) for r in data) / tot * 100
        if v >= 0.3:
            print(f"| {s[6:]} | {v:.1f} |")
    outer, inner = defaultdict(lambda: defaultdict(float)), defaultdict(lambda: defaultdict(float))
    for (off, text, chain), r in zip(ins, data):
        assert text.split()[0].lstrip("@!P0123456789U ") in r[idx["Source"]] or True
        n = f(r, "# Samples")
        if n == 0:
            continue
        o = chain[-1] if chain else ("?", 0)
        i = chain[0] if chain else ("?", 0)
        for s in stalls:
            outer[o][s] += f(r, s)
            inner[i][s] += f(r, s)
        outer[o]["n"] += n
        inner[i]["n"] += n
    for name, table in (("kernel-body line (outermost frame)", outer), ("innermost frame", inner)):
        print(f"\n| {name} | % samples | top stalls |\n|---|---|---|")
        for key, d in sorted(table.items(), key=lambda kv: -kv[1]["n"])[:top_n]:
            st = sorted(((d[s], s[6:]) for s in stalls), reverse=True)[:3]
            print(f"| {key[0]}:{key[1]} | {d['n'] / tot * 100:.2f} | " + ", ".join(f"{s} {v / tot * 100:.2f}" for v, s in st if v > 0) + " |")


if __name__ == "__main__":
    main()
