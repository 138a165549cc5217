"""Bring-up tool: joins the per-instruction warp-state samples of an ncu --set full capture of ar_decode_kernel
(ncu -i X.ncu-rep --page source --csv --print-source sass > src.csv) with nvdisasm -gi line info of the same build
(cuobjdump -xelf all libtts_ar.so; nvdisasm -gi -c ar_decode.sm_100a.cubin > di.txt) and prints the share of warp
samples and the top stall reasons per top-level source line of the kernel body.
usage: python scripts/phase_samples.py src.csv di.txt"""
import collections
import csv
import re
import sys

src_csv, di_txt = sys.argv[1], sys.argv[2]
outer, infunc, block, last, prev = [], False, [], None, False
for line in open(di_txt):
    if line.startswith('.text.'):
        infunc = 'ar_decode_kernel' in line
    if '//## File' in line:
        if not prev:
            block = []
        block.append(line)
        prev = True
        continue
    if re.match(r'\s+/\*[0-9a-f]{4,}\*/', line):
        if prev and block:
            m = re.search(r'inlined at "([^"]+)", line (\d+)', block[-1]) or re.search(r'File "([^"]+)", line (\d+)', block[-1])
            last = (m.group(1).split('/')[-1], int(m.group(2)))
        if infunc:
            outer.append(last)
    prev = False
rows = list(csv.reader(open(src_csv)))
hdr, data = rows[1], rows[2:]
assert len(data) == len(outer), (len(data), len(outer))
idx = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
agg = collections.defaultdict(collections.Counter)
for o, r in zip(outer, data):
    for s_ in stalls:
        v = int(r[idx[s_]] or 0)
        if v:
            agg[o][s_] += v
    agg[o]['n_instr'] += 1
    agg[o]['samples'] += int(r[idx['# Samples']] or 0)
tot = sum(a['samples'] for a in agg.values())
print('instructions', len(outer), 'samples', tot)
for o, a in sorted(agg.items(), key=lambda x: (x[0][0], x[0][1])):
    if a['samples'] > tot * 0.004:
        top = sorted(((v, k) for k, v in a.items() if k.startswith('stall_')), reverse=True)[:5]
        print(f"{100 * a['samples'] / tot:5.1f}% n={a['n_instr']:5d} {o[0]}:{o[1]:4d} " + ' '.join(f"{k[6:]}={100 * v / a['samples']:.0f}" for v, k in top))
