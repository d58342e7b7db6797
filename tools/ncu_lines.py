"""Top source lines by warp-stall samples from `ncu --page source --csv --print-source cuda,sass`."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 30
hdr = None
agg = {}
fname = ''
for r in rows:
    if not r: continue
    if r[0] == 'File Path': fname = r[1].split('/')[-1]; continue
    if r[0] == 'Line No': hdr = r; continue
    if hdr is None or not r[0].isdigit() or len(r) < len(hdr): continue
    if r[2] != '-': continue            # per-SASS rows carry an address; per-line aggregates have '-'
    H = {h: i for i, h in enumerate(hdr)}
    samples = int(r[H['# Samples']])
    if samples == 0: continue
    st = sorted(((h, int(r[i])) for h, i in H.items() if h.startswith('stall_') and 'Not Issued' not in h and r[i].isdigit() and int(r[i]) > 0), key=lambda x: -x[1])[:2]
    key = (fname, int(r[0]))
    if key in agg: agg[key][0] += samples
    else: agg[key] = [samples, r[1].strip()[:90], st]
tot = sum(v[0] for v in agg.values())
print('total samples', tot)
for (f, ln), (s, src, st) in sorted(agg.items(), key=lambda x: -x[1][0])[:n]:
    print(f'{f}:{ln:4d} {s:6d} {100.0*s/tot:5.1f}%  {src}  {st}')
