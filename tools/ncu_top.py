"""Summarise an `ncu --page source --csv --print-source sass` export: top stalled SASS instructions."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out, k = [], 0
for r in rows:
    if r and r[0] == 'Address':
        hdr = r; k += 1
        if k == 2: break
        continue
    if k == 1 and r and r[0].startswith('0x'): out.append(r)
H = {h: i for i, h in enumerate(hdr)}
tot = sum(int(r[H['# Samples']]) for r in out)
stalls = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
agg = {s: sum(int(r[H[s]]) for r in out) for s in stalls}
print('total samples', tot, 'instructions', len(out))
print(sorted(agg.items(), key=lambda x: -x[1])[:8])
top = sorted(enumerate(out), key=lambda x: -int(x[1][H['# Samples']]))[:n]
for idx, r in sorted(top):
    st = sorted(((s, int(r[H[s]])) for s in stalls if int(r[H[s]]) > 0), key=lambda x: -x[1])[:3]
    print(idx, r[H['Source']].strip()[:64], '| samples', r[H['# Samples']], '| exec', r[H['Instructions Executed']], st)
