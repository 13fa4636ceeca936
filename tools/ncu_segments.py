"""Summarise the source page of an ncu report: stall-reason totals, samples between landmark
instructions (barriers, reductions, divides) and the most-sampled instructions.
usage: ncu -i X.ncu-rep --page source --csv > X.csv ; python tools/ncu_segments.py X.csv [norm]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
h = rows[1]; data = [dict(zip(h, r)) for r in rows[2:]]
stalls = [c for c in h if c.startswith('stall_') and 'Not Issued' not in c]
N = sum(int(d['# Samples'] or 0) for d in data)
inst = sum(int(d['Instructions Executed'] or 0) for d in data)
norm = float(sys.argv[2]) if len(sys.argv) > 2 else max(int(d['Instructions Executed'] or 0) for d in data)
print('kernel', rows[0][1][:80]); print('samples', N, 'warp-inst', inst, 'per unit', round(inst / norm, 1))
tot = {s: sum(int(d[s] or 0) for d in data) for s in stalls}
print(' '.join(f"{s[6:]}={v / N * 100:.1f}%" for s, v in sorted(tot.items(), key=lambda x: -x[1])[:9]))
marks = ('BAR.SYNC', 'REDUX', 'CREDUX', 'WARPSYNC.ALL', 'MUFU.RCP64H', 'CALL', 'RET', 'SYNCS', 'UBLKCP')
acc = ins = 0
for i, d in enumerate(data):
    s = d['Source'].strip(); smp = int(d['# Samples'] or 0); ex = int(d['Instructions Executed'] or 0)
    t = s.split(); op = t[1] if t[0].startswith('@') else t[0]
    if any(op.startswith(m) for m in marks):
        print(f"{i:5d} {acc / N * 100:5.1f}% {ins / norm:7.1f} inst -> {s[:50]} [{smp / N * 100:.1f}%, x{ex / norm:.2f}]")
        acc = ins = 0
    acc += smp; ins += ex
print(f"tail {acc / N * 100:.1f}% {ins / norm:.1f}")
print('--- top instructions')
for d in sorted(data, key=lambda d: -int(d['# Samples'] or 0))[:int(sys.argv[3]) if len(sys.argv) > 3 else 14]:
    st = sorted(((s[6:], int(d[s] or 0)) for s in stalls), key=lambda x: -x[1])[:2]
    print(f"{int(d['# Samples']) / N * 100:5.1f}% x{int(d['Instructions Executed'] or 0) / norm:.2f} {d['Source'].strip()[:60]} {st}")
