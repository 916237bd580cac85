"""Developer tool: opcode mix and stall mix of one kernel from `ncu --page source --csv --print-source sass`.
    python tools/ncu_sass_mix.py file.csv [top_n_instructions]"""
import csv, collections, sys
rows = list(csv.reader(open(sys.argv[1])))
hi = [i for i, r in enumerate(rows) if r and r[0] == 'Address'][0]
hdr = rows[hi]; idx = {h: i for i, h in enumerate(hdr)}
data = [r for r in rows[hi + 1:] if len(r) > idx['Instructions Executed'] and r[idx['Source']].strip()]
f = lambda r, k: float(r[idx[k]] or 0)
tot_s = sum(f(r, '# Samples') for r in data); tot_i = sum(f(r, 'Instructions Executed') for r in data)
print(rows[0][1][:90], '| warp instr', int(tot_i), 'samples', int(tot_s))
stalls = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
agg = {h: sum(f(r, h) for r in data) for h in stalls}
print('stalls %:', ' '.join(f"{h[6:]}={v / tot_s * 100:.1f}" for h, v in sorted(agg.items(), key=lambda x: -x[1])[:8]))
ops = collections.Counter(); opsS = collections.Counter()
for r in data:
    src = r[idx['Source']].strip().split()
    op = (src[1] if src[0].startswith('@') else src[0]).split('.')[0]
    ops[op] += f(r, 'Instructions Executed'); opsS[op] += f(r, '# Samples')
print('op: %instr %samples')
for op, v in ops.most_common(16):
    print(f"  {op:8s} {v / tot_i * 100:5.1f} {opsS[op] / tot_s * 100:5.1f}")
n = int(sys.argv[2]) if len(sys.argv) > 2 else 0
if n:
    print('hottest instructions by samples:')
    for r in sorted(data, key=lambda r: -f(r, '# Samples'))[:n]:
        print(f"  {f(r, '# Samples') / tot_s * 100:5.2f}% {int(f(r, 'Instructions Executed')):>10d} {r[idx['Address']]} {r[idx['Source']].strip()[:90]}")
