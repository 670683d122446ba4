"""Summarise an `ncu --page source --csv` dump: opcode mix, instructions per warp iteration, top stall rows."""
import csv, sys
from collections import Counter
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]; data = rows[2:]
ix = {h: i for i, h in enumerate(hdr)}
tot = sum(int(r[ix['Instructions Executed']]) for r in data)
c = Counter(); s = Counter()
for r in data:
    op = r[ix['Source']].split()
    o = (op[0] if not op[0].startswith('@') else op[1]).split('.')[0]
    c[o] += int(r[ix['Instructions Executed']]); s[o] += int(r[ix['# Samples']])
print('warp instructions', tot, 'sass lines', len(data))
for o, v in c.most_common(14):
    print(f'  {o:10s} {v / tot * 100:5.1f}%  samples {s[o]}')
print('avg active threads', round(sum(int(r[ix['Thread Instructions Executed']]) for r in data) / tot, 2))
it = [int(r[ix['Instructions Executed']]) for r in data if 'WARPSYNC' in r[ix['Source']]]
if it:
    print('warp iterations', max(it), 'instructions / warp iteration', round(tot / max(it), 1))
for r in sorted(data, key=lambda r: -int(r[ix['# Samples']]))[:int(sys.argv[2]) if len(sys.argv) > 2 else 12]:
    print(' ', r[ix['# Samples']], r[ix['Instructions Executed']], r[ix['Avg. Threads Executed']], r[ix['Source']][:90])
