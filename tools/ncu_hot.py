"""Top stall sites from an ncu --page source --csv dump (first kernel only)."""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
# find header row (starts with Address)
hi = next(i for i, r in enumerate(rows) if r and r[0] == 'Address')
hdr = rows[hi]
body = []
for r in rows[hi + 1:]:
    if not r or r[0] == 'Kernel Name' or r[0] == 'Address': break
    body.append(dict(zip(hdr, r)))
tot = sum(int(b['# Samples'] or 0) for b in body)
print("instructions:", len(body), "samples:", tot, "warp-inst executed:", sum(int(b['Instructions Executed'] or 0) for b in body))
ops = collections.Counter()
for b in body:
    op = b['Source'].split()[0] if not b['Source'].strip().startswith('@') else b['Source'].split()[1]
    ops[op.split('.')[0]] += int(b['Instructions Executed'] or 0)
print("executed by opcode:", ", ".join(f"{k}={v}" for k, v in ops.most_common(25)))
N = int(sys.argv[2]) if len(sys.argv) > 2 else 30
keys = ['stall_long_sb', 'stall_short_sb', 'stall_wait', 'stall_barrier', 'stall_math', 'stall_mio', 'stall_no_inst', 'stall_not_selected', 'stall_selected', 'stall_dispatch', 'stall_lg', 'stall_membar']
for i, b in sorted(enumerate(body), key=lambda x: -int(x[1]['# Samples'] or 0))[:N]:
    s = int(b['# Samples'] or 0)
    why = ", ".join(f"{k[6:]}={b[k]}" for k in keys if k in b and int(b[k] or 0) > 0)
    print(f"{i:5d} {s:5d} {100*s/tot:5.1f}%  {b['Source'][:70]:70s} ex={b['Instructions Executed']:>7s} | {why}")
