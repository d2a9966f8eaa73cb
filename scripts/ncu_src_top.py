"""Dev aid: top stall sites of a kernel from an .ncu-rep source page (SASS view).  python scripts/ncu_src_top.py rep kernel_regex [n]"""
import csv, subprocess, sys
rep, pat = sys.argv[1], sys.argv[2]
n = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--kernel-name', 'regex:' + pat], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
# several kernels may follow each other; take the first block
hdr = rows[1]
body = []
for r in rows[2:]:
    if len(r) < len(hdr) or r[0] == 'Kernel Name' or r[0] == 'Address':
        if r and r[0] == 'Kernel Name':
            break
        continue
    body.append(r)
iS = hdr.index('# Samples'); iSrc = hdr.index('Source'); iEx = hdr.index('Instructions Executed')
stall_cols = [i for i, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
tot = sum(int(r[iS] or 0) for r in body)
print('total samples', tot, 'instructions', len(body))
agg = {}
for i in stall_cols:
    agg[hdr[i]] = sum(int(r[i] or 0) for r in body)
print('by reason:', {k: v for k, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v})
idx = sorted(range(len(body)), key=lambda k: -int(body[k][iS] or 0))[:n]
for k in sorted(idx):
    r = body[k]
    why = {hdr[i][6:]: int(r[i]) for i in stall_cols if r[i] and int(r[i]) > 0}
    top = sorted(why.items(), key=lambda kv: -kv[1])[:3]
    print(f'{k:6d} {int(r[iS]):7d} {100.0 * int(r[iS]) / max(tot, 1):5.1f}%  ex {r[iEx]:>9}  {r[iSrc].strip()[:70]:70s} {top}')
