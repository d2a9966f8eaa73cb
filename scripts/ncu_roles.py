"""Dev aid: warp-stall samples / instruction counts / stall reasons of a warp-specialised kernel aggregated per role.
python scripts/ncu_roles.py rep kernel_regex source.cu fn1=role1,fn2=role2,..."""
import csv, subprocess, sys
rep, pat, srcfile, spec = sys.argv[1:5]
fnmap = [kv.split('=') for kv in spec.split(',')]
src = open(srcfile).read().splitlines()
base = srcfile.split('/')[-1]
def role(ln):
    for i in range(min(ln, len(src)) - 1, -1, -1):
        for fn, r in fnmap:
            if fn in src[i] and '(' in src[i] and ';' not in src[i].split('(')[0]:
                return r
    return 'other'
out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'cuda,sass', '--kernel-name', 'regex:' + pat],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
cur = hdr = key = None
samp, inst, sass, stalls = {}, {}, {}, {}
for r in rows:
    if not r:
        continue
    if r[0] == 'File Path':
        cur = r[1].split('/')[-1]
    elif r[0] == 'Line No':
        hdr = r
    elif hdr and r[0] not in ('', 'Function Name'):
        try:
            ln = int(r[0])
        except ValueError:
            continue
        key = cur if cur != base else base + ':' + role(ln)
        iE, iS = hdr.index('Instructions Executed'), hdr.index('# Samples')
        off = len(r) - len(hdr)     # a source line containing commas inside quotes can shift the columns
        try:
            samp[key] = samp.get(key, 0) + int(r[iS + off] or 0)
            inst[key] = inst.get(key, 0) + int(r[iE + off] or 0)
        except (ValueError, IndexError):
            continue
        d = stalls.setdefault(key, {})
        for i, h in enumerate(hdr):
            if h.startswith('stall_') and 'Not Issued' not in h and 0 <= i + off < len(r) and (r[i + off] or '0').isdigit():
                d[h[6:]] = d.get(h[6:], 0) + int(r[i + off] or 0)
    elif hdr and r[0] == '' and len(r) > 3 and r[2].startswith('0x'):
        sass[key] = sass.get(key, 0) + 1
tot = sum(samp.values())
for k, v in sorted(samp.items(), key=lambda kv: -kv[1])[:14]:
    top = sorted(stalls.get(k, {}).items(), key=lambda kv: -kv[1])[:4]
    print(f"{k:26s} samples {100 * v / tot:5.1f}%  warp-inst {inst[k] / 1e6:9.1f}M  sass {sass.get(k, 0):5d}  {[(a, round(100 * b / max(v, 1))) for a, b in top]}")
