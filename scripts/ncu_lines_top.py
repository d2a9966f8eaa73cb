"""Dev aid: top CUDA source lines of a kernel by warp-stall samples (needs -lineinfo and --import-source on).
python scripts/ncu_lines_top.py rep kernel_regex [n]"""
import csv, subprocess, sys
rep, pat = sys.argv[1], sys.argv[2]
n = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'cuda,sass', '--kernel-name', 'regex:' + pat],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
cur_file, hdr, lines = None, None, []
for r in rows:
    if not r:
        continue
    if r[0] == 'File Path':
        cur_file = r[1].split('/')[-1]
    elif r[0] == 'Line No':
        hdr = r
    elif hdr and r[0] not in ('', 'Function Name') and len(r) >= len(hdr) - 2:
        try:
            iS = hdr.index('# Samples')
            lines.append((int(r[iS] or 0), cur_file, r[0], r[1].strip(), r))
        except ValueError:
            pass
tot = sum(l[0] for l in lines)
print('total samples', tot)
stall_cols = [i for i, h in enumerate(hdr) if h.startswith('stall_')] if hdr else []
for s, f, ln, src, r in sorted(lines, key=lambda l: -l[0])[:n]:
    why = sorted(((hdr[i][6:], int(r[i] or 0)) for i in stall_cols if i < len(r) and (r[i] or '0').isdigit()), key=lambda kv: -kv[1])[:2]
    print(f'{s:8d} {100.0 * s / max(tot, 1):5.1f}%  {f}:{ln:>4}  {src[:90]:90s} {why}')
