"""Tensor-path evidence per kernel of libngp_b200.so: counts of the SASS mnemonics that prove (or disprove) the Blackwell
path (B200_PROFILING.md): UTC*MMA = tcgen05.mma, LDTM / STTM = tcgen05.ld / st, UTMALDG = TMA, HMMA = legacy mma.sync.
  python scripts/sass_opcounts.py > profiles/r2_sass_opcounts.txt"""
import collections, os, re, subprocess, sys
so = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "nerf_navigation_b200", "libngp_b200.so")
out = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
pats = {"UTC*MMA": r"\bUTC\w*MMA\b", "LDTM": r"\bLDTM\b", "STTM": r"\bSTTM\b", "UTMALDG": r"\bUTMALDG\b", "UBLKCP": r"\bUBLKCP\b", "HMMA": r"\bHMMA\b",
        "SYNCS(mbarrier)": r"\bSYNCS\b", "LDG": r"\bLDG\b", "LDS": r"\bLDS\b", "ATOMS": r"\bATOMS\b"}
cur, cnt, tot = None, collections.OrderedDict(), {}
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
        cnt[cur] = collections.Counter(); tot[cur] = 0
        continue
    if cur and re.match(r"\s+/\*[0-9a-f]{4,5}\*/", line):
        tot[cur] += 1
        for k, p in pats.items():
            if re.search(p, line):
                cnt[cur][k] += 1
print(f"{'kernel':58s} {'instr':>6s} " + " ".join(f"{k:>9s}" for k in pats))
for k in sorted(cnt, key=lambda k: (-cnt[k]['UTC*MMA'], -cnt[k]['HMMA'], k)):
    if tot[k] < 40 and not cnt[k]['UTC*MMA'] and not cnt[k]['HMMA']:
        continue
    print(f"{k[:58]:58s} {tot[k]:6d} " + " ".join(f"{cnt[k][p]:9d}" for p in pats))
