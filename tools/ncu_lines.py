"""Instruction and stall-sample share per source line of one kernel in an .ncu-rep (needs -lineinfo + --import-source on)."""
import collections, csv, subprocess, sys
rep, kern, src = sys.argv[1], sys.argv[2], sys.argv[3]
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name", "regex:" + kern],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
starts = [i for i, r in enumerate(rows) if r and r[0] == 'File Path']
blocks = [(starts[i], starts[i + 1] if i + 1 < len(starts) else len(rows)) for i in range(len(starts))]
want = [b for b in blocks if rows[b[0]][1].endswith(src)]
b = want[int(sys.argv[4]) if len(sys.argv) > 4 else 0]      # which captured launch of that kernel
blk = rows[b[0]:b[1]]
hdr = blk[2]
ie = hdr.index('Instructions Executed'); ws = hdr.index('Warp Stall Sampling (All Samples)')
lines = open(rows[b[0]][1]).read().split('\n')
agg, cur, seen = collections.OrderedDict(), None, set()
for r in blk[3:]:
    if r[0] != '':
        cur = int(r[0]); continue
    if cur is None or len(r) <= ie: continue
    key = (cur, r[2])
    if key in seen: continue
    seen.add(key)
    try: inst = float(r[ie] or 0); samp = float(r[ws] or 0)
    except ValueError: continue
    a = agg.setdefault(cur, [0, 0, 0]); a[0] += inst; a[1] += samp; a[2] += 1
ti = sum(a[0] for a in agg.values()); ts = sum(a[1] for a in agg.values())
print("total warp-instructions %.4g, stall samples %.4g" % (ti, ts))
for ln, a in sorted(agg.items()):
    if a[0] / ti > 0.006 or a[1] / ts > 0.006:
        print("%4d sass=%3d inst=%5.1f%% stall=%5.1f%% | %s" % (ln, a[2], 100 * a[0] / ti, 100 * a[1] / ts, lines[ln - 1].strip()[:100]))
