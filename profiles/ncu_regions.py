"""Sum ncu source-page counters over line ranges of gather.cu: python profiles/ncu_regions.py rep nq name:lo-hi ..."""
import csv, sys, subprocess
rep = sys.argv[1]; nq = float(sys.argv[2])
out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--print-source', 'cuda,sass', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
for i, r in enumerate(rows[:10]):
    if 'Instructions Executed' in r: hi = i; break
hdr = rows[hi]; ia = hdr.index('Instructions Executed'); isamp = hdr.index('# Samples')
lines = {}; other = [0, 0]; cur = None
for r in rows:
    if r and r[0] == 'File Path': cur = r[1]; continue
    if r and r[0].isdigit():
        try: v = (int(r[ia]), int(r[isamp]))
        except Exception: continue
        if cur.endswith('gather.cu'): lines[int(r[0])] = v
        else: other[0] += v[0]; other[1] += v[1]
lines[100000] = tuple(other)          # everything attributed to CUDA headers (intrinsics)
tot = sum(v[0] for v in lines.values()); ts = sum(v[1] for v in lines.values())
print('total warp-instr per query %.0f' % (tot / nq))
for spec in sys.argv[3:]:
    name, rng = spec.split(':'); lo, hi2 = map(int, rng.split('-'))
    a = sum(v[0] for l, v in lines.items() if lo <= l <= hi2); s = sum(v[1] for l, v in lines.items() if lo <= l <= hi2)
    print('%-12s L%d-%d: %6.0f instr/query %5.1f%% inst %5.1f%% samples' % (name, lo, hi2, a / nq, 100 * a / tot, 100 * s / ts))
