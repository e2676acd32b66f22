"""Per source line of gather.cu: how often it ran (max over its SASS instructions of 'Instructions Executed') per query.
python profiles/ncu_linecounts.py rep nq lo hi"""
import csv, sys, subprocess
rep = sys.argv[1]; nq = float(sys.argv[2]); lo = int(sys.argv[3]); hi = int(sys.argv[4])
out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--print-source', 'cuda,sass', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = [r for r in rows[:10] if 'Instructions Executed' in r][0]; ia = hdr.index('Instructions Executed')
cur = None; line = None; src = {}; mx = {}; n = {}; tot = {}
for r in rows:
    if r and r[0] == 'File Path': cur = r[1]; continue
    if not r or cur is None or not cur.endswith('gather.cu'): continue
    if r[0].isdigit(): line = int(r[0]); src[line] = r[1]; tot[line] = int(r[ia]); continue
    if r[0] == '' and line is not None and len(r) > ia and r[ia].isdigit():
        mx[line] = max(mx.get(line, 0), int(r[ia])); n[line] = n.get(line, 0) + 1
for l in sorted(src):
    if lo <= l <= hi: print('L%3d runs/query %7.2f  sass %3d  instr/query %7.1f | %s' % (l, mx.get(l, 0) / nq, n.get(l, 0), tot[l] / nq, src[l][:110]))
