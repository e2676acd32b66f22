import csv, sys, subprocess
rep=sys.argv[1]; nq=float(sys.argv[2]) if len(sys.argv)>2 else 1048576
out=subprocess.run(['ncu','-i',rep,'--page','source','--print-source','cuda,sass','--csv'],capture_output=True,text=True).stdout
rows=list(csv.reader(out.splitlines()))
for i,r in enumerate(rows[:10]):
    if 'Instructions Executed' in r: hi=i; break
hdr=rows[hi]; ia=hdr.index('Instructions Executed'); isamp=hdr.index('# Samples')
lines=[]
for r in rows[hi+1:]:
    if r[0].isdigit():
        try: lines.append((int(r[ia]), int(r[isamp]), int(r[0]), r[1]))
        except: pass
tot=sum(l[0] for l in lines); ts=sum(l[1] for l in lines)
print('total warp-instr', tot, 'per query', tot/nq, 'samples', ts)
for l in sorted(lines, reverse=True)[:int(sys.argv[3]) if len(sys.argv)>3 else 30]:
    print(f'{100*l[0]/tot:5.1f}% inst {100*l[1]/ts:5.1f}% samp  L{l[2]:4d}: {l[3][:120]}')
raw=subprocess.run(['ncu','-i',rep,'--page','raw','--csv'],capture_output=True,text=True).stdout
rows=list(csv.reader(raw.splitlines())); h=rows[0]
for w in ['gpu__time_duration.sum','dram__bytes_read.sum','dram__bytes_write.sum','gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed','sm__warps_active.avg.pct_of_peak_sustained_active','smsp__issue_active.avg.pct_of_peak_sustained_active','l1tex__t_sector_hit_rate.pct','lts__t_sector_hit_rate.pct','smsp__thread_inst_executed_per_inst_executed.ratio','sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active','launch__registers_per_thread','launch__grid_size','lts__t_bytes.sum','l1tex__t_bytes.sum']:
    if w in h: print(w, rows[2][h.index(w)], rows[1][h.index(w)])
