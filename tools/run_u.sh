# r02u: where a 1/8 tile partition of C5 spends a frame (one frame at a time), without and with the cooperative tail
for c in 0 64; do
MCRT_COOP_ITERS=$c ncu --metrics gpu__time_duration.sum,sm__cycles_active.min,sm__cycles_active.avg,sm__cycles_active.max,smsp__inst_executed.sum --clock-control none -s 60 -c 60 --csv --log-file gpurun_out/r02u_c$c.csv python tools/bench_c5_tiles.py --parts 8 --frames-in-flight 1 --steps 2 --warmup 2 > /dev/null 2>&1
grep -v "^==" gpurun_out/r02u_c$c.csv | python -c "
import csv,sys,collections
rows=collections.OrderedDict()
for r in csv.DictReader(sys.stdin):
    rows.setdefault((r['ID'], r['Kernel Name'].split('(')[0].replace('void ','')), {})[r['Metric Name']] = float(r['Metric Value'].replace(',',''))
tot=collections.defaultdict(float)
for (i,k),m in rows.items():
    tot[k]+=m['gpu__time_duration.sum']
    if 'k_trace' in k or 'k_shade' in k: print(i, '%-22s %8.1f us  cyc min/avg/max %8.0f %8.0f %8.0f  insts %.3g' % (k, m['gpu__time_duration.sum']/1e3, m['sm__cycles_active.min'], m['sm__cycles_active.avg'], m['sm__cycles_active.max'], m['smsp__inst_executed.sum']))
print('coop $c totals (us):', {k: round(v/1e3,1) for k,v in tot.items()})
"
done
