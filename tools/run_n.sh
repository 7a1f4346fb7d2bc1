# r02n: sorted scatter stage with shared-memory staging -- equality tests, same-run A/B
python -m pytest tests/test_pt_gpu.py -m gpu -q -x -k "split_shade or coherence_bucketed or fused_wavefront or frames_in_flight" > gpurun_out/r02n_pytest.log 2>&1; tail -8 gpurun_out/r02n_pytest.log
B="python bench.py --steps 30 --no-cpu-baseline --no-extras"
MCRT_SHADE_SPLIT=0 $B > gpurun_out/r02n_single.json 2>gpurun_out/r02n_single.err
MCRT_SHADE_SPLIT=1 MCRT_SCATTER_SORT=0 MCRT_COHERENCE=0 $B > gpurun_out/r02n_split_unsorted.json 2>/dev/null
MCRT_SHADE_SPLIT=1 $B > gpurun_out/r02n_split_sorted.json 2>gpurun_out/r02n_split_sorted.err
MCRT_SHADE_SPLIT=1 $B --frames-in-flight 3 > gpurun_out/r02n_split_sorted_fif3.json 2>/dev/null
for f in gpurun_out/r02n_*.json; do python -c "
import json,sys
try:
  d=json.load(open(sys.argv[1])); print('%-45s %.1f Mp/s %.3f ms e2e %.3f' % (sys.argv[1], d['value']/1e6, d['ms_per_step'], d['e2e']['ms_per_step']), {k: round(v,3) for k,v in d['roofline']['stage_ms_per_step'].items()})
except Exception as e: print(sys.argv[1], 'ERR', e)" $f; done
export MCRT_SHADE_SPLIT=1
ncu --metrics gpu__time_duration.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__inst_executed.sum,smsp__issue_active.avg.pct --clock-control none -k regex:k_shade -s 20 -c 10 --csv --log-file gpurun_out/r02n_shade.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-extras --frames-in-flight 1 --halves 1 > /dev/null 2>&1
grep -v "^==" gpurun_out/r02n_shade.csv | python -c "
import csv,sys
for r in csv.DictReader(sys.stdin): print(r['ID'], r['Kernel Name'][:22], r['Metric Name'][:40], r['Metric Value'])"
