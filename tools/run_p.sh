# r02p: sorted scatter with pulled runs; scatter at higher occupancy (variants, unsorted)
python -m pytest tests/test_pt_gpu.py -m gpu -q -x -k "split_shade" > gpurun_out/r02p_pytest.log 2>&1; tail -3 gpurun_out/r02p_pytest.log
B="python bench.py --steps 30 --no-cpu-baseline --no-extras"
MCRT_SHADE_SPLIT=1 MCRT_SCATTER_SORT=0 MCRT_COHERENCE=0 $B > gpurun_out/r02p_split_unsorted.json 2>/dev/null
MCRT_SHADE_SPLIT=1 $B > gpurun_out/r02p_split_sorted.json 2>gpurun_out/r02p_split_sorted.err
MCRT_SHADE_SPLIT=1 MCRT_COHERENCE=0 $B > gpurun_out/r02p_split_sorted_coh0.json 2>/dev/null
MCRT_LIB_VARIANT=b256x3 MCRT_SHADE_SPLIT=1 MCRT_SCATTER_SORT=0 MCRT_COHERENCE=0 MCRT_SHADE_CTAS_PER_SM=3 $B > gpurun_out/r02p_b256x3.json 2>/dev/null
MCRT_LIB_VARIANT=b256x2 MCRT_SHADE_SPLIT=1 MCRT_SCATTER_SORT=0 MCRT_COHERENCE=0 MCRT_SHADE_CTAS_PER_SM=2 $B > gpurun_out/r02p_b256x2.json 2>/dev/null
for f in gpurun_out/r02p_*.json; do python -c "
import json,sys
try:
  d=json.load(open(sys.argv[1])); print('%-45s %.1f Mp/s %.3f ms e2e %.3f' % (sys.argv[1], d['value']/1e6, d['ms_per_step'], d['e2e']['ms_per_step']), {k: round(v,3) for k,v in d['roofline']['stage_ms_per_step'].items()})
except Exception as e: print(sys.argv[1], 'ERR', e)" $f; done
