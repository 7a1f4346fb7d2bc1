# r02k: split shade stage -- equality test, then same-run A/B against the single kernel
python -m pytest tests/test_pt_gpu.py -m gpu -q -x -k "split_shade or coherence_bucketed or fused_wavefront" > gpurun_out/r02k_pytest.log 2>&1; tail -15 gpurun_out/r02k_pytest.log
python -m pytest tests/test_bdpt_strategies_gpu.py -m gpu -q -x -s 2>&1 | grep "frame\|passed\|failed" 
B="python bench.py --steps 30 --no-cpu-baseline --no-extras"
MCRT_SHADE_SPLIT=0 $B > gpurun_out/r02k_single.json 2>gpurun_out/r02k_single.err
MCRT_SHADE_SPLIT=1 $B > gpurun_out/r02k_split.json 2>gpurun_out/r02k_split.err
MCRT_SHADE_SPLIT=1 MCRT_COHERENCE=0 $B > gpurun_out/r02k_split_coh0.json 2>/dev/null
MCRT_SHADE_SPLIT=1 MCRT_SURF_CTAS_PER_SM=6 $B > gpurun_out/r02k_split_surf6.json 2>/dev/null
MCRT_SHADE_SPLIT=1 $B --frames-in-flight 3 > gpurun_out/r02k_split_fif3.json 2>/dev/null
MCRT_SHADE_SPLIT=0 $B --frames-in-flight 3 > gpurun_out/r02k_single_fif3.json 2>/dev/null
for f in gpurun_out/r02k_*.json; do python -c "
import json,sys
try:
  d=json.load(open(sys.argv[1])); print('%-45s %.1f Mp/s %.3f ms e2e %.3f' % (sys.argv[1], d['value']/1e6, d['ms_per_step'], d['e2e']['ms_per_step']), {k: round(v,3) for k,v in d['roofline']['stage_ms_per_step'].items()})
except Exception as e: print(sys.argv[1], 'ERR', e)" $f; done
