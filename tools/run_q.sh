# r02q: the split shade stage with the scatter kernel at 3 x 256 threads per SM -- tests, coherence and occupancy A/B
python -m pytest tests/test_pt_gpu.py -m gpu -q -x -k "split_shade" > gpurun_out/r02q_pytest.log 2>&1; tail -3 gpurun_out/r02q_pytest.log
B="python bench.py --steps 30 --no-cpu-baseline --no-extras"
MCRT_SHADE_SPLIT=0 $B > gpurun_out/r02q_single.json 2>/dev/null
MCRT_SHADE_SPLIT=1 MCRT_COHERENCE=0 $B > gpurun_out/r02q_split_coh0.json 2>/dev/null
MCRT_SHADE_SPLIT=1 MCRT_COHERENCE=3 $B > gpurun_out/r02q_split_coh3.json 2>gpurun_out/r02q_split_coh3.err
MCRT_SHADE_SPLIT=1 MCRT_COHERENCE=0 MCRT_LIB_VARIANT=scat4 MCRT_SCAT_CTAS_PER_SM=4 $B > gpurun_out/r02q_scat4.json 2>/dev/null
MCRT_SHADE_SPLIT=1 MCRT_COHERENCE=0 MCRT_LIB_VARIANT=scat128x6 MCRT_SCAT_CTAS_PER_SM=6 $B > gpurun_out/r02q_scat128x6.json 2>/dev/null
MCRT_SHADE_SPLIT=1 MCRT_COHERENCE=0 $B --frames-in-flight 3 > gpurun_out/r02q_split_coh0_fif3.json 2>/dev/null
MCRT_SHADE_SPLIT=1 MCRT_COHERENCE=0 $B --frames-in-flight 1 > gpurun_out/r02q_split_coh0_fif1.json 2>/dev/null
for f in gpurun_out/r02q_*.json; do python -c "
import json,sys
try:
  d=json.load(open(sys.argv[1])); print('%-45s %.1f Mp/s %.3f ms e2e %.3f' % (sys.argv[1], d['value']/1e6, d['ms_per_step'], d['e2e']['ms_per_step']), {k: round(v,3) for k,v in d['roofline']['stage_ms_per_step'].items()})
except Exception as e: print(sys.argv[1], 'ERR', e)" $f; done
