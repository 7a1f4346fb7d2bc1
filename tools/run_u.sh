# r02u2: host builder A/B -- optimal (DP) collapse + optimal slot assignment vs greedy, same box
python -m pytest tests/test_rr_gpu.py tests/test_two_level_gpu.py tests/test_pt_gpu.py -m gpu -x -q > gpurun_out/r02v_pytest.log 2>&1; tail -3 gpurun_out/r02v_pytest.log
B="python bench.py --steps 60 --no-cpu-baseline --no-extras"
MCRT_BVH_COLLAPSE=greedy MCRT_BVH_SLOTS=greedy $B > gpurun_out/r02v_greedy.json 2>/dev/null
$B > gpurun_out/r02v_optimal.json 2>/dev/null
MCRT_BVH_COLLAPSE=greedy $B > gpurun_out/r02v_slots_only.json 2>/dev/null
MCRT_BVH_COLLAPSE=greedy MCRT_BVH_SLOTS=greedy $B > gpurun_out/r02v_greedy2.json 2>/dev/null
$B > gpurun_out/r02v_optimal2.json 2>/dev/null
for f in gpurun_out/r02v_*.json; do python -c "
import json,sys
try:
  d=json.load(open(sys.argv[1])); print('%-40s %.1f Mp/s %.3f ms e2e %.3f' % (sys.argv[1], d['value']/1e6, d['ms_per_step'], d['e2e']['ms_per_step']), {k: round(v,3) for k,v in d['roofline']['stage_ms_per_step'].items()})
except Exception as e: print(sys.argv[1], 'ERR', e)" $f; done
