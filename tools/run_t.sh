# r02t: cooperative tail -- equality tests, then its effect on C3 (bench) and on a 1/8 tile partition of C5
python -m pytest tests/test_coop_tail_gpu.py -m gpu -q -x > gpurun_out/r02t_pytest.log 2>&1; tail -12 gpurun_out/r02t_pytest.log
B="python bench.py --steps 30 --no-cpu-baseline --no-extras"
MCRT_COOP_ITERS=0 $B > gpurun_out/r02t_off.json 2>/dev/null
MCRT_COOP_ITERS=128 $B > gpurun_out/r02t_c128.json 2>/dev/null
MCRT_COOP_ITERS=64 $B > gpurun_out/r02t_c64.json 2>/dev/null
MCRT_COOP_ITERS=0 $B --frames-in-flight 1 > gpurun_out/r02t_off_fif1.json 2>/dev/null
MCRT_COOP_ITERS=128 $B --frames-in-flight 1 > gpurun_out/r02t_c128_fif1.json 2>/dev/null
for f in gpurun_out/r02t_*.json; do python -c "
import json,sys
try:
  d=json.load(open(sys.argv[1])); print('%-45s %.1f Mp/s %.3f ms e2e %.3f' % (sys.argv[1], d['value']/1e6, d['ms_per_step'], d['e2e']['ms_per_step']), {k: round(v,3) for k,v in d['roofline']['stage_ms_per_step'].items()})
except Exception as e: print(sys.argv[1], 'ERR', e)" $f; done
# what one of 8 GPUs does under the tile partition of C5 (10 M triangles, 4K), one frame at a time
for c in 0 32 64 128 256; do MCRT_COOP_ITERS=$c python tools/bench_c5_tiles.py --parts 8 --frames-in-flight 1 --steps 12 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print('C5 tiles 1/8, coop $c:', {k: d[k] for k in d if k in ('value','ms_per_step','mrays_per_s','frames_in_flight')})"; done
MCRT_COOP_ITERS=128 python tools/bench_c5_tiles.py --parts 8 --frames-in-flight 3 --steps 12 2>/dev/null | cut -c1-300
MCRT_COOP_ITERS=0 python tools/bench_c5_tiles.py --parts 8 --frames-in-flight 3 --steps 12 2>/dev/null | cut -c1-300
