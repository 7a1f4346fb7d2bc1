for c in 0 -64 -256 64; do MCRT_COOP_ITERS=$c python tools/bench_c5_tiles.py --parts 8 --frames-in-flight 1 --steps 12 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print('C5 tiles 1/8, coop $c:', {k: d[k] for k in d if k in ('ms_per_step','mrays_per_s')})"; done
