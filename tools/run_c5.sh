T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29513"
$T tools/bench_c5_tiles.py --steps 16 --frames-in-flight 1 --verify 2>/dev/null | tail -1 | cut -c1-600
$T tools/bench_c5_tiles.py --steps 16 --frames-in-flight 3 2>/dev/null | tail -1 | cut -c1-600
$T tools/bench_c5_tiles.py --steps 16 --frames-in-flight 3 --partition samples 2>/dev/null | tail -1 | cut -c1-600
python tools/bench_c5_tiles.py --steps 8 --frames-in-flight 1 2>/dev/null | tail -1 | cut -c1-400
python tools/bench_c5_tiles.py --steps 8 --frames-in-flight 3 2>/dev/null | tail -1 | cut -c1-400
