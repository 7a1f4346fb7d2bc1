T8="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511"
T4="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29512"
$T8 bench.py --gpus 8 --steps 30 --warmup 3 > gpurun_out/r02y_n8.json 2> gpurun_out/r02y_n8.err; tail -2 gpurun_out/r02y_n8.err; cut -c1-330 gpurun_out/r02y_n8.json
$T4 bench.py --gpus 4 --steps 30 --warmup 3 > gpurun_out/r02y_n4.json 2> gpurun_out/r02y_n4.err; cut -c1-330 gpurun_out/r02y_n4.json
python bench.py --gpus 1 --steps 30 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/r02y_n1.json 2>/dev/null; cut -c1-330 gpurun_out/r02y_n1.json
