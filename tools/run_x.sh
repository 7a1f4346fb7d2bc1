python -c "import __graft_entry__ as g; g.smoke(); print('SMOKE OK')" 2>&1 | tail -3
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
$T bench.py --gpus 2 --steps 30 --warmup 3 > gpurun_out/r02x_n2.json 2> gpurun_out/r02x_n2.err; tail -2 gpurun_out/r02x_n2.err; cut -c1-400 gpurun_out/r02x_n2.json
$T bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/r02x_ref_n2.json 2> gpurun_out/r02x_ref_n2.err; cut -c1-300 gpurun_out/r02x_ref_n2.json
python -m pytest tests/test_round2_gpu.py -m gpu -q -k "nccl" 2>&1 | tail -2
