# r02j (2 GPUs): the two-rank NCCL merge through the C ABI, bench.py under torchrun (own arm + reference arm)
python -m pytest tests/test_round2_gpu.py -m gpu -q -k "nccl or merge" > gpurun_out/r02j_pytest.log 2>&1; tail -5 gpurun_out/r02j_pytest.log
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
$T bench.py --gpus 2 --steps 30 --warmup 3 --no-extras > gpurun_out/r02j_n2.json 2> gpurun_out/r02j_n2.err; tail -3 gpurun_out/r02j_n2.err; cat gpurun_out/r02j_n2.json
python bench.py --gpus 1 --steps 30 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/r02j_n1.json 2> gpurun_out/r02j_n1.err; cat gpurun_out/r02j_n1.json
$T bench.py --impl reference --gpus 2 --steps 1 --warmup 0 > gpurun_out/r02j_ref_n2.json 2> gpurun_out/r02j_ref_n2.err; cat gpurun_out/r02j_ref_n2.json
