# r02h: the new no-alloc entry point alone in a fresh process, the WHOLE gpu suite (full-size parity included),
# the full default bench line, the ncu launch list of the same command, the per-kernel counters of this build
python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/r02h_noalloc_alone.log 2>&1; tail -3 gpurun_out/r02h_noalloc_alone.log
python -m pytest tests -m gpu -x -q --durations=15 > gpurun_out/r02h_pytest.log 2>&1; tail -25 gpurun_out/r02h_pytest.log
python bench.py > gpurun_out/r02h_bench.json 2> gpurun_out/r02h_bench.err; tail -2 gpurun_out/r02h_bench.err; cat gpurun_out/r02h_bench.json
python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/r02h_ref.json 2> gpurun_out/r02h_ref.err; cat gpurun_out/r02h_ref.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02h_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/r02h_ncu.log 2>&1; tail -2 gpurun_out/r02h_ncu.log
python tools/measure_counters.py > gpurun_out/r02h_counters.log 2>&1; tail -5 gpurun_out/r02h_counters.log
