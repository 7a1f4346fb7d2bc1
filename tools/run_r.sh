# r02r: final build of the round -- whole gpu suite, default bench line, launch list, per-kernel counters, one full-frame ncu capture
python -m pytest tests -m gpu -x -q -s > gpurun_out/r02r_pytest.log 2>&1; grep -n "C1 image RMSE\|records,\|visibility differs" gpurun_out/r02r_pytest.log | head; tail -4 gpurun_out/r02r_pytest.log
python tools/measure_counters.py > gpurun_out/r02r_counters.log 2>&1; tail -3 gpurun_out/r02r_counters.log; cp gpurun_out/kernel_counters.json profiles/kernel_counters.json
python bench.py > gpurun_out/r02r_bench.json 2> gpurun_out/r02r_bench.err; tail -2 gpurun_out/r02r_bench.err; cut -c1-1200 gpurun_out/r02r_bench.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02r_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/r02r_ncu.log 2>&1
python tools/launch_summary.py gpurun_out/r02r_launches.csv
ncu --set full --clock-control none --import-source on -k "regex:k_trace|k_shade" -s 48 -c 16 -o gpurun_out/r02r_frame -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-extras --frames-in-flight 1 --halves 1 > gpurun_out/r02r_full.log 2>&1; tail -1 gpurun_out/r02r_full.log
python tools/bench_configs.py C1 C2 C4 C5 > gpurun_out/r02_configs.jsonl 2>/dev/null; MCRT_FRAMES_IN_FLIGHT=3 python tools/bench_configs.py C1 C2 C5 > gpurun_out/r02_configs_fif3.jsonl 2>/dev/null; cut -c1-40,225-300 gpurun_out/r02_configs.jsonl gpurun_out/r02_configs_fif3.jsonl
