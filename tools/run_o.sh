export MCRT_SHADE_SPLIT=1
ncu --set full --clock-control none --import-source on -k regex:k_shade_scatter -s 5 -c 2 -o gpurun_out/r02o_sorted -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-extras --frames-in-flight 1 --halves 1 > gpurun_out/r02o_full.log 2>&1; tail -2 gpurun_out/r02o_full.log
