# r02l: where the split shade stage spends its time -- launch list, then one full ncu capture of each new kernel
export MCRT_SHADE_SPLIT=1 MCRT_COHERENCE=0
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02l_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/r02l_ncu.log 2>&1; tail -2 gpurun_out/r02l_ncu.log
python tools/launch_summary.py gpurun_out/r02l_launches.csv
ncu --set full --clock-control none --import-source on -k regex:k_shade_s -s 10 -c 4 -o gpurun_out/r02l_split -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-extras --frames-in-flight 1 --halves 1 > gpurun_out/r02l_full.log 2>&1; tail -2 gpurun_out/r02l_full.log
ls -la gpurun_out/*.ncu-rep
