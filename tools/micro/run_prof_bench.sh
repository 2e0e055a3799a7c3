# B200_PROFILING.md recipe on the bench command: plain run, then the launch list, then the dominant kernel with --set full
python bench.py --steps 2 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/r2k_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2k_launches.csv python bench.py --steps 2 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/r2k_ncu1.log 2>&1
python bench.py --steps 1 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/r2k_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:attn_spatial_tc -s 3 -c 1 -f -o gpurun_out/r2k_spatial python bench.py --steps 1 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/r2k_ncu2.log 2>&1
tail -2 gpurun_out/r2k_ncu2.log
timeout 200 python tools/micro/flowtime.py 2>&1 | grep "warp\|fewkeys"
