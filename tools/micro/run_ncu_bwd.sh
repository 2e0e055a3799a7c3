export TVI_BWD_ONLY_TC=1
timeout 200 python tools/time_spatial_bwd.py > gpurun_out/r2i_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:attn_bwd_tc_kernel -s 4 -c 2 -f -o gpurun_out/r2i_bwd python tools/time_spatial_bwd.py > gpurun_out/r2i_ncu.log 2>&1
tail -2 gpurun_out/r2i_ncu.log
