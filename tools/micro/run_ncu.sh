cd tools/micro
./tc_time 4 28 9216 3 > ../../gpurun_out/r2d_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:attn_spatial_tc -s 3 -c 1 -f -o ../../gpurun_out/r2g_tc_singleS ./tc_time 4 28 9216 3 > ../../gpurun_out/r2d_ncu.log 2>&1
tail -2 ../../gpurun_out/r2d_ncu.log
