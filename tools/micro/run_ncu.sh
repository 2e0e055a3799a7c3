cd tools/micro
export TVI_TC2_STAGGER=0
./tc_time 4 28 9216 3 3 > ../../gpurun_out/r2d_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:attn_spatial_tc2 -s 3 -c 1 -f -o ../../gpurun_out/r2d_tc2_p4 ./tc_time 4 28 9216 3 3 > ../../gpurun_out/r2d_ncu.log 2>&1
tail -3 ../../gpurun_out/r2d_ncu.log
