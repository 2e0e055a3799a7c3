timeout 300 python -m pytest tests/test_gpu_backward.py -x -q -k "spatial_bwd" > gpurun_out/r2h_bwd_tests.log 2>&1; echo "rc $?"
tail -3 gpurun_out/r2h_bwd_tests.log
TVI_BWD_ONLY_TC=1 timeout 200 python tools/time_spatial_bwd.py
