timeout 300 python -m pytest tests/test_gpu_backward.py -x -q -s -k "spatial_bwd" > gpurun_out/r2h_bwd_tests.log 2>&1; echo "rc $?"
tail -15 gpurun_out/r2h_bwd_tests.log
