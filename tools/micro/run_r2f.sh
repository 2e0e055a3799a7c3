timeout 900 python -m pytest tests -m gpu -x -q -s -k "feed_forward or block or spatio or model" > gpurun_out/r2f_pytest_tol.log 2>&1; echo "pytest rc $?" >> gpurun_out/r2f_pytest_tol.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r2f_bench.json 2> gpurun_out/r2f_bench.err; echo "bench rc $?"
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2f_smoke.log 2>&1; echo "smoke rc $?"
grep -c "max abs err" gpurun_out/r2f_pytest_tol.log; tail -3 gpurun_out/r2f_pytest_tol.log; tail -3 gpurun_out/r2f_smoke.log; tail -5 gpurun_out/r2f_bench.err
