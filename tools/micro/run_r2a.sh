set -x
cd tools/micro
nvidia-smi --query-gpu=clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active --format=csv -lms 100 > ../../gpurun_out/r2a_clocks.csv &
SMI=$!
./pipe_rates > ../../gpurun_out/r2a_pipe_rates.txt 2>&1
./softmax_step_bench > ../../gpurun_out/r2a_softmax_step.txt 2>&1
for p in 0 2 3 4; do ./tc_time $p 28 9216 200 >> ../../gpurun_out/r2a_tc_time.txt 2>&1; done
./tc_timeline 0 > ../../gpurun_out/r2a_tc_timeline.txt 2>&1
kill $SMI
cd ../..
python bench.py --steps 20 --warmup 5 > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err
tail -3 gpurun_out/r2a_tc_time.txt
