cd tools/micro
O=../../gpurun_out/r2b_tc2.txt
: > $O
for cfg in "3 1111" "40 1111" "200 300" "1 64" "28 9216"; do
  set -- $cfg
  echo "== dbg batch $1 S $2" >> $O
  timeout 20 ./tc_time_dbg 4 $1 $2 3 >> $O 2>&1 || echo "rc $?" >> $O
done
if grep -q "TIMEOUT\|rc 1" $O; then echo "debug run failed; release skipped" >> $O; else
for p in 0 4 5 6; do
  echo "== release poly $p" >> $O
  timeout 30 ./tc_time $p 28 9216 100 >> $O 2>&1 || { echo "rc $?" >> $O; break; }
done
timeout 60 ./tc_timeline 4 28 9216 5 > ../../gpurun_out/r2c_tc2_timeline.txt 2>&1
fi
grep -v "WAIT TIMEOUT" $O | grep -v checksum | tail -40
grep "TIMEOUT" $O | awk '{print $6,$7,$8,$9,$10,$11,$12,$13}' | sort | uniq -c | head
