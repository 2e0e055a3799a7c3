cd tools/micro
O=../../gpurun_out/r2b_tc2.txt
: > $O
for cfg in "3 1111" "200 300" "28 9216"; do
  set -- $cfg
  echo "== dbg batch $1 S $2" >> $O
  timeout 60 ./tc_time_dbg 0 $1 $2 3 3 >> $O 2>&1 || echo "rc $?" >> $O
done
if grep -q "TIMEOUT\|rc 1" $O; then echo "debug run failed; release skipped" >> $O; else
for st in 1 0; do
for p in 0 4 5 6; do
  echo "== release poly $p stagger $st" >> $O
  TVI_TC2_STAGGER=$st timeout 30 ./tc_time $p 28 9216 100 3 >> $O 2>&1 || { echo "rc $?" >> $O; break; }
done
done
fi
grep -v "WAIT TIMEOUT" $O | grep -v checksum | tail -40
grep "TIMEOUT" $O | awk '{print $6,$7,$8,$9,$10,$11,$12,$13}' | sort | uniq -c | head
