cd tools/micro
for p in 2 3 4 5; do echo "== poly $p"; timeout 30 ./tc_time $p 28 9216 30 | grep -v "checksum\|vs mma"; done
cd ../..
timeout 300 python bench.py --steps 20 --warmup 5 --no-extras --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['roofline']['ms_per_launch'], d['roofline']['achieved'], d['e2e']['value'])"
