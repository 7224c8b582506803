# A/B of two builds of libtts_b200.so on ONE box: new / old / new.  Usage: bash tools/ab_lib.sh  (expects libtts_b200_old.so)
L=transformertts_b200/libtts_b200.so
run() { python tools/time_ops.py PE 2>&1 | grep -i "PE"; for i in 1 2; do python bench.py --steps 20 --warmup 5 --no-ar --no-cpu --no-extra 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('forward ms', round(d['ms_per_step'],4), 'e2e ms', round(d['e2e']['ms_per_step'],4))"; done; }
echo NEW; run; cp $L /tmp/new.so; cp transformertts_b200/libtts_b200_old.so $L; echo OLD; run; cp /tmp/new.so $L; echo NEW2; run
