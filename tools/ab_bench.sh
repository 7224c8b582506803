#!/bin/bash
# A/B of two builds of libtts_b200.so on ONE box (new / old / new / old): device ms per cfg-2 forward.
cd "$(dirname "$0")/.."
L=transformertts_b200/libtts_b200.so
run() { python bench.py --steps 30 --warmup 5 --no-ar --no-cpu --no-extra 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['ms_per_step'],4), round(d['e2e']['ms_per_step'],4))"; }
cp $L /tmp/new.so
for i in 1 2; do echo -n "NEW "; run; cp transformertts_b200/libtts_b200_old.so $L; echo -n "OLD "; run; cp /tmp/new.so $L; done
