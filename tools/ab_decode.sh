#!/bin/bash
# A/B of two builds of libtts_b200.so on ONE box (new / old / new) for the AR decode (1000 frames).
cd "$(dirname "$0")/.."
L=transformertts_b200/libtts_b200.so
run() { for a in "1 1001" "16 1001" "32 1001" "64 1001" "128 1001"; do python tools/profile_decode.py $a 2>&1 | tail -1; done; }
echo NEW; run; cp $L /tmp/new.so; cp transformertts_b200/libtts_b200_old.so $L; echo OLD; run; cp /tmp/new.so $L; echo NEW2; run
