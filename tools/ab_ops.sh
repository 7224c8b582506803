#!/bin/bash
# A/B of two builds of libtts_b200.so on ONE box (new / old / new) for operator shapes of tools/time_ops.py.
# Usage: bash tools/ab_ops.sh <filter>   (expects transformertts_b200/libtts_b200_old.so)
cd "$(dirname "$0")/.."
L=transformertts_b200/libtts_b200.so
run() { python tools/time_ops.py "$1" 2>&1 | grep -v "^sum"; }
echo NEW; run "$1"; cp $L /tmp/new.so; cp transformertts_b200/libtts_b200_old.so $L; echo OLD; run "$1"; cp /tmp/new.so $L; echo NEW2; run "$1"
