#!/bin/bash
# rows-per-cluster x cluster-size sweep of the AR decode kernel (1000 frames), behind decode_plan()'s cost model
cd "$(dirname "$0")/.."
for a in "1 1001 1 16" "1 1001 1 8" "1 1001 1 4" "16 1001 0 0" "16 1001 1 8" "16 1001 1 16" "16 1001 2 16" \
         "32 1001 0 0" "32 1001 1 8" "32 1001 2 8" "32 1001 2 16" "32 1001 1 4" "32 1001 4 16" \
         "64 1001 0 0" "64 1001 2 8" "64 1001 4 8" "64 1001 2 4" "64 1001 4 16" "64 1001 8 8" \
         "128 1001 0 0" "128 1001 4 8" "128 1001 8 8" "128 1001 4 4" "128 1001 8 4"; do
  timeout 120 python tools/profile_decode.py $a 2>&1 | tail -1
done
