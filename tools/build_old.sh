#!/bin/bash
# Build the library of the last commit as transformertts_b200/libtts_b200_old.so (for tools/ab_*.sh), leaving the
# working tree and its library as they are.
set -e
cd "$(dirname "$0")/.."
cp transformertts_b200/libtts_b200.so /tmp/lib_new.so
git stash -q
python -m transformertts_b200.build --force > /dev/null
cp transformertts_b200/libtts_b200.so transformertts_b200/libtts_b200_old.so
git stash pop -q
cp /tmp/lib_new.so transformertts_b200/libtts_b200.so
touch transformertts_b200/libtts_b200.so
echo "old = $(git log --oneline | head -1)"
