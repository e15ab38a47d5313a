#!/bin/bash
# Same-box A/B of two builds of libevr.so on any probe script:  scripts/ab_any.sh <libA> <libB> <script.py> [rounds]
cd "$(dirname "$0")/.."
L=$PWD/$(ls -d event-*_b200)/lib
for round in $(seq 1 ${4:-2}); do
  for v in $1 $2; do
    echo "== $v (round $round)"
    EVR_LIB_PATH=$L/libevr_$v.so python $3 2>&1 | tail -${AB_TAIL:-14}
  done
done
