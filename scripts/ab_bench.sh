#!/bin/bash
# Same-box A/B of two builds of libevr.so (EVR_LIB_PATH): alternating encode-bench runs, three rounds.
cd "$(dirname "$0")/.."
L=$(ls -d event-*_b200)/lib
for round in 1 2 3; do
  for v in old new; do
    EVR_LIB_PATH=$PWD/$L/libevr_$v.so python bench.py --workload encode --no-cpu-baseline --steps 40 --warmup 10 > gpurun_out/ab_$v.json 2>/dev/null
    echo -n "$v: "; python scripts/show_bench.py gpurun_out/ab_$v.json | cut -c1-60
  done
done
