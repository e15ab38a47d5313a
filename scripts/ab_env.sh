#!/bin/bash
# Same-box A/B of an environment switch:  scripts/ab_env.sh EVR_NO_TAILSPLIT [rounds]   (alternating encode benches)
cd "$(dirname "$0")/.."
for round in $(seq 1 ${2:-3}); do
  for v in 1 0; do
    env $1=$v python bench.py --workload encode --no-cpu-baseline --steps 40 --warmup 10 > gpurun_out/ab_env$v.json 2>/dev/null
    echo -n "$1=$v: "; python scripts/show_bench.py gpurun_out/ab_env$v.json | head -1
  done
done
