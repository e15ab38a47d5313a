#!/bin/bash
# Round-2 evidence (run under gpurun, one GPU): GPU tests, smoke, the driver's bench commands, then the ncu passes --
# each ncu run follows the same command run plain (B200_PROFILING.md).  Outputs land in gpurun_out/ with the r02_<tag>_
# prefix; summaries are copied into profiles/ here afterwards (scripts/ncu_summary.py).
#   bash scripts/ncu_round2.sh <tag> [quick]
cd "$(dirname "$0")/.."
T=${1:-g}
O=gpurun_out
P=r02_${T}
M="--metrics gpu__time_duration.sum --clock-control none --csv"
if [ "$2" != "quick" ]; then
  timeout 1500 python -m pytest tests -m gpu -x -q > $O/${P}_gpu_tests.txt 2>&1; tail -3 $O/${P}_gpu_tests.txt
  python -c "import __graft_entry__ as g; g.smoke()" > $O/${P}_smoke.txt 2>&1; tail -2 $O/${P}_smoke.txt
fi
python bench.py > $O/${P}_bench.json 2> $O/${P}_bench.err
python bench.py --impl reference --steps 2 --warmup 1 > $O/${P}_bench_reference.json 2> $O/${P}_bench_reference.err
python scripts/prof_case.py encode 2 > $O/plain_encode.log 2>&1 &&
ncu $M --log-file $O/${P}_launches_encode.csv python scripts/prof_case.py encode 2 > $O/ncu_encode.log 2>&1
python scripts/prof_case.py scan_small 2 > $O/plain_scan_small.log 2>&1 &&
ncu $M -k regex:topk --log-file $O/${P}_launches_scan_small.csv python scripts/prof_case.py scan_small 2 > $O/ncu_scan_small.log 2>&1
python scripts/prof_case.py scan_c3 2 > $O/plain_scan_c3.log 2>&1 &&
ncu $M -k regex:topk --log-file $O/${P}_launches_scan_c3.csv python scripts/prof_case.py scan_c3 2 > $O/ncu_scan_c3.log 2>&1
python scripts/prof_case.py encode 1 > $O/plain_encode1.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gemm_tc -s 5 -c 4 -o $O/${P}_prof_gemm -f python scripts/prof_case.py encode 1 > $O/ncu_gemm.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:attention_tc -s 1 -c 1 -o $O/${P}_prof_attn_b32 -f python scripts/prof_case.py encode 1 > $O/ncu_attn_b32.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:layernorm -s 3 -c 1 -o $O/${P}_prof_ln -f python scripts/prof_case.py encode 1 > $O/ncu_ln.log 2>&1
python scripts/prof_case.py scan_small 1 > $O/plain_scan_small1.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:topk_scan -s 1 -c 1 -o $O/${P}_prof_scan_small -f python scripts/prof_case.py scan_small 1 > $O/ncu_scan_small_full.log 2>&1
bash scripts/scan_l2_breakdown.sh c3_${T} 10000 > $O/${P}_scan_c3_l2.txt 2>&1
ls -la $O | grep ${P}
