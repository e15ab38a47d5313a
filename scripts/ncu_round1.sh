#!/bin/bash
# ncu evidence for round 1 (run under gpurun; every ncu run is preceded by the same command run plain).
set -x
cd "$(dirname "$0")/.."
O=gpurun_out
M="--metrics gpu__time_duration.sum --clock-control none --csv"
python scripts/prof_case.py encode 2 > $O/plain_encode.log 2>&1 &&
ncu $M --log-file $O/launches_encode.csv python scripts/prof_case.py encode 2 > $O/ncu_encode.log 2>&1
python scripts/prof_case.py scan_big 2 > $O/plain_scan_big.log 2>&1 &&
ncu $M -k regex:topk --log-file $O/launches_scan_big.csv python scripts/prof_case.py scan_big 2 > $O/ncu_scan_big.log 2>&1
python scripts/prof_case.py scan_small 2 > $O/plain_scan_small.log 2>&1 &&
ncu $M -k regex:topk --log-file $O/launches_scan_small.csv python scripts/prof_case.py scan_small 2 > $O/ncu_scan_small.log 2>&1
# full sets: the GEMM of the MLP (c_fc: M=12800 N=3072 K=768) is launch #... pick 3 gemm launches after the first layer
python scripts/prof_case.py encode 1 > $O/plain_encode1.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gemm_tc -s 6 -c 4 -o $O/prof_gemm -f python scripts/prof_case.py encode 1 > $O/ncu_gemm.log 2>&1
python scripts/prof_case.py scan_big 1 > $O/plain_scan_big1.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:topk_scan -c 1 -o $O/prof_scan_big -f python scripts/prof_case.py scan_big 1 > $O/ncu_scan_big_full.log 2>&1
python scripts/prof_case.py scan_small 1 > $O/plain_scan_small1.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:topk_scan -c 1 -o $O/prof_scan_small -f python scripts/prof_case.py scan_small 1 > $O/ncu_scan_small_full.log 2>&1
ls -la $O
