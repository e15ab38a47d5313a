#!/bin/bash
# ncu evidence, checkpoint r01_e (run under gpurun; each ncu run follows the same command run plain).
cd "$(dirname "$0")/.."
O=gpurun_out
M="--metrics gpu__time_duration.sum --clock-control none --csv"
python bench.py > $O/r01_e_bench.json 2> $O/r01_e_bench.err
python scripts/prof_case.py encode 2 > $O/plain_encode.log 2>&1 &&
ncu $M --log-file $O/r01_e_launches_encode.csv python scripts/prof_case.py encode 2 > $O/ncu_encode.log 2>&1
python scripts/prof_case.py scan_small 2 > $O/plain_scan_small.log 2>&1 &&
ncu $M -k regex:topk --log-file $O/r01_e_launches_scan_small.csv python scripts/prof_case.py scan_small 2 > $O/ncu_scan_small.log 2>&1
python scripts/prof_case.py scan_big 2 > $O/plain_scan_big.log 2>&1 &&
ncu $M -k regex:topk --log-file $O/r01_e_launches_scan_big.csv python scripts/prof_case.py scan_big 2 > $O/ncu_scan_big.log 2>&1
python scripts/prof_case.py encode 1 > $O/plain_encode1.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gemm_tc -s 5 -c 4 -o $O/r01_e_prof_gemm -f python scripts/prof_case.py encode 1 > $O/ncu_gemm.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:attention -s 1 -c 1 -o $O/r01_e_prof_attn -f python scripts/prof_case.py encode 1 > $O/ncu_attn.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:topk_scan -s 1 -c 1 -o $O/r01_e_prof_scan_small -f python scripts/prof_case.py scan_small 1 > $O/ncu_scan_small_full.log 2>&1
ls -la $O | tail -20
ncu --set full --clock-control none --import-source on -k regex:topk_scan -s 1 -c 1 -o $O/r01_e_prof_scan_big -f python scripts/prof_case.py scan_big 1 > $O/ncu_scan_big_full.log 2>&1
python bench.py --impl reference --steps 2 --warmup 1 > $O/r01_e_bench_reference.json 2> $O/r01_e_bench_reference.err
python scripts/torch_eager_baseline.py > $O/r01_e_torch_eager.json 2> $O/r01_e_torch_eager.err
python scripts/graph_probe.py > $O/r01_e_small_batch.txt 2>&1
EVR_NO_GRAPH=1 python scripts/graph_probe.py >> $O/r01_e_small_batch.txt 2>&1
python scripts/scenecut_probe.py > $O/plain_scene.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:scene_walk -s 2 -c 1 -o $O/r01_e_prof_scene -f python scripts/scenecut_probe.py > $O/ncu_scene.log 2>&1
