#!/bin/bash
# Developer probe: C3 scan time and DRAM traffic vs the in-flight footprint the slice plan aims for / tile rotation.
cd "$(dirname "$0")/.."
for cfg in "48 0" "48 1" "24 0" "12 0"; do
  set -- $cfg
  echo "== in-flight target $1 MB, EVR_SCAN_NO_ROTATE=$2"
  EVR_SCAN_INFLIGHT_MB=$1 EVR_SCAN_NO_ROTATE=$2 python scripts/hbm_regime_probe.py 2>&1 | grep strided | tail -1
  EVR_SCAN_INFLIGHT_MB=$1 EVR_SCAN_NO_ROTATE=$2 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,lts__t_sector_hit_rate.pct --clock-control none -k regex:topk_scan -s 1 -c 1 --csv --log-file gpurun_out/inflight_$1_$2.csv python scripts/prof_case.py scan_c3 1 > /dev/null 2>&1
  grep -v "^==" gpurun_out/inflight_$1_$2.csv | cut -d, -f13- | tail -3
done
