#!/bin/bash
# DRAM bytes of the full-scan launch (10 M x 512 store, ncu, one launch) under the plan knobs given as VAR=value args.
#   bash scripts/scan_traffic_sweep.sh "<label>" Q [VAR=value ...]
cd "$(dirname "$0")/.."
O=gpurun_out
label=$1; q=$2; shift 2
env "$@" ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_sectors_srcunit_tex_op_read.sum,lts__t_sectors_srcunit_tex_op_read_lookup_miss.sum \
    --clock-control none -k regex:topk_scan --csv --log-file $O/scan_traffic_$label.csv python scripts/prof_case.py scan:10000000:$q 1 > $O/scan_traffic_$label.log 2>&1
grep -v "^==" $O/scan_traffic_$label.csv | python -c "
import csv,sys
rows=list(csv.DictReader(l for l in sys.stdin if l.startswith('\"')))
by={}
for r in rows: by.setdefault(r['ID'],{})[r['Metric Name']]=float(r['Metric Value'].replace(',',''))
m=by[sorted(by)[-1]]
print('$label Q=$q $*: dram_read %.1f GB  write %.2f GB  time %.2f ms  lts_read %.0f GB  miss %.0f GB' % (m['dram__bytes_read.sum']/1e9, m['dram__bytes_write.sum']/1e9, m['gpu__time_duration.sum']/1e6, m['lts__t_sectors_srcunit_tex_op_read.sum']*32/1e9, m['lts__t_sectors_srcunit_tex_op_read_lookup_miss.sum']*32/1e9))
"
