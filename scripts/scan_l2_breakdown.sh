#!/bin/bash
# Where do the DRAM reads of the full-scan launch come from?  L2 sectors by operation, eviction class and source.
#   bash scripts/scan_l2_breakdown.sh "<label>" Q [VAR=value ...]
cd "$(dirname "$0")/.."
O=gpurun_out
label=$1; q=$2; shift 2
MET=dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__d_sectors_fill_device.sum
for op in read write; do for pol in evict_first evict_last evict_normal evict_normal_demote; do MET=$MET,lts__t_sectors_op_${op}_${pol}.sum,lts__t_sectors_op_${op}_${pol}_lookup_miss.sum; done; done
MET=$MET,lts__t_sectors_op_red.sum,lts__t_sectors_op_red_lookup_miss.sum,lts__t_sectors_op_atom.sum
MET=$MET,lts__t_sectors_srcunit_ltcfabric.sum,lts__t_sectors_srcunit_ltcfabric_op_read.sum,lts__t_sectors_srcunit_ltcfabric_op_read_lookup_miss.sum,lts__t_sectors_srcunit_ltcfabric_op_write.sum
MET=$MET,lts__t_sectors_srcunit_tex_op_read.sum,lts__t_sectors_srcunit_tex_op_read_lookup_miss.sum,lts__t_sectors_srcunit_l1_op_read.sum,lts__t_sectors_srcunit_l1_op_read_lookup_miss.sum,lts__t_sectors_srcunit_l1_op_write.sum,lts__t_sectors_srcunit_l1_op_write_lookup_miss.sum
env "$@" ncu --metrics $MET --clock-control none -k regex:topk_scan --csv --log-file $O/scan_l2_$label.csv python scripts/prof_case.py scan:10000000:$q 1 > $O/scan_l2_$label.log 2>&1
grep -v "^==" $O/scan_l2_$label.csv | python -c "
import csv,sys
rows=list(csv.DictReader(l for l in sys.stdin if l.startswith('\"')))
by={}
for r in rows: by.setdefault(r['ID'],{})[r['Metric Name']]=(float(r['Metric Value'].replace(',','')), r['Metric Unit'])
m=by[sorted(by)[-1]]
print('== $label Q=$q $*')
for k,(v,u) in sorted(m.items()):
    if u=='sector': print('  %-70s %9.2f GB' % (k, v*32/1e9))
    elif u=='byte': print('  %-70s %9.2f GB' % (k, v/1e9))
    else: print('  %-70s %12.1f %s' % (k, v, u))
"
