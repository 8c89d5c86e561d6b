#!/bin/bash
# usage: tools/ncu_summary.sh <file.ncu-rep>   -> a text summary on stdout (details page + the raw metrics quoted in DESIGN.md)
REP=$1
echo "# ncu --set full --clock-control none --import-source on; summary made by tools/ncu_summary.sh from $(basename $REP)"
ncu -i $REP --page details --print-units base 2>/dev/null | grep -E "^  [A-Za-z_0-9:<>]+.*\(|Section:|Duration|Throughput|Pipe|Issue|Registers|Stall|Hit Rate|Occupancy|Executed Ipc|L2 Cache|DRAM|Shared Memory|Bank|Block Limit|Grid Size|Block Size|No Eligible|Eligible Warps|Spill"
echo "# raw metrics"
ncu -i $REP --page raw --csv 2>/dev/null | python3 -c "
import csv, sys, re
rows = list(csv.reader(sys.stdin))
hdr, units = rows[0], rows[1]
want = re.compile(r'^(dram__bytes_(read|write)\.sum|lts__t_bytes\.sum|lts__t_sector_hit_rate\.pct|l1tex__data_bank_conflicts_pipe_lsu\.sum|l1tex__t_sector_hit_rate\.pct|sm__inst_executed_pipe_(alu|fma|fp64|xu|lsu|uniform)\.avg\.pct_of_peak_sustained_active|sm__pipe_(alu|fma|fmaheavy|fp64|tensor|shared)_cycles_active\.avg\.pct_of_peak_sustained_active|sm__pipe_tensor_subpipe_.*pct_of_peak_sustained_active|smsp__average_warps_issue_stalled_.*_per_issue_active\.ratio|smsp__issue_active\.avg\.pct_of_peak_sustained_active|smsp__inst_executed\.sum|sm__warps_active\.avg\.pct_of_peak_sustained_active|launch__registers_per_thread|smsp__thread_inst_executed_per_inst_executed\.ratio|gpu__time_duration\.sum)$')
for r in rows[2:]:
    print('## launch:', r[hdr.index('Kernel Name')] if 'Kernel Name' in hdr else '')
    for h, u, v in zip(hdr, units, r):
        if want.match(h): print(' %s %s %s' % (h, u, v))
"
