#!/bin/bash
# Developer tool: per-source-line profile of one kernel of an `ncu --set full --import-source on` report.
#   tools/hot_lines.sh report.ncu-rep kernel_regex object_name(fused|encode|scan) mangled_name_substring lines > profiles/x_hot_lines.txt
# (the first launch that matches the regex; the substring picks the template instance in the object, e.g. fused_count_kernelILb0ELb0E)
root=$(cd $(dirname $0)/.. && pwd); rep=$1; k=$2; obj=$3; mangled=$4; n=${5:-40}
ncu -i $rep --page raw --csv -k regex:$k -c 1 | python -c "
import csv,sys
rows=list(csv.reader(sys.stdin)); hdr=rows[0]; r=rows[2]
g=lambda m: r[hdr.index(m)] if m in hdr else '?'
print('kernel', g('Kernel Name').split('(')[0], '| duration', g('gpu__time_duration.sum'), rows[1][hdr.index('gpu__time_duration.sum')], '| warp instructions', g('smsp__inst_executed.sum'),
      '| active lanes per instruction', g('smsp__thread_inst_executed_per_inst_executed.ratio'), '| issue active %', g('smsp__issue_active.avg.pct_of_peak_sustained_active'),
      '| registers', g('launch__registers_per_thread'), '| grid', g('launch__grid_size'), 'x', g('launch__block_size'))
out=[]
for kk in hdr:
    if 'issue_stalled' in kk and 'per_issue_active' in kk: out.append((float(r[hdr.index(kk)]), kk.replace('smsp__average_warps_issue_stalled_','').replace('_per_issue_active.ratio','')))
print('stalled warps per issued instruction, by reason:', ', '.join(f'{nm} {v:.2f}' for v,nm in sorted(out, reverse=True)[:9]))
"
ncu -i $rep --page source --csv -k regex:$k -c 1 2>/dev/null > /tmp/hl_src.csv
rm -rf /tmp/hl_xelf && mkdir /tmp/hl_xelf && (cd /tmp/hl_xelf && cuobjdump -xelf all $root/bpe_transformer_b200/csrc/$obj.o > /dev/null 2>&1)
echo "share of executed warp instructions | mean active lanes | stall samples | SASS instructions | (file, line)"
python $root/tools/sass_lines.py /tmp/hl_src.csv /tmp/hl_xelf/$obj.sm_100a.cubin "$mangled" 2000 | head -$n
