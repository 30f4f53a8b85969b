#!/bin/bash
# Developer tool: summary of an ncu report of fused.cu kernels.  $1 = report, $2 = kernel name substring, $3 = object (default fused.o)
set -e
rep=$1; k=$2; obj=${3:-fused}
python tools/ncu_summary.py $rep /tmp/prof_sum.csv > /dev/null
python - <<PY
import csv
rows=list(csv.reader(open('/tmp/prof_sum.csv')))
for h,v in zip(rows[0],rows[1]): print(f"{h[:70]:70s} {v}")
PY
ncu -i $rep --page raw --csv | python -c "
import csv,sys
rows=list(csv.reader(sys.stdin)); hdr=rows[0]; r=rows[2]
out=[]
for k in hdr:
    if 'issue_stalled' in k and 'per_issue_active' in k: out.append((float(r[hdr.index(k)]), k.replace('smsp__average_warps_issue_stalled_','').replace('_per_issue_active.ratio','')))
print('stalls per issue:', ', '.join(f'{n} {v:.2f}' for v,n in sorted(out, reverse=True)[:8]))
"
ncu -i $rep --page source --csv 2>/dev/null > /tmp/prof_src.csv
rm -rf /tmp/xelf && mkdir /tmp/xelf && (cd /tmp/xelf && cuobjdump -xelf all /root/repo/bpe_transformer_b200/csrc/$obj.o > /dev/null 2>&1)
python tools/sass_lines.py /tmp/prof_src.csv /tmp/xelf/$obj.sm_100a.cubin "$k" 2000 > /tmp/prof_lines.txt
head -${4:-28} /tmp/prof_lines.txt
