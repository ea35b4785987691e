#!/bin/bash
# per-config median duration of the attention launches of one eager bench step (ncu, cold cache)
tag=$1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:attention_stream -c 140 --csv --log-file gpurun_out/${tag}_att.csv python bench.py --steps 1 --warmup 3 --lanes 1 --no-graph --no-extras --no-cpu-baseline > gpurun_out/${tag}_att.log 2>&1
python - <<EOF2
import csv,collections,re
rows=list(csv.reader(open("gpurun_out/${tag}_att.csv")))
hi=[i for i,r in enumerate(rows) if r and r[0]=="ID"][0]
hdr=rows[hi]; ki,vi=hdr.index("Kernel Name"),hdr.index("Metric Value")
agg=collections.defaultdict(list)
for r in rows[hi+1:]:
    if len(r)>vi: agg[re.sub(r"\(.*","",r[ki])[:60]].append(float(r[vi].replace(",",""))/1e3)
for k,v in sorted(agg.items()): print(k, len(v), round(sorted(v)[len(v)//2],1))
EOF2
