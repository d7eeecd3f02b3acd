#!/bin/bash
# per-kernel times of one step of config 3 under given params:  tools/explore_kern.sh name=val ...
export CMSBWT_PROFILE=1
export CMSBWT_PARAMS="$*"
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file /tmp/l.csv python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline > /tmp/l.log 2>&1
python - <<'PY'
import csv,collections,re
rows=list(csv.reader(open('/tmp/l.csv')))
hi=[i for i,r in enumerate(rows) if 'Kernel Name' in r][0]
h=rows[hi];kn=h.index('Kernel Name');mv=h.index('Metric Value')
seq=[(re.sub(r'^void ','',r[kn]).split('(')[0][:44],float(r[mv].replace(',',''))/1e6) for r in rows[hi+1:] if len(r)>mv]
agg=collections.OrderedDict();tot=0
for n,ms in seq:
    agg.setdefault(n,[0,0]);agg[n][0]+=ms;agg[n][1]+=1;tot+=ms
for n,(ms,c) in sorted(agg.items(),key=lambda kv:-kv[1][0])[:22]: print("%8.3f ms %3d  %s"%(ms,c,n))
print("%8.3f ms total, %d launches"%(tot,len(seq)))
PY
