ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/batch_launches.csv python tools/batch_stats.py 128 2048 > gpurun_out/batch_ncu.log 2>&1
python - <<'PY'
import csv,io
lines=open('gpurun_out/batch_launches.csv').read().splitlines()
st=next(i for i,l in enumerate(lines) if l.startswith('"ID"'))
rows=list(csv.DictReader(io.StringIO("\n".join(lines[st:]))))
agg={}
for r in rows[-200:]:
    n=r["Kernel Name"].split("(")[0]; v=float(r["Metric Value"].replace(",","")); u=r["Metric Unit"]
    us=v*{"ns":1e-3,"us":1,"ms":1e3}.get(u,1e-3)
    a=agg.setdefault(n,[0,0.0]); a[0]+=1; a[1]+=us
for n,(c,us) in sorted(agg.items(), key=lambda kv:-kv[1][1]): print("%-40s %4d  avg %.1f us"%(n[:40],c,us/c))
PY
