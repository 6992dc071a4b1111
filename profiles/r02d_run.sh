mkdir -p gpurun_out
KERNEL=3 REPS=1 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name regex:k_st_ -c 1500 --csv --log-file gpurun_out/r02d_launches_c2.csv python profiles/quick.py c2 16 "k3|" > gpurun_out/r02d_ncu_c2.log 2>&1
python - <<'PY'
import csv, collections
rows=[r for r in csv.reader(open('gpurun_out/r02d_launches_c2.csv')) if len(r)>5]
hdr=rows[0]; ki=hdr.index('Kernel Name'); vi=hdr.index('Metric Value')
tot=collections.defaultdict(float); cnt=collections.Counter(); seq=[]
for r in rows[1:]:
    try: v=float(r[vi].replace(',',''))
    except: continue
    n=r[ki].split('(')[0].split('<')[0]; tot[n]+=v; cnt[n]+=1; seq.append((n,v))
for n in tot: print(n, cnt[n], 'launches', round(tot[n]/1e6,2),'ms total', round(tot[n]/cnt[n]/1e3,1),'us avg')
# per-wave profile: print every 10th wave
tr=[v for n,v in seq if 'trace' in n]; sh=[v for n,v in seq if 'shade' in n]; lo=[v for n,v in seq if 'logic' in n]
for i in range(0,len(tr),8): print(i, 'logic',round(lo[i]/1e3,1),'trace',round(tr[i]/1e3,1),'shade',round(sh[i]/1e3,1) if i<len(sh) else '-')
PY
