"""Top SASS instructions by warp-stall samples (and all barrier / MMA / TMA / TMEM ops) of a kernel, from
`ncu -i rep.ncu-rep --page source --csv --kernel-name regex:<kernel> > src.csv`.
Usage: python profiles/sass_top.py src.csv [min share]"""
import csv, sys
csv.field_size_limit(10**9)
rows=list(csv.reader(open(sys.argv[1])))
hi=next(i for i,r in enumerate(rows) if 'Address' in r)
h=rows[hi]
ai=h.index('Address'); si=h.index('Source'); ii=h.index('Instructions Executed'); wi=h.index('Warp Stall Sampling (All Samples)')
data=[]
for r in rows[hi+1:]:
    if len(r)<=max(wi,ii):
        if data: break
        continue
    try: data.append((int(r[wi] or 0), int(r[ii] or 0), r[ai], r[si]))
    except ValueError: continue
tot=sum(d[0] for d in data); ti=sum(d[1] for d in data)
print(len(data), 'samples', tot, 'warp insts', ti)
thr=float(sys.argv[2]) if len(sys.argv)>2 else 0.01
for idx,d in enumerate(data):
    if d[0] > tot*thr or any(k in d[3] for k in ('SYNCS.PHASECHK','UTCHMMA','LDTM','UTMALDG')):
        print(idx, f"{100*d[0]/tot:5.1f}%", d[1], d[3][:100])
