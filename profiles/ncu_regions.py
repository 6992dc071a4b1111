import csv,sys,re
rows=list(csv.reader(open(sys.argv[1])))
# region boundaries from function names in er_wavefront.cuh
src=open('/root/repo/elevenrender_b200/csrc/er_wavefront.cuh').read().split('\n')
marks=[]
for i,l in enumerate(src,1):
    for name,tag in (('ER_D void wf_push','queue'),('ER_D void wf_pop','descend'),('wf_exact_dist(','leaf/cand'),('ER_D int wf_shade','handlers'),('ER_D void wf_trav_idle','mainloop'),('ER_D unsigned wf_lane','queue')):
        if name in l and not l.strip().startswith('//'): marks.append((i,tag))
marks.sort()
def wf_region(l):
    r='queue'
    for i,t in marks:
        if l>=i: r=t
    return r
hdr=None; cur=None; groups={}
def grp(f,l):
    if f=='er_wavefront.cuh': return wf_region(int(l))
    if f in('er_shade.cuh','er_device.cuh','er_path.cuh'): return 'shade'
    if f=='er_trace.cuh': return 'mt/hitpos'
    return f
for r in rows:
    if len(r)>=2 and r[0]=="File Path": cur=r[1].split('/')[-1]; continue
    if r and r[0]=="Line No": hdr=r; continue
    if hdr and len(r)==len(hdr) and r[0] not in ("","-"):
        d=dict(zip(hdr,r)); g=grp(cur,r[0]); G=groups.setdefault(g,{})
        for k in ('# Samples','stall_no_inst','stall_long_sb','stall_wait','stall_short_sb','stall_branch_resolving','stall_selected','Instructions Executed','Thread Instructions Executed'):
            try: G[k]=G.get(k,0)+int(d[k])
            except: pass
tot=sum(G['# Samples'] for G in groups.values()); ti=sum(G['Instructions Executed'] for G in groups.values())
print(f"{'group':12s} samp%  noinst long_sb wait short br  sel | inst% thr")
for g,G in sorted(groups.items(), key=lambda kv:-kv[1]['# Samples'])[:10]:
    s=G['# Samples']
    print(f"{g:12s} {100*s/tot:5.1f}  {100*G['stall_no_inst']/s:5.1f} {100*G['stall_long_sb']/s:5.1f} {100*G['stall_wait']/s:5.1f} {100*G['stall_short_sb']/s:5.1f} {100*G['stall_branch_resolving']/s:4.1f} {100*G['stall_selected']/s:4.1f} | {100*G['Instructions Executed']/ti:5.1f} {G['Thread Instructions Executed']/max(1,G['Instructions Executed']):5.1f}")
