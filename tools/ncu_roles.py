"""Warp-stall sampling of the chain kernel by warp role, from an `ncu --set full --import-source on` report (read here, no GPU
needed): the kernel's three roles are separated by their setmaxnreg instructions in the SASS; within a role the
instructions are grouped by how often a warp executed them (stage loop body, per-stage bookkeeping, epilogues).
usage: python tools/ncu_roles.py gpurun_out/prof.ncu-rep [panel-detail]   (profiles/r02f_stall_summary_by_role.txt)"""
import csv, collections, sys, subprocess
rep=sys.argv[1]
out=subprocess.run(['ncu','-i',rep,'--page','source','--csv'],capture_output=True,text=True).stdout
rows=list(csv.reader(out.splitlines()))
hdr=rows[1]; data=rows[2:]
ix={h:i for i,h in enumerate(hdr)}
stalls=[h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
bounds=[i for i,r in enumerate(data) if 'USETMAXREG' in r[ix['Source']]]+[len(data)]
print(rows[0][1][:90], bounds)
def region(a,b,name,detail=False):
    rs=data[a:b]
    tot=sum(int(r[ix['# Samples']]) for r in rs)
    agg={s:sum(int(r[ix[s]]) for r in rs) for s in stalls}
    print(f"== {name}: samples {tot}")
    print("   ", [(s[6:],round(100*v/tot,1)) for s,v in sorted(agg.items(), key=lambda kv:-kv[1])[:7]])
    if detail:
        by=collections.defaultdict(lambda:[0,0,collections.Counter()])
        for r in rs:
            k=int(r[ix['Instructions Executed']])
            by[k][0]+=int(r[ix['# Samples']]); by[k][1]+=1
            for s in stalls: by[k][2][s]+=int(r[ix[s]])
        for k,(smp,cnt,st) in sorted(by.items(), key=lambda kv:-kv[1][0])[:8]:
            print(f"   exec {k:>10d} instrs {cnt:4d} samples {100*smp/tot:5.1f}%  {[(s[6:],round(100*v/max(smp,1))) for s,v in st.most_common(3)]}")
region(bounds[0],bounds[1],'DMMA',True)
region(bounds[1],bounds[2],'producer')
region(bounds[2],bounds[3],'panel',len(sys.argv)>2)
