#!/usr/bin/env python
"""Per-source-line view of an `ncu --set full --import-source on` capture: stall mix, opcode mix and the top CUDA-C
lines by warp-stall samples with their main stall reasons (what drove the round-2 changes to s4.cu and dense.cu).

    python tools/ncu_source_lines.py gpurun_out/prof_x.ncu-rep [top_n] [kernel_regex]

(also accepts the CSV of `ncu -i x.ncu-rep --page source --print-source cuda,sass --csv`)."""
import csv, io, subprocess, sys
path=sys.argv[1]; topn=int(sys.argv[2]) if len(sys.argv)>2 else 40
if path.endswith('.ncu-rep'):
    cmd=['ncu','-i',path,'--page','source','--print-source','cuda,sass','--csv','--launch-count','1']
    if len(sys.argv)>3: cmd+=['--kernel-name','regex:'+sys.argv[3]]
    rows=list(csv.reader(io.StringIO(subprocess.run(cmd,capture_output=True,text=True).stdout)))
else:
    rows=list(csv.reader(open(path)))
hdr=rows[2]
iS=hdr.index('# Samples'); iI=hdr.index('Instructions Executed'); iW=hdr.index('L1 Wavefronts Shared'); iG=hdr.index('L1 Tag Requests Global')
stalls=[i for i,k in enumerate(hdr) if k.startswith('stall_') and 'Not Issued' not in k]
def f(v):
    try: return float(v.replace(',',''))
    except: return 0.0
lines=[];cur=None;curfile=None
for r in rows:
    if len(r)==2 and r[0]=='File Path': curfile=r[1].split('/')[-1]; continue
    if len(r)<10 or r[0]=='Line No': continue
    if r[0]!='':
        cur={'f':curfile,'n':r[0],'src':r[1],'s':f(r[iS]),'i':f(r[iI]),'w':f(r[iW]),'g':f(r[iG]),'st':{hdr[i]:f(r[i]) for i in stalls},'ops':{}}
        lines.append(cur)
    elif cur is not None:
        op=r[3].split()
        if not op: continue
        o=op[1] if op[0].startswith('@') else op[0]
        o='.'.join(o.rstrip(';').split('.')[:2])
        cur['ops'][o]=cur['ops'].get(o,0)+f(r[iI])
T=sum(l['s'] for l in lines); TI=sum(l['i'] for l in lines); TW=sum(l['w'] for l in lines)
tot={}
for l in lines:
    for k,v in l['st'].items(): tot[k]=tot.get(k,0)+v
ts=sum(tot.values())
print('stall mix', {k[6:]:round(100*v/ts,1) for k,v in sorted(tot.items(),key=lambda kv:-kv[1]) if v/ts>0.01})
ops={}
for l in lines:
    for k,v in l['ops'].items(): ops[k]=ops.get(k,0)+v
print('op mix', {k:round(100*v/TI,1) for k,v in sorted(ops.items(),key=lambda kv:-kv[1])[:18]})
print('total shared wavefronts',TW)
for l in sorted(lines,key=lambda l:-l['s'])[:topn]:
    st=l['st']; t=sum(st.values()) or 1
    top=', '.join(f"{k[6:]}:{round(100*v/t)}" for k,v in sorted(st.items(),key=lambda kv:-kv[1])[:3] if v/t>0.08)
    print(f"{l['f'][:10]}:{l['n']:>4} smp {100*l['s']/T:5.1f}% inst {100*l['i']/TI:5.1f}% shwf {100*l['w']/max(TW,1):5.1f}% [{top}] | {l['src'].strip()[:95]}")
