import csv,sys
rows=list(csv.reader(open(sys.argv[1])))
hdr=rows[1]; data=rows[2:]
ix={h:i for i,h in enumerate(hdr)}
def f(r,k):
    try: return float(r[ix[k]])
    except: return 0.0
tot=sum(f(r,'# Samples') for r in data); toti=sum(f(r,'Instructions Executed') for r in data)
seg=0; acc={}
for r in data:
    src=r[ix['Source']]
    a=acc.setdefault(seg,{'s':0,'i':0,'n':0,'fp64':0,'lds':0,'bc':0,'wf':0})
    a['s']+=f(r,'# Samples'); a['i']+=f(r,'Instructions Executed'); a['n']+=1
    if any(op in src for op in ('DFMA','DADD','DMUL')): a['fp64']+=f(r,'Instructions Executed')
    a['wf']+=f(r,'L1 Wavefronts Shared'); a['bc']+=f(r,'L1 Wavefronts Shared Excessive')
    if 'BAR.SYNC' in src: seg+=1
for k,a in acc.items():
    print('seg',k,'static',a['n'],'samples %5.1f%%'%(100*a['s']/tot),'inst %5.1f%%'%(100*a['i']/toti),'inst/1e6 %.1f'%(a['i']/1e6),'fp64 %.1f'%(a['fp64']/1e6),'smem wf %.1f excess %.1f'%(a['wf']/1e6,a['bc']/1e6))
