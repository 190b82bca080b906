import csv,sys
rows=list(csv.reader(open(sys.argv[1])))
hdr=rows[1]; data=rows[2:]
ix={h:i for i,h in enumerate(hdr)}
def f(r,k):
    try: return float(r[ix[k]])
    except: return 0.0
tot=sum(f(r,'# Samples') for r in data)
print('total samples',tot,'total inst',sum(f(r,'Instructions Executed') for r in data))
# cumulative by region: print top 40 instructions by samples
top=sorted(data,key=lambda r:-f(r,'# Samples'))[:int(sys.argv[2]) if len(sys.argv)>2 else 40]
for r in top:
    st={k:f(r,k) for k in hdr if k.startswith('stall_') and '(' not in k}
    main=sorted(st.items(),key=lambda kv:-kv[1])[:2]
    print(r[ix['Address']][-5:], '%5.1f%%'%(100*f(r,'# Samples')/tot), r[ix['Source']][:70].ljust(70), int(f(r,'Instructions Executed')), main)
