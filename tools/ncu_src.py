import csv, sys
rows=list(csv.reader(open(sys.argv[1])))
# a report with several kernels has one block per kernel, each opened by a "Kernel Name" row: take the first
starts=[i for i,r in enumerate(rows) if r and r[0]=='Kernel Name']
end=starts[1] if len(starts)>1 else len(rows)
print('kernel:',rows[0][1] if len(rows[0])>1 else '?')
hdr=rows[1]; data=[r for r in rows[2:end] if len(r)==len(hdr)]
ix={h:i for i,h in enumerate(hdr)}
reasons=[h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
def I(x):
    try: return int(x)
    except: return 0
tot=sum(I(d[ix['# Samples']]) for d in data)
agg={r:sum(I(d[ix[r]]) for d in data) for r in reasons}
print('total samples',tot)
for r,v in sorted(agg.items(),key=lambda x:-x[1])[:10]: print(' ',r,v,f'{100*v/tot:.1f}%')
print('warp instr executed',sum(I(d[ix['Instructions Executed']]) for d in data))
n=int(sys.argv[2]) if len(sys.argv)>2 else 30
top=sorted(data,key=lambda d:-I(d[ix['# Samples']]))[:n]
for d in top:
    rs=sorted(((I(d[ix[r]]),r[6:]) for r in reasons),reverse=True)[:2]
    print(d[ix['Address']][-5:],d[ix['# Samples']].rjust(6),d[ix['Instructions Executed']].rjust(9),d[ix['Source']][:64].ljust(64),rs)
