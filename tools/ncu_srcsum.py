import csv,sys
rows=list(csv.reader(open(sys.argv[1])))
hdr=rows[1]
idx={h:i for i,h in enumerate(hdr)}
stalls=[h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
tot=0
data=[]
for r in rows[2:]:
    if len(r)<len(hdr): continue
    n=int(r[idx['# Samples']]); tot+=n
    data.append((r[idx['Address']],r[idx['Source']].strip(),n,int(r[idx['Instructions Executed']]),{s[6:]:int(r[idx[s]]) for s in stalls if int(r[idx[s]])}))
print('total samples',tot)
thr=int(sys.argv[2]) if len(sys.argv)>2 else 100000
for a,src,n,ex,st in data:
    if ex>thr:
        print(f"{a[-4:]} {n:6d} {ex:9d} {src[:58]:58s} {st}")
