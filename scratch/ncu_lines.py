import csv, sys
rows=list(csv.reader(open(sys.argv[1])))
hi=[i for i,r in enumerate(rows[:6]) if 'Source' in r][0]
hdr=rows[hi]
ci=hdr.index('Instructions Executed'); st=hdr.index('Warp Stall Sampling (All Samples)')
data=[]
fname=None
for r in rows[hi+1:]:
    if len(r)<ci+1: continue
    if r[0]=='' : continue   # sass rows
    k=len(hdr)
    try: n=int(r[ci-k]); sv=int(r[st-k] or 0)
    except: continue
    data.append((n, sv, r[0], ','.join(r[1:len(r)-k+2]).strip()[:100]))
tot=sum(d[0] for d in data); stt=sum(d[1] for d in data)
print('total inst', tot, 'stall samples', stt)
for n,s,l,src in sorted(data, reverse=True)[:int(sys.argv[2]) if len(sys.argv)>2 else 40]:
    print('%5.1f%% st%5.1f%%  L%s  %s'%(100*n/tot, 100*s/max(stt,1), l, src))
