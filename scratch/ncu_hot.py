"""Top stalling SASS instructions of the hot chunk bodies (ncu source page, cuda,sass view) with their index in
address order, source line and dominant stall reason."""
import csv, sys
rows=list(csv.reader(open(sys.argv[1])))
thr=int(sys.argv[2]) if len(sys.argv)>2 else 300
hdr=None; sass={}; cur=None
for r in rows:
    if r and r[0]=="Line No": hdr=r; ci=hdr.index("Instructions Executed"); st=hdr.index("Warp Stall Sampling (All Samples)"); idx={h:i for i,h in enumerate(hdr)}; continue
    if hdr and len(r)==len(hdr):
        if r[2]=="-": cur=r[0]
        else:
            try:
                reasons={k[6:]:int(r[idx[k]] or 0) for k in idx if k.startswith('stall_') and '(' not in k}
                top=max(reasons,key=reasons.get) if reasons else ''
                sass[int(r[2],16)]=(r[3].strip(), int(r[st] or 0), int(r[ci] or 0), cur, top)
            except: pass
tot=sum(v[1] for v in sass.values())
ks=[k for k in sorted(sass) if sass[k][2]>=250000]
print('total samples',tot,'hot instrs',len(ks))
for i,k in enumerate(ks):
    s=sass[k]
    if s[1]>=thr: print('%5d L%-4s %5.2f%% %-12s %s'%(i,s[3],100*s[1]/tot,s[4][:12],s[0][:80]))
