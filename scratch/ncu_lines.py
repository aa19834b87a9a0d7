"""Aggregate an `ncu --page source --csv --print-source cuda,sass` dump by CUDA source line:
instructions executed, stall samples and shared-memory wavefronts per line."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hdr = None
data = []
fname = ""
for r in rows:
    if r and r[0] == "File Path":
        fname = r[1].split("/")[-1]
    elif r and r[0] == "Line No":
        hdr = r
        ci = hdr.index("Instructions Executed"); st = hdr.index("Warp Stall Sampling (All Samples)")
        wf = hdr.index("L1 Wavefronts Shared"); wi = hdr.index("L1 Wavefronts Shared Ideal")
    elif hdr and len(r) == len(hdr) and r[2] == "-":
        try:
            data.append((int(r[ci]), int(r[st] or 0), int(r[wf] or 0), int(r[wi] or 0), fname, r[0], r[1].strip()[:90]))
        except ValueError:
            pass
tot = sum(d[0] for d in data); stt = sum(d[1] for d in data); wft = sum(d[2] for d in data)
print("total inst %d  stall samples %d  smem wavefronts %d (ideal %d)" % (tot, stt, wft, sum(d[3] for d in data)))
for n, s, w, wi, f, l, src in sorted(data, reverse=True)[:top]:
    print("%5.1f%% st%5.1f%% wf%5.1f%% (x%.2f) %s:%s  %s" % (100 * n / tot, 100 * s / max(stt, 1), 100 * w / max(wft, 1), w / max(wi, 1), f[:12], l, src))
