"""Per-SASS-instruction stall samples of one region of an ncu report: prints instructions in order with stall counts."""
import csv, io, subprocess, sys
rep, lo, hi = sys.argv[1], int(sys.argv[2], 16), int(sys.argv[3], 16)
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
h = None
for r in rows:
    if r and r[0] == "Address": h = r; ci = {n: i for i, n in enumerate(h)}; continue
    if not h or len(r) < len(h) - 2: continue
    a = int(r[0], 16) & 0xffffff
    if lo <= a <= hi:
        st = {k: int(r[ci[k]] or 0) for k in ("stall_wait", "stall_short_sb", "stall_long_sb", "stall_mio", "stall_math", "stall_not_selected", "stall_selected", "stall_dispatch", "stall_branch_resolving", "stall_no_inst")}
        tot = int(r[ci["# Samples"]])
        print("%06x %-64s n=%3d  %s" % (a, r[ci["Source"]].strip()[:64], tot, " ".join("%s=%d" % (k[6:], v) for k, v in st.items() if v)))
