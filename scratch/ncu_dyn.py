"""Dynamic instruction mix from an ncu report's SASS source page: executed warp instructions, stall samples and shared
wavefronts aggregated by opcode and by contiguous region (split where the executed count changes by > 5 %)."""
import csv, io, re, subprocess, sys, collections
rep = sys.argv[1]
kern = sys.argv[2] if len(sys.argv) > 2 else None
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
# split per kernel
secs = []
cur = None
for r in rows:
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "hdr": None, "rows": []}
        secs.append(cur)
    elif cur is not None and r and r[0] == "Address":
        cur["hdr"] = r
    elif cur is not None and cur["hdr"] and len(r) >= len(cur["hdr"]) - 2:
        cur["rows"].append(r)
for s in secs:
    if kern and kern not in s["name"]: continue
    h = s["hdr"]
    ci = {n: i for i, n in enumerate(h)}
    def op(t):
        t = re.sub(r'^\s*@!?U?P\d+\s+', '', t.strip())
        return t.split()[0].split('.')[0] if t.split() else "?"
    tot = 0; byop = collections.Counter(); stall = collections.Counter(); wf = collections.Counter()
    regions = []; reg = None
    for r in s["rows"]:
        n = int(r[ci["Instructions Executed"]]); st = int(r[ci["# Samples"]]); w = int(r[ci["L1 Wavefronts Shared"]] or 0)
        o = op(r[ci["Source"]])
        tot += n; byop[o] += n; stall[o] += st; wf[o] += w
        if reg is None or abs(n - reg["n"]) > 0.05 * max(n, reg["n"], 1):
            reg = {"n": n, "cnt": 0, "ops": collections.Counter(), "st": 0, "wf": 0, "start": r[ci["Address"]]}
            regions.append(reg)
        reg["cnt"] += 1; reg["ops"][o] += 1; reg["st"] += st; reg["wf"] += w
    print("kernel", s["name"][:100]); print("total executed", tot)
    ts = sum(stall.values())
    for o, n in byop.most_common(30):
        print("  %-14s %6.2f%% inst  %6.2f%% stall  wf %d" % (o, 100.0 * n / tot, 100.0 * stall[o] / max(ts, 1), wf[o]))
    print("regions (static count x executions), top by dynamic instructions:")
    regions.sort(key=lambda g: -g["n"] * g["cnt"])
    for g in regions[:25]:
        print("  %s static %4d x exec %9d = %5.2f%% inst, %5.2f%% stall, wf %10d  %s" % (g["start"][-6:], g["cnt"], g["n"], 100.0 * g["n"] * g["cnt"] / tot,
              100.0 * g["st"] / max(ts, 1), g["wf"], ", ".join("%s:%d" % kv for kv in g["ops"].most_common(8))))
