import json, sys
tag = sys.argv[1] if len(sys.argv) > 1 else ""
for line in sys.stdin:
    line = line.strip()
    if not line.startswith("{"):
        continue
    d = json.loads(line)
    r = d.get("roofline", {})
    print(tag, "value=%.4g ms=%.3f frac=%.4f e2e=%.4g launches=%s clocks=%s" % (d["value"], d["ms_per_step"], r.get("frac", 0), d["e2e"]["value"], d.get("gpu_launches"), d.get("clocks")))
