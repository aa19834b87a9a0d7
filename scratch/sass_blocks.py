"""Histogram the opcodes of a SASS dump per straight-line region between branch targets (loop bodies show up as the
big regions that end in a backward BRA)."""
import re, sys, collections
lines = open(sys.argv[1]).read().splitlines()
ins = []
for l in lines:
    m = re.match(r'\s+/\*([0-9a-f]{4,})\*/\s+(.*?);', l)
    if m:
        ins.append((int(m.group(1), 16), m.group(2)))
targets = set()
for a, t in ins:
    m = re.search(r'\bBRA\b.*?(0x[0-9a-f]+)', t)
    if m:
        targets.add(int(m.group(1), 16))
def op(t):
    t = re.sub(r'^@!?U?P\d+\s+', '', t)
    return t.split()[0].split('.')[0]
blocks = []
cur = []
for a, t in ins:
    if a in targets and cur:
        blocks.append(cur); cur = []
    cur.append((a, t))
    if re.search(r'\b(BRA|EXIT|RET)\b', t):
        blocks.append(cur); cur = []
if cur: blocks.append(cur)
print("instructions", len(ins), "blocks", len(blocks))
for b in blocks:
    if len(b) < int(sys.argv[2]) if len(sys.argv) > 2 else 150: continue
    h = collections.Counter(op(t) for a, t in b)
    print("block 0x%x..0x%x n=%d last=%s" % (b[0][0], b[-1][0], len(b), b[-1][1][:40]))
    print("   ", ", ".join("%s:%d" % kv for kv in h.most_common()))
