import sys
sys.path.insert(0, '.'); sys.path.insert(0, 'scratch')
import os
os.environ.setdefault("QUIET", "1")
import importlib, io, contextlib
buf = io.StringIO()
with contextlib.redirect_stdout(buf):
    import time_team
T, B, C, L = (int(x) for x in sys.argv[1:5])
time_team.run(T, B, C, L, steps=40)
