import sys
sys.path.insert(0, '.'); sys.path.insert(0, 'scratch')
import io, contextlib
with contextlib.redirect_stdout(io.StringIO()):
    import time_team
for B in (512, 592, 640, 768, 896, 1024, 1184, 1280):
    time_team.run(1000, B, 42, 120)
