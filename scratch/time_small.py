"""mdd_ctc_loss at T=1000, L=120 for growing batch sizes (how the concurrent-sweep path degrades as CTAs share SMs)."""
import sys
sys.path.insert(0, '.'); sys.path.insert(0, 'scratch')
import time_team  # noqa (runs its default shapes first)
for B in (2, 64, 128, 256, 296, 384, 512, 592):
    time_team.run(1000, B, 42, 120)
