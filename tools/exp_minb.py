"""Dp = 128 sweeps at several batch sizes (AGG_MINB2_GRID switch): stage times of c2-sized jobs."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from stage_times import run
for B in (4, 8, 12, 16, 20, 32):
    run("c2 SGBM", 720, 1280, 128, 5, 0, B, n=2)
    run("c2 HH  ", 720, 1280, 128, 5, 1, B, n=2)
