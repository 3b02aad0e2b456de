import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.argv = sys.argv[:1]
import importlib.util
spec = importlib.util.spec_from_file_location("st", os.path.join(os.path.dirname(__file__), "stage_times.py"))
src = open(spec.origin).read().split("\nB = int(")[0]
exec(src)
for B in (1, 8, 8, 32):
    run("c2 SGBM", 720, 1280, 128, 5, 0, B)
for B in (8, 32):
    run("c2 HH", 720, 1280, 128, 5, 1, B)
    run("c1 HH", 480, 640, 64, 5, 1, B)
