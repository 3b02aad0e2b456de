import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
src = open(os.path.join(os.path.dirname(__file__), "stage_times.py")).read().split("\nB = int(")[0]
exec(src)
run("c1 SGBM", 480, 640, 64, 5, 0, 1)
run("c2 SGBM", 720, 1280, 128, 5, 0, 1)
run("c3 HH", 1080, 1920, 256, 5, 1, 1)
run("c3 HH", 1080, 1920, 256, 5, 1, 2)
