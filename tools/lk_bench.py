"""GPU box: time the Lucas-Kanade alignment alone (bench.py's lk_align leg) and print it as JSON"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import bench
print(json.dumps(bench.lk_leg(0, cpu="--no-cpu" not in sys.argv), indent=1))
