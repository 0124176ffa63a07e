"""Single-process multi-device context legs only (bench.single_process_leg): python tools/multi_bench.py N [points [configs]]"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "point-cloud-signed-distance_b200")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench, flash
from flash import _lib
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2
pts = int(sys.argv[2]) if len(sys.argv) > 2 else 10_000_000
cfg = int(sys.argv[3]) if len(sys.argv) > 3 else 64
print(json.dumps(bench.single_process_leg(flash, _lib.lib(), list(range(n)), 3, pts, cfg)))
