"""K2 / K4 legs only (bench.configs_leg).  Prints one JSON line."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "point-cloud-signed-distance_b200")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench, flash
from flash import _lib
print(json.dumps(bench.configs_leg(flash, _lib.lib(), 3), indent=1))
