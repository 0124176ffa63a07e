"""Runs ON THE GPU BOX: only bench.py's K3 tracking leg (squishable, 50 NaiveSolver iterations per frame, on-device loop).
usage: [FLASHSDF_LIB=...] python tools/k3_probe.py  -> one JSON line"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "point-cloud-signed-distance_b200"))
import bench  # noqa: E402
import flash  # noqa: E402
from flash import _lib  # noqa: E402

out = bench.tracking_leg(flash, _lib.lib())
print(json.dumps({k: ({kk: vv for kk, vv in v.items() if kk in ("points", "ms_per_frame", "frames_per_sec", "frames_per_sec_with_upload", "cost_end_of_each_repeat")} if isinstance(v, dict) else v) for k, v in out.items() if k.startswith("k3_")}))
