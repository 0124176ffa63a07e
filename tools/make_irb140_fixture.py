"""Regenerates point-cloud-signed-distance_b200/flash/data/irb140_convhull.json from the reference's
URDF + binary STL hulls with flash.models.load_urdf (run in the build container, where
/root/reference exists; the GPU box only sees the committed JSON)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "point-cloud-signed-distance_b200"))
import flash  # noqa: E402

REF = "/root/reference/examples/data"
m = flash.models.load_urdf(os.path.join(REF, "IRB140/urdf/irb_140_convhull.urdf"), package_path=[REF])
print(m)
out = os.path.join(ROOT, "point-cloud-signed-distance_b200/flash/data/irb140_convhull.json")
with open(out, "w") as f:
    json.dump(flash.models.manipulator_to_dict(m), f)
print("wrote", out, os.path.getsize(out), "bytes;", [len(s.vertices) for s in m.surfaces], "vertices")
