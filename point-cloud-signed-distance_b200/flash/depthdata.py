"""Mirror of src/depthdata.jl: the text point-cloud reader (render_lcmgl is out of scope)."""
import numpy as np


class PointCloud:  # src/depthdata.jl:13-16
    def __init__(self, camera_origin, positions, colors):
        self.camera_origin = camera_origin
        self.positions = positions  # (n, 3) float64 -- [p.position for p in point_cloud.points]
        self.colors = colors        # (n, 3) float64 RGB

    def __repr__(self):  # src/depthdata.jl:17
        return "PointCloud with origin: %s containing %d points" % (list(self.camera_origin), len(self.positions))


def read_point_cloud(file):
    """src/depthdata.jl:19-30: first line `x, y, z` camera origin; then rows `x, y, z, r, g, b`."""
    close = False
    if isinstance(file, str):
        file, close = open(file), True
    try:
        origin = np.array([float(v) for v in file.readline().split(",")[:3]])
        data = np.loadtxt(file, delimiter=",", ndmin=2)
    finally:
        if close:
            file.close()
    if data.shape[1] < 6:
        raise ValueError("point cloud rows must have 6 comma-separated columns")
    return PointCloud(origin, np.ascontiguousarray(data[:, :3]), np.ascontiguousarray(data[:, 3:6]))
