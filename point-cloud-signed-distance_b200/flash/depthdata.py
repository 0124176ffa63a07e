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
    """src/depthdata.jl:19-30: first line `x, y, z` camera origin; then rows `x, y, z, r, g, b`.

    A path is read by the library (flash_read_point_cloud, csrc/flash_urdf.h -- the reader that also feeds
    flash_points_set_file); an open text stream is parsed here."""
    if isinstance(file, (str, bytes)):
        import ctypes as C
        import os
        from . import _lib
        from ._desc import c_double_p
        L = _lib.lib()
        path = os.fsencode(file)
        n = C.c_int64()
        origin = np.zeros(3)
        rc = L.flash_read_point_cloud(path, origin.ctypes.data_as(c_double_p), None, None, 0, C.byref(n))
        if rc == 0:
            xyz = np.empty((n.value, 3)); rgb = np.empty((n.value, 3))
            rc = L.flash_read_point_cloud(path, origin.ctypes.data_as(c_double_p), xyz.ctypes.data_as(c_double_p), rgb.ctypes.data_as(c_double_p), n.value, C.byref(n))
        if rc != 0:
            msg = L.flash_last_error(None)
            raise IOError("flash_read_point_cloud failed (%d): %s" % (rc, msg.decode() if msg else "?"))
        return PointCloud(origin, xyz, rgb)
    origin = np.array([float(v) for v in file.readline().split(",")[:3]])
    data = np.loadtxt(file, delimiter=",", ndmin=2)
    if data.shape[1] < 6:
        raise ValueError("point cloud rows must have 6 comma-separated columns")
    return PointCloud(origin, np.ascontiguousarray(data[:, :3]), np.ascontiguousarray(data[:, 3:6]))
