"""Mirror of src/depthsensors.jl (ray fan + batched ray casting; LCMGL drawing is out of scope)."""
import ctypes as C

import numpy as np

from ._desc import c_double_p
from . import _lib
from .core import Skin, skin as _skin


def generateKinectRays(rows, cols, vertical_fov=0.4682, horizontal_fov=0.5449):  # src/depthsensors.jl:10-30
    rays = np.empty((rows, cols, 3))
    rc = _lib.lib().flash_kinect_rays(rows, cols, vertical_fov, horizontal_fov, rays.ctypes.data_as(c_double_p))
    if rc != 0:
        raise ValueError("flash_kinect_rays failed (%d)" % rc)
    return rays


class DepthSensor:  # src/depthsensors.jl:32-34
    def __init__(self, rays):
        self.rays = np.ascontiguousarray(rays, dtype=np.float64)


def Kinect(rows, cols, vertical_fov=0.4682, horizontal_fov=0.5449):  # src/depthsensors.jl:54
    return DepthSensor(generateKinectRays(rows, cols, vertical_fov, horizontal_fov))


def _cam12(sensor_tform):
    T = np.asarray(sensor_tform, float)
    if T.shape == (4, 4):
        T = T[:3]
    if T.shape != (3, 4):
        raise ValueError("sensor_tform must be a 3x4 [R|t] (AffineMap) matrix")
    return np.ascontiguousarray(T).reshape(12)


def _cast(surface, sensor, sensor_tform, want_hits):
    if not isinstance(surface, Skin):
        raise TypeError("surface must be the closure returned by flash.skin(state)")
    m = surface.manipulator
    rays = sensor.rays.reshape(-1, 3)
    R = len(rays)
    cam = _cam12(sensor_tform)
    depths = np.empty(R)
    hits = np.empty((R, 3)) if want_hits else None
    nh = C.c_int64(0)
    m.check(_lib.lib().flash_raycast(m.ctx(), surface.x.ctypes.data_as(c_double_p), rays.ctypes.data_as(c_double_p), R,
                                     cam.ctypes.data_as(c_double_p), depths.ctypes.data_as(c_double_p),
                                     hits.ctypes.data_as(c_double_p) if want_hits else None, C.byref(nh)))
    return depths.reshape(sensor.rays.shape[:2]), (hits[:nh.value] if want_hits else None)


def raycast_depths(surface, sensor, sensor_tform):  # src/depthsensors.jl:88-97
    return _cast(surface, sensor, sensor_tform, False)[0]


def raycast_points(surface, sensor, sensor_tform):  # src/depthsensors.jl:99-113
    return _cast(surface, sensor, sensor_tform, True)[1]


def raycast(state, sensor, sensor_tform):  # src/depthsensors.jl:115-118
    return raycast_points(_skin(state), sensor, sensor_tform)


def affine(translation=(0, 0, 0), axis=None, angle=0.0):
    """compose(Translation(t), LinearMap(AngleAxis(angle, axis...))) as a 3x4 matrix."""
    T = np.hstack([np.eye(3), np.zeros((3, 1))])
    if axis is not None:
        a = np.asarray(axis, float)
        a = a / np.linalg.norm(a)
        K = np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
        T[:, :3] = np.eye(3) + np.sin(angle) * K + (1 - np.cos(angle)) * (K @ K)
    T[:, 3] = translation
    return T
