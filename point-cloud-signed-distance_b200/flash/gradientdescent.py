"""Mirror of src/gradientdescent.jl."""
import numpy as np

from ._desc import c_double_p, c_int32_p
from . import _lib
from .core import ManipulatorState, _flatten, num_states

default_deformation_cost_weight = 10  # src/gradientdescent.jl:7


def flatten(state):  # src/gradientdescent.jl:9-11
    return _flatten(state)


def unflatten_(state, data):  # unflatten!, src/gradientdescent.jl:13-17
    data = np.asarray(data, float)
    nq = state.mechanism_state.q.size
    if data.size != nq + state.deformation_data.size:
        raise ValueError("state vector has %d entries, expected %d" % (data.size, nq + state.deformation_data.size))
    state.mechanism_state.q[:] = data[:nq]
    state.deformation_data[:] = data[nq:]


def cost(state, sensed_points, deformation_cost_weight=default_deformation_cost_weight):  # src/gradientdescent.jl:28-39
    f = CostFunctor(state.manipulator, sensed_points)
    return f(flatten(state), deformation_cost_weight=deformation_cost_weight)


class CostFunctor:  # src/gradientdescent.jl:41-57
    """cost = CostFunctor(manipulator, sensed_points); cost(x) -> raw cost (not divided by N).

    The functor captures its points, as in the reference (src/gradientdescent.jl:41-47).  The manipulator's
    flash_ctx holds ONE frame at a time, so the manipulator remembers which functor uploaded last
    (`_points_owner`) and a functor whose cloud was displaced re-uploads it before its next call: two functors
    on one manipulator never evaluate on each other's points.  Every call is one flash_cost_grad launch.
    Batched forms (`batch`) evaluate C candidate states at once.
    """

    def __init__(self, manipulator, sensed_points):
        self.manipulator = manipulator
        pts = np.ascontiguousarray(sensed_points, dtype=np.float64).reshape(-1, 3)
        self.sensed_points = pts
        self.upload()

    def upload(self):
        m = self.manipulator
        m.check(_lib.lib().flash_points_set(m.ctx(), self.sensed_points.ctypes.data_as(c_double_p), len(self.sensed_points)))
        m._points_owner = self

    def batch(self, X, gradient=True, argmin=False, deformation_cost_weight=default_deformation_cost_weight):
        m = self.manipulator
        X = np.ascontiguousarray(X, dtype=np.float64)
        if X.ndim == 1:
            X = X.reshape(1, -1)
        nx = num_states(m)
        if X.shape[1] != nx:
            raise ValueError("state vector has %d entries, expected %d" % (X.shape[1], nx))
        Cn = X.shape[0]
        if m._points_owner is not self or m._ctx is None:
            self.upload()          # another functor (or a rebuilt ctx) replaced this functor's cloud
        if _lib.lib().flash_num_points(m.ctx()) != len(self.sensed_points):
            raise RuntimeError("flash ctx holds %d points, this functor %d" % (_lib.lib().flash_num_points(m.ctx()), len(self.sensed_points)))
        cost = np.empty(Cn)
        grad = np.empty((Cn, nx)) if gradient else None
        arg = np.empty((Cn, len(self.sensed_points)), dtype=np.int32) if argmin else None
        m.check(_lib.lib().flash_cost_grad(
            m.ctx(), X.ctypes.data_as(c_double_p), Cn, float(deformation_cost_weight),
            cost.ctypes.data_as(c_double_p),
            grad.ctypes.data_as(c_double_p) if gradient else None,
            arg.ctypes.data_as(c_int32_p) if argmin else None))
        return cost, grad, arg

    def __call__(self, x, deformation_cost_weight=default_deformation_cost_weight):
        c, _, _ = self.batch(x, gradient=False, deformation_cost_weight=deformation_cost_weight)
        return float(c[0])


def CostAndGradientFunctor(cost):  # src/gradientdescent.jl:59-65: (g, x) -> value, gradient written into g
    def f(g, x):
        c, grad, _ = cost.batch(x, gradient=True)
        g[:] = grad[0]
        return float(c[0])
    return f


class FramesCostFunctor:
    """A batch of F frames evaluated with C candidate states per frame in one launch (BASELINE.json configs[3]:
    8 Kinect frames x 16 candidates): flash_frames_set + flash_cost_grad_frames.  X is [F][C][nx]."""

    def __init__(self, manipulator, frames):
        import ctypes as C
        self.manipulator = manipulator
        self.frames = [np.ascontiguousarray(f, dtype=np.float64).reshape(-1, 3) for f in frames]
        self._ptrs = (c_double_p * len(self.frames))(*[f.ctypes.data_as(c_double_p) for f in self.frames])
        self._counts = (C.c_int64 * len(self.frames))(*[len(f) for f in self.frames])
        self.upload()

    def upload(self):
        m = self.manipulator
        m.check(_lib.lib().flash_frames_set(m.ctx(), self._ptrs, self._counts, len(self.frames)))
        m._points_owner = self

    def batch(self, X, gradient=True, deformation_cost_weight=default_deformation_cost_weight):
        m = self.manipulator
        nx = num_states(m)
        X = np.ascontiguousarray(X, dtype=np.float64)
        F = len(self.frames)
        if X.ndim != 3 or X.shape[0] != F or X.shape[2] != nx:
            raise ValueError("X must be [%d frames][C][%d states]" % (F, nx))
        if m._points_owner is not self or m._ctx is None:
            self.upload()
        Cn = X.shape[1]
        cost = np.empty((F, Cn))
        grad = np.empty((F, Cn, nx)) if gradient else None
        m.check(_lib.lib().flash_cost_grad_frames(m.ctx(), X.ctypes.data_as(c_double_p), Cn, float(deformation_cost_weight),
                                                  cost.ctypes.data_as(c_double_p), grad.ctypes.data_as(c_double_p) if gradient else None))
        return cost, grad
