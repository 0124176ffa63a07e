"""Mirror of src/tracking.jl (+ the NaiveSolver of SimpleGradientDescent.jl it defaults to)."""
import ctypes as C

import numpy as np

from ._desc import SolverParams, c_double_p, c_int32_p
from . import _lib
from .core import num_states
from .gradientdescent import CostFunctor, default_deformation_cost_weight


class NaiveSolver:
    """SimpleGradientDescent.NaiveSolver(n; rate, max_step, iteration_limit, ...).

    That package (@0fcc1f9, REQUIRE.dev:24) is not in the reference tree; the step rule is restated
    from the keyword names used at src/tracking.jl:12-15, examples/irb140.ipynb cell 9 and
    examples/squishable.ipynb cell 9:  g = grad(cost/N); stop if |g| < tolerance;
    step = -rate * g ./ precondition_divisors, rescaled so |step| <= max_step; x += step.
    """

    def __init__(self, num_vars, rate=0.1, max_step=0.5, iteration_limit=30,
                 gradient_convergence_tolerance=0.0, precondition_divisors=None):
        self.num_vars = int(num_vars)
        self.rate = float(rate)
        self.max_step = float(max_step)
        self.iteration_limit = int(iteration_limit)
        self.gradient_convergence_tolerance = float(gradient_convergence_tolerance)
        self.precondition_divisors = None if precondition_divisors is None else np.asarray(precondition_divisors, float)


def estimate_state(manipulator, sensed_points, x_estimated, callback=None, solver=None, on_device=False):
    """src/tracking.jl:8-27.  Returns the solver's solution vector.

    Default: the loop runs here (one flash_cost_grad per iteration) and `callback(x, c)` sees every
    evaluation with the RAW cost, as in the reference (src/tracking.jl:17-21).  on_device=True runs the
    whole loop inside flash_estimate_state (no host round trip per iteration; no callback).
    """
    nx = num_states(manipulator)
    if solver is None:
        solver = NaiveSolver(nx, rate=0.1, max_step=0.5, iteration_limit=30)  # src/tracking.jl:12-15
    x = np.array(x_estimated, dtype=np.float64)
    if x.shape != (nx,):
        raise ValueError("x_estimated has %s entries, expected %d" % (x.shape, nx))
    cost = CostFunctor(manipulator, sensed_points)
    n = len(cost.sensed_points)
    div = solver.precondition_divisors
    if on_device:
        if callback is not None:
            raise ValueError("callback is not available with on_device=True")
        X = x.reshape(1, nx).copy()
        sp = SolverParams(solver.rate, solver.max_step, solver.iteration_limit,
                          solver.gradient_convergence_tolerance, float(default_deformation_cost_weight))
        manipulator.check(_lib.lib().flash_estimate_state(
            manipulator.ctx(), X.ctypes.data_as(c_double_p), 1, C.byref(sp),
            None if div is None else np.ascontiguousarray(div).ctypes.data_as(c_double_p), None, None))
        return X[0]
    for _ in range(solver.iteration_limit):
        c, g, _a = cost.batch(x, gradient=True)
        if callback is not None:
            callback(x.copy(), float(c[0]))
        g = g[0] / n  # wrapped_cost divides by length(sensed_points), src/tracking.jl:20
        if solver.gradient_convergence_tolerance > 0 and np.linalg.norm(g) < solver.gradient_convergence_tolerance:
            break
        step = -solver.rate * g / (div if div is not None else 1.0)
        sn = np.linalg.norm(step)
        if sn > solver.max_step:
            step *= solver.max_step / sn
        x = x + step
    return x
