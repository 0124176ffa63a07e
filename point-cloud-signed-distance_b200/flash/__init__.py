"""flash -- host-side mirror of the Julia package `Flash` over the libflashsdf C ABI.

The reference's host language (Julia 0.5) is not installed in this image, so the interface of
src/Flash.jl, src/models.jl, src/gradientdescent.jl, src/tracking.jl, src/depthsensors.jl and
src/depthdata.jl is mirrored here in Python with the same names, argument meaning and error
behaviour; every numeric call goes through the C ABI in include/flashsdf.h (ctypes), i.e. through the
CUDA kernels.  There is no CPU path: importing works anywhere, computing needs a B200.
The Julia `ccall` shim a maintainer would add is julia/FlashCUDA.jl.
"""
from .core import (  # noqa: F401
    BodyGeometry, ConvexGeometry, DeformableInterpolatingSkin, RigidInterpolatingSkin,
    Manipulator, ManipulatorState, Mechanism, Body, num_states, num_deformations, num_positions,
    set_configuration_, skin, surface_points, FlashError, drawing_region, contour_mesh, draw,
    JOINT_FIXED, JOINT_REVOLUTE, JOINT_QUAT_FLOATING,
)
from . import models, gradientdescent, tracking, depthsensors, depthdata  # noqa: F401

Models = models
GradientDescent = gradientdescent
Tracking = tracking
DepthSensors = depthsensors
DepthData = depthdata
