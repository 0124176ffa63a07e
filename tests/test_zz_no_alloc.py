"""flash_skin_eval_static: the entry point that allocates nothing (include/flashsdf.h; no counterpart in the
reference).  Its device temporaries live in a file-scope
`static __device__` array; the test calls it through the C ABI, twice in a row with different states and sizes,
and checks each result against flash_skin_eval, the existing entry point, within the tolerance of the parity
tests (1e-5 relative).  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

import flash
from flash import _lib

pytestmark = pytest.mark.gpu

RTOL = 1e-5
MAX_QUERY = 65536          # FLASH_STATIC_MAX_QUERY
ERR_TOO_LARGE = -4


def _state(m, rng):
    nx, nq = flash.num_states(m), flash.num_positions(m)
    x = flash.gradientdescent.flatten(flash.ManipulatorState(m))
    return np.ascontiguousarray(x + np.r_[0.2 * rng.standard_normal(nq), 0.02 * rng.standard_normal(nx - nq)])


def _existing(m, x, pts):
    n = len(pts)
    dist, arg = np.empty(n), np.empty(n, dtype=np.int32)
    m.check(_lib.lib().flash_skin_eval(m.ctx(), x.ctypes.data_as(_lib.c_double_p), pts.ctypes.data_as(_lib.c_double_p), n,
                                       dist.ctypes.data_as(_lib.c_double_p), arg.ctypes.data_as(_lib.c_int32_p)))
    return dist, arg


def _static(m, x, pts):
    n = len(pts)
    dist, arg = np.full(n, np.nan), np.full(n, -7, dtype=np.int32)
    m.check(_lib.lib().flash_skin_eval_static(m.ctx(), x.ctypes.data, pts.ctypes.data, n, dist.ctypes.data, arg.ctypes.data))
    return dist, arg


def _same(a, b):
    (da, aa), (db, ab) = a, b
    assert np.isfinite(da).all()
    assert np.abs(da - db).max() <= RTOL * max(1.0, np.abs(db).max()), np.abs(da - db).max()
    assert (aa == ab).all()


def test_static_scratch_entry_point_matches_flash_skin_eval_on_two_calls_in_a_row():
    rng = np.random.default_rng(41)
    m = flash.models.irb_and_squishable()          # seven hulls + a table + a deformable RBF skin: every evaluator
    # first call: the largest size the array holds; second call: another state and fewer points, so anything the
    # array kept from the first call (tables, points, distances, the status word) would show
    x1, x2 = _state(m, rng), _state(m, rng)
    p1 = rng.uniform([-0.8, -0.8, -0.3], [1.0, 0.8, 1.2], (MAX_QUERY, 3))
    p2 = rng.uniform([-0.5, -0.5, -0.1], [0.7, 0.5, 0.9], (12345, 3))
    s1 = _static(m, x1, p1)
    s2 = _static(m, x2, p2)
    _same(s1, _existing(m, x1, p1))
    _same(s2, _existing(m, x2, p2))
    assert np.abs(s1[0][:12345] - s2[0]).max() > 1e-3   # the two calls really differ
    # a model with no hull at all, on the same array right after
    b = flash.models.beanbag()
    xb = flash.gradientdescent.flatten(flash.ManipulatorState(b))
    q = np.array([[100.0, 0.0, 0.0], [0.3, 0.2, -0.1]])
    sb = _static(b, xb, q)
    _same(sb, _existing(b, xb, q))
    assert abs(sb[0][0] - 99.0) < 0.02 * 99.0          # test/runtests.jl:17


def test_static_scratch_entry_point_rejects_sizes_beyond_its_maximum_and_allocates_nothing():
    import torch
    rng = np.random.default_rng(42)
    m = flash.models.irb140()
    x = _state(m, rng)
    L = _lib.lib()
    pts = rng.uniform(-1, 1, (MAX_QUERY + 1, 3))
    dist = np.empty(len(pts))
    rc = L.flash_skin_eval_static(m.ctx(), x.ctypes.data, pts.ctypes.data, len(pts), dist.ctypes.data, None)
    assert rc == ERR_TOO_LARGE
    msg = L.flash_last_error(m.ctx()).decode()
    assert "FLASH_STATIC_MAX_QUERY" in msg and "65536" in msg, msg
    assert L.flash_skin_eval_static(m.ctx(), x.ctypes.data, None, 0, None, None) == 0      # empty query
    # device memory in use does not move across a call (the first call has loaded the module and looked the array up)
    pts = pts[:40000]
    ref = _existing(m, x, pts)
    _same(_static(m, x, pts), ref)
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    got = _static(m, x, pts)
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0, (free0, free1)
    _same(got, ref)
    # the caller's buffers may live on the device
    dx, dp = torch.from_numpy(x).cuda(), torch.from_numpy(pts).cuda()
    dd, da = torch.empty(len(pts), dtype=torch.float64, device="cuda"), torch.empty(len(pts), dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    m.check(L.flash_skin_eval_static(m.ctx(), dx.data_ptr(), dp.data_ptr(), len(pts), dd.data_ptr(), da.data_ptr()))
    _same((dd.cpu().numpy(), da.cpu().numpy()), ref)
