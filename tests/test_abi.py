"""The C-ABI shared library: loads on a CPU-only box, exports every symbol include/flashsdf.h
declares, and refuses to compute without a device (no CPU fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import flash
from flash import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _have_gpu():
    try:
        cuda = C.CDLL("libcuda.so.1")
        n = C.c_int()
        return cuda.cuInit(0) == 0 and cuda.cuDeviceGetCount(C.byref(n)) == 0 and n.value > 0
    except OSError:
        return False


def test_header_symbols_all_exported():
    hdr = open(os.path.join(ROOT, "include", "flashsdf.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = sorted(set(re.findall(r"\b(flash_[a-z0-9_]+)\s*\(", hdr)))
    assert len(declared) >= 24
    L = _lib.lib()
    missing = [s for s in declared if not hasattr(L, s)]
    assert not missing, missing
    assert sorted(_lib.SYMBOLS) == declared
    assert L.flash_abi_version() == 2


def test_library_has_no_torch_or_nccl_link_dependency():
    import subprocess
    out = subprocess.run(["ldd", _lib.LIB_PATH], capture_output=True, text=True).stdout
    assert "torch" not in out and "nccl" not in out and "libcudart" not in out  # static cudart, NCCL via dlopen


def test_kinect_rays_host_helper():
    r = flash.depthsensors.generateKinectRays(3, 5)
    assert r.shape == (3, 5, 3) and np.allclose(np.linalg.norm(r, axis=2), 1)
    assert np.allclose(r[1, 2], [0, 0, 1])


@pytest.mark.skipif(_have_gpu(), reason="CPU-only behaviour")
def test_compute_fails_loudly_without_a_device():
    m = flash.models.squishable()
    with pytest.raises(flash.FlashError, match="no CUDA device"):
        flash.gradientdescent.CostFunctor(m, np.zeros((4, 3)))


def test_bad_descriptions_are_rejected_before_touching_the_device():
    from flash._desc import PackedModel
    L = _lib.lib()
    p = PackedModel(flash.models.beanbag())
    p.desc.abi_version = 7
    out = C.c_void_p()
    assert L.flash_model_create(C.byref(p.desc), None, C.byref(out)) == -1
    assert b"abi_version" in L.flash_last_error(None)
    assert L.flash_model_create(None, None, C.byref(out)) == -1
    assert L.flash_num_states(None) == -1
