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


def _driver():
    """tests/abi_driver: the plain-C caller of every ABI entry point (built by __graft_entry__.build())"""
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = os.path.join(root, "tests", "abi_driver")
    src = os.path.join(root, "tests", "abi_driver.c")
    libdir = os.path.join(root, "point-cloud-signed-distance_b200", "lib")
    if not os.path.exists(exe) or os.path.getmtime(exe) < os.path.getmtime(src):
        subprocess.check_call(["gcc", "-O1", "-std=c99", "-Wall", "-o", exe, src, "-L" + libdir, "-lflashsdf", "-lm", "-Wl,-rpath," + libdir])
    gold = os.path.join(root, "tests", "golden")
    return [exe, os.path.join(gold, "IRB140", "urdf", "irb_140_convhull.urdf"), gold]


def test_plain_c_driver_host_side(tmp_path):
    """Every header symbol is callable from C; without a device the file-ingestion entry points work and every compute
    entry point fails loudly (FLASH_ERR_NO_DEVICE) -- there is no CPU path to fall back to."""
    import subprocess
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present: covered by test_plain_c_driver_on_the_gpu")
    r = subprocess.run(_driver() + [str(tmp_path)], capture_output=True, text=True)
    assert r.returncode == 0 and "FAIL " not in r.stdout and "PASSED (host-side checks only" in r.stdout, r.stdout + r.stderr


@pytest.mark.gpu
def test_plain_c_driver_on_the_gpu(tmp_path):
    """The same driver on a GPU box: cost / gradient / argmin against the closed form of a box, the beanbag known answer
    of test/runtests.jl:17, frames, async upload + swap, files, ray casting, the on-device solver, iso-surface
    extraction, grid mode, deterministic + multi-device contexts, URDF loading -- all from plain C."""
    import subprocess
    r = subprocess.run(_driver() + [str(tmp_path)], capture_output=True, text=True)
    print(r.stdout)
    assert r.returncode == 0 and "FAIL " not in r.stdout and r.stdout.strip().endswith("PASSED"), r.stdout + r.stderr


def test_documented_limits_match_the_sources():
    """INTEGRATION.md's table of compiled-in limits quotes the constants of the sources."""
    src = open(os.path.join(ROOT, "point-cloud-signed-distance_b200", "csrc", "flash_dev.h")).read()
    ing = open(os.path.join(ROOT, "point-cloud-signed-distance_b200", "csrc", "flash_ingest.cuh")).read()
    hdr = open(os.path.join(ROOT, "include", "flashsdf.h")).read()
    doc = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    const = lambda text, name: int(re.search(r"constexpr int %s = (\d+);" % name, text).group(1))
    assert "| bodies / surfaces per model | %d / %d |" % (const(src, "MAX_BODIES"), const(src, "MAX_SURFACES")) in doc
    assert "| convex surfaces | %d |" % const(src, "MAX_CONVEX") in doc
    assert "| RBF skins; control points per skin | %d; %d |" % (const(src, "MAX_SKINS"), const(src, "MAX_SKIN_PTS")) in doc
    assert "| frames per batch | %d |" % const(ing, "MAX_FRAMES") in doc
    assert int(re.search(r"#define FLASH_MAX_FRAMES (\d+)", hdr).group(1)) == const(ing, "MAX_FRAMES")
    q = int(re.search(r"#define FLASH_STATIC_MAX_QUERY (\d+)", hdr).group(1))
    assert "{:,} query points".format(q) in doc and "devices[16]" in hdr and "| devices per context | 16 |" in doc
