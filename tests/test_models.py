"""Host-side mirror of src/models.jl / src/Flash.jl / src/depthdata.jl (no device needed)."""
import io
import os
import struct

import numpy as np
import pytest

import flash


def test_models_construct():
    """test/runtests.jl:6-11"""
    beanbag = flash.models.beanbag()
    squishable = flash.models.squishable()
    arm = flash.models.two_link_arm()
    deformable_arm = flash.models.two_link_arm(True)
    assert flash.num_states(beanbag) == 7 + 3 * 6           # src/Flash.jl:90
    assert flash.num_states(squishable) == 43               # examples/squishable.ipynb cell 9
    assert flash.num_states(arm) == 2 == flash.num_states(deformable_arm)
    s = arm.surfaces[0]
    assert isinstance(s, flash.RigidInterpolatingSkin)      # `deformable` is ignored, src/models.jl:68
    assert len(s.surface_points) == 40 and len(s.skeleton_points) == 6   # 46 centres
    assert repr(arm) == "Manipulator with 3 links and 1 surfaces"


def test_squishable_control_points():
    """SURVEY 9 values computed from src/models.jl:114-129."""
    pts = np.array([p for _, p in flash.models.squishable().surfaces[0].surface_points])
    assert pts.shape == (12, 3)
    assert np.allclose(sorted(np.abs(pts[:4]).ravel()), [0] * 4 + [0.15] * 8, atol=1e-3)
    assert np.allclose(np.abs(pts[4:8, [0, 2]]), 0.15492, atol=1e-4)
    assert np.allclose(np.abs(pts[8:, :2]), 0.18499, atol=1e-4)


def test_irb140_fixture_and_combined_scene():
    m = flash.models.irb140()
    assert repr(m) == "Manipulator with 8 links and 7 surfaces"     # examples/irb140.ipynb cell 3
    assert flash.num_states(m) == 6
    assert [len(s.vertices) for s in m.surfaces] == [52] * 7
    b = m.mechanism.bodies
    assert np.allclose(b[1].joint_to_parent[:, 3], [0, 0, 0.1095]) and np.allclose(b[1].axis, [0, -1, 0])
    assert np.allclose(b[6].joint_to_parent[:, 3], [0.065, 0, 0]) and np.allclose(b[6].axis, [1, 0, 0])
    c = flash.models.irb_and_squishable()
    assert repr(c) == "Manipulator with 10 links and 9 surfaces"   # examples/irb_and_squishable.ipynb cell 4
    assert flash.num_states(c) == 63 and flash.num_positions(c) == 27
    assert [type(s).__name__ for s in c.surfaces] == ["ConvexGeometry"] * 7 + ["DeformableInterpolatingSkin", "ConvexGeometry"]
    assert c.surfaces[7].surface_points[0][0] == 7 and c.surfaces[8].body == 8   # body indices shifted by merge!


def test_state_layout_flatten_unflatten():
    """src/gradientdescent.jl:9-17 and src/Flash.jl:92-125."""
    m = flash.models.irb_and_squishable()
    st = flash.ManipulatorState(m)
    x = flash.gradientdescent.flatten(st)
    assert x.shape == (63,)
    assert x[0] == 1 and x[13] == 1 and x[20] == 1 and x.sum() == 3     # identity quaternions [w x y z | t]
    x2 = np.arange(63.0)
    flash.gradientdescent.unflatten_(st, x2)
    assert np.array_equal(st.mechanism_state.q, x2[:27]) and np.array_equal(st.deformation_data, x2[27:])
    skin = m.surfaces[7]
    assert np.array_equal(st.deformations[skin][2], x2[27 + 6:27 + 9])   # 3-wide views in surface-point order
    with pytest.raises(ValueError):
        flash.gradientdescent.unflatten_(st, np.zeros(5))
    with pytest.raises(ValueError):
        flash.set_configuration_(st, np.zeros(3))


def _write_stl(path, tris):
    with open(path, "wb") as f:
        f.write(b"\0" * 80 + struct.pack("<I", len(tris)))
        for t in tris:
            f.write(struct.pack("<12fH", 0, 0, 0, *np.asarray(t, np.float32).ravel(), 0))


def test_load_urdf_with_binary_stl(tmp_path):
    """src/models.jl:154-171 on a synthetic two-link URDF with a tetrahedron mesh."""
    tet = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [0, 0, 1]], float)
    tris = [tet[[0, 2, 1]], tet[[0, 1, 3]], tet[[0, 3, 2]], tet[[1, 2, 3]]]
    os.makedirs(tmp_path / "pkg" / "meshes")
    _write_stl(tmp_path / "pkg" / "meshes" / "tet.stl", tris)
    urdf = tmp_path / "robot.urdf"
    urdf.write_text("""<robot name="r">
      <link name="base"><visual><origin xyz="0 0 0.5" rpy="0 0 1.57079"/><geometry><mesh filename="package://pkg/meshes/tet.stl"/></geometry></visual></link>
      <link name="arm"><visual><geometry><mesh filename="package://pkg/meshes/tet.stl" scale="2 2 2"/></geometry></visual></link>
      <joint name="j" type="revolute"><parent link="base"/><child link="arm"/><origin xyz="1 2 3" rpy="0 0 0"/><axis xyz="0 0 1"/></joint>
    </robot>""")
    m = flash.models.load_urdf(str(urdf), package_path=[str(tmp_path)])
    assert repr(m) == "Manipulator with 3 links and 2 surfaces"
    assert flash.num_states(m) == 1
    assert len(m.surfaces[0].vertices) == 4 and np.allclose(m.surfaces[1].vertices.max(), 2.0)   # welded 12 -> 4, scaled
    assert np.allclose(m.surfaces[0].frame[:, 3], [0, 0, 0.5]) and abs(m.surfaces[0].frame[0, 1] + 1) < 1e-4
    assert np.allclose(m.mechanism.bodies[1].joint_to_parent[:, 3], [1, 2, 3])
    with pytest.raises(IOError, match="unsupported URDF joint"):
        bad = tmp_path / "bad.urdf"
        bad.write_text(urdf.read_text().replace('type="revolute"', 'type="prismatic"'))
        flash.models.load_urdf(str(bad), package_path=[str(tmp_path)])


def test_read_point_cloud():
    """src/depthdata.jl:19-30"""
    txt = "1.38246, 0.768824, 1.48581\n0.414323, -0.0773639, 0.959196, 0.207843, 0.207843, 0.219608\n0.412916, -0.0752139, 0.959346, 0.2, 0.3, 0.4\n"
    pc = flash.depthdata.read_point_cloud(io.StringIO(txt))
    assert np.allclose(pc.camera_origin, [1.38246, 0.768824, 1.48581])
    assert pc.positions.shape == (2, 3) and np.allclose(pc.colors[1], [0.2, 0.3, 0.4])
    assert "containing 2 points" in repr(pc)
    gold = os.path.join(os.path.dirname(__file__), "golden", "squishable_unsquished_xyz.npy")
    assert np.load(gold).shape == (25571, 3)      # examples/squishable.ipynb cell 4


def test_read_point_cloud_through_the_c_abi(tmp_path):
    """flash_read_point_cloud (csrc/flash_urdf.h), the reader behind flash_points_set_file: same numbers as the
    stream reader, FLASH_ERR_IO for missing / malformed files."""
    rng = np.random.default_rng(3)
    xyz = rng.uniform(-1, 1, (257, 3)); rgb = rng.uniform(0, 1, (257, 3))
    path = tmp_path / "cloud_xyzrgb.txt"
    with open(path, "w") as f:
        f.write("1.38246, 0.768824, 1.48581\n")
        for p, c in zip(xyz, rgb):
            f.write(", ".join(repr(float(v)) for v in list(p) + list(c)) + "\n")
    pc = flash.depthdata.read_point_cloud(str(path))
    ref = flash.depthdata.read_point_cloud(open(path))
    assert np.array_equal(pc.camera_origin, ref.camera_origin) and np.array_equal(pc.positions, ref.positions) and np.array_equal(pc.colors, ref.colors)
    assert np.array_equal(pc.positions, xyz) and "containing 257 points" in repr(pc)
    with pytest.raises(IOError, match="cannot open"):
        flash.depthdata.read_point_cloud(str(tmp_path / "missing.txt"))
    bad = tmp_path / "bad.txt"
    bad.write_text("0, 0, 0\n1, 2, 3\n")
    with pytest.raises(IOError, match="does not hold"):
        flash.depthdata.read_point_cloud(str(bad))


GOLD_URDF = os.path.join(os.path.dirname(__file__), "golden", "IRB140", "urdf", "irb_140_convhull.urdf")


def test_c_urdf_loader_reproduces_the_committed_fixture_and_the_survey_tables():
    """flash_urdf_load (csrc/flash_urdf.h) on the reference's own irb_140_convhull.urdf + *_chull.stl (fixtures copied
    from examples/data/IRB140/): the description equals the committed JSON fixture the GPU tests use, the joint chain
    and the visual origins are those of SURVEY 9 (examples/data/IRB140/urdf/irb_140_convhull.urdf:290-330), every mesh
    welds to 52 vertices with the tabulated AABB, and the *_chull.obj twins hold the same vertex sets."""
    m = flash.models.load_urdf(GOLD_URDF, package_path=[os.path.join(os.path.dirname(__file__), "golden")])
    ref = flash.models.irb140()
    assert repr(m) == "Manipulator with 8 links and 7 surfaces"      # examples/irb140.ipynb:261
    assert [b.name for b in m.mechanism.bodies] == ["base_link"] + ["link_%d" % k for k in range(1, 7)]
    for b, r in zip(m.mechanism.bodies, ref.mechanism.bodies):
        assert (b.name, b.parent, b.joint_type) == (r.name, r.parent, r.joint_type)
        assert np.array_equal(b.axis, r.axis) and np.allclose(b.joint_to_parent, r.joint_to_parent, rtol=0, atol=1e-15)
    for s_, r in zip(m.surfaces, ref.surfaces):
        assert s_.body == r.body and np.allclose(s_.frame, r.frame, rtol=0, atol=1e-15)
        assert np.array_equal(s_.vertices, r.vertices)                # fp32 STL values, bit for bit, same welding order
    # joints
    joints = [((0, 0, 0.1095), (-1.57079, 0, 0), (0, -1, 0)), ((0.07, -0.2425, 0), (0, 0, 0), (0, 0, 1)), ((0, -0.360, 0), (0, 0, 0), (0, 0, 1)),
              ((0.2185, 0, 0), (0, 0, 0), (1, 0, 0)), ((0.1615, 0, 0), (0, 0, 0), (0, 0, -1)), ((0.065, 0, 0), (0, 0, 0), (1, 0, 0))]
    for k, (xyz, rpy, axis) in enumerate(joints):
        b = m.mechanism.bodies[k + 1]
        assert b.parent == k and b.joint_type == flash.JOINT_REVOLUTE and np.allclose(b.axis, axis) and np.allclose(b.joint_to_parent[:, 3], xyz)
        c, s = np.cos(rpy[0]), np.sin(rpy[0])
        assert np.allclose(b.joint_to_parent[:, :3], [[1, 0, 0], [0, c, -s], [0, s, c]], atol=1e-15)
    # visual origins and mesh AABBs (metres, mesh frame)
    vis = [(0, 0, 0), (0, 0.1095, 0), (-0.07, 0.352, 0), (-0.07, 0.712, 0), (-0.2885, 0.712, 0), (-0.450, 0.712, 0), (-0.515, 0.712, 0)]
    aabb = [((-0.3132, -0.1525, -0.2019), (0.1784, 0.0005, 0.2018)), ((-0.1556, -0.4543, -0.1365), (0.1943, -0.0889, 0.2044)),
            ((-0.0293, -0.8027, -0.1755), (0.1540, -0.2831, 0.0989)), ((-0.0790, -0.8091, -0.0832), (0.3088, -0.6090, 0.0850)),
            ((0.2885, -0.7681, -0.0645), (0.5078, -0.6559, 0.0705)), ((0.4079, -0.7543, -0.0155), (0.4922, -0.6697, 0.0166)),
            ((0.4696, -0.7370, -0.0250), (0.5150, -0.6867, 0.0250))]
    names = ["base_link", "link1", "link2", "link3", "link4", "link5", "link6"]
    for k, s_ in enumerate(m.surfaces):
        assert s_.body == k and np.allclose(s_.frame[:, 3], vis[k]) and len(s_.vertices) == 52
        assert np.allclose(s_.vertices.min(0), aabb[k][0], atol=6e-5) and np.allclose(s_.vertices.max(0), aabb[k][1], atol=6e-5)
        obj = np.array([[float(v) for v in l.split()[1:4]] for l in open(os.path.join(os.path.dirname(GOLD_URDF), "meshes", names[k] + "_chull.obj")) if l.startswith("v ")])
        obj = np.unique(obj, axis=0)
        assert len(obj) == 52
        d = np.abs(obj[:, None, :] - s_.vertices[None, :, :]).max(2)
        assert d.min(1).max() < 1e-6 and d.min(0).max() < 1e-6      # the same 52 points (OBJ text keeps 6 digits)
    # the base link's visual origin carries the literal -1.57079 (not -pi/2)
    c, s = np.cos(-1.57079), np.sin(-1.57079)
    assert np.allclose(m.surfaces[0].frame[:, :3], [[1, 0, 0], [0, c, -s], [0, s, c]], atol=1e-15)


def test_c_urdf_loader_error_paths(tmp_path):
    import ctypes as C
    from flash import _lib
    L = _lib.lib()
    h = C.c_void_p()
    assert L.flash_urdf_load(b"/nonexistent/robot.urdf", None, 0, C.byref(h)) == -7 and b"cannot open" in L.flash_last_error(None)   # FLASH_ERR_IO
    bad = tmp_path / "bad.urdf"
    bad.write_text("<robot><link name='a'></robot>")
    assert L.flash_urdf_load(str(bad).encode(), None, 0, C.byref(h)) == -7 and b"does not match" in L.flash_last_error(None)
    two = tmp_path / "two.urdf"
    two.write_text("<robot><link name='a'/><link name='b'/></robot>")
    assert L.flash_urdf_load(str(two).encode(), None, 0, C.byref(h)) == -7 and b"exactly one root" in L.flash_last_error(None)
    nomesh = tmp_path / "nomesh.urdf"
    nomesh.write_text("<robot><link name='a'><visual><geometry><mesh filename='package://p/x.stl'/></geometry></visual></link></robot>")
    assert L.flash_urdf_load(str(nomesh).encode(), None, 0, C.byref(h)) == -7 and b"cannot resolve" in L.flash_last_error(None)
    ok = C.c_void_p()
    gold_dir = os.path.join(os.path.dirname(__file__), "golden").encode()
    arr = (C.c_char_p * 1)(gold_dir)
    assert L.flash_urdf_load(GOLD_URDF.encode(), arr, 1, C.byref(ok)) == 0
    assert L.flash_urdf_set_joint_type(ok, 0, 2) == 0 and L.flash_urdf_desc(ok).contents.joint_type[0] == 2     # floating base
    assert L.flash_urdf_set_joint_type(ok, 99, 2) == -1
    assert L.flash_urdf_surface_mesh(ok, 3).endswith(b"link3_chull.stl")
    L.flash_urdf_free(ok)


def test_packed_model_layout():
    from flash._desc import PackedModel
    p = PackedModel(flash.models.irb_and_squishable())
    assert p.desc.n_bodies == 9 and p.desc.n_surfaces == 9
    assert list(p.surface_kind) == [0] * 7 + [2, 0]
    assert list(p.point_offset) == [0, 52, 104, 156, 208, 260, 312, 364, 377, 385]
    assert list(p.n_surface_points[7:]) == [12, 8]
    assert list(p.joint_type) == [2, 1, 1, 1, 1, 1, 1, 2, 2]
