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
    with pytest.raises(ValueError, match="unsupported URDF joint"):
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


def test_packed_model_layout():
    from flash._desc import PackedModel
    p = PackedModel(flash.models.irb_and_squishable())
    assert p.desc.n_bodies == 9 and p.desc.n_surfaces == 9
    assert list(p.surface_kind) == [0] * 7 + [2, 0]
    assert list(p.point_offset) == [0, 52, 104, 156, 208, 260, 312, 364, 377, 385]
    assert list(p.n_surface_points[7:]) == [12, 8]
    assert list(p.joint_type) == [2, 1, 1, 1, 1, 1, 1, 2, 2]
