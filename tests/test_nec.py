"""NEC deck handling (fdtd_b200/nec.py) against the behaviour of the reference's nec-file crate and of
cem-app/src/composer/file_formats/nec.rs, plus the CPU restatement of the scene rasteriser (oracle/scene_ref.py).
The reference ships no deck and no tests for this crate (nec-file/examples/read_file.rs only prints), so the expected
values below are derived by hand from the cited lines."""
import numpy as np
import pytest

from fdtd_b200 import configs, nec
from fdtd_b200.scene import CoordinateTransformations, SceneDomain, Shape
from fdtd_b200.solver import Material, Resolution

F = np.float32


def test_dipole_deck_parses_like_the_reference():
    d = nec.NecFile.from_text(configs.NEC_DIPOLE_DECK)
    assert d.comments == [" half-wave dipole"]            # remainder after the card identifier, leading blank kept (parser.rs:122)
    assert d.ground_plane is None and d.symmetry == ("none",)
    (tag, g), = d.geometry
    assert tag == 1 and g.kind == "wire" and g.num_segments == 21
    assert g.length == F(0.5)
    # interpreter.rs:322-331: translation = first end, rotation = axis-angle(…, 0.0) = identity
    assert np.array_equal(g.transform, np.array([[1, 0, 0, 0], [0, 1, 0, 0], [0, 0, 1, -0.25], [0, 0, 0, 1]], dtype=F))
    assert g.segments.kind == "fixed" and g.segments.radius == F(0.005)


def test_cards_after_ge_are_not_read_and_missing_ge_is_an_error():
    nec.NecFile.from_text("CE\nGW 1 3 0 0 0 0 0 1 0.01\nGE\nthis line is never tokenised\n")
    with pytest.raises(nec.UnexpectedEnd):
        nec.NecFile.from_text("CM x\nCE\nGW 1 3 0 0 0 0 0 1 0.01\n")
    with pytest.raises(nec.UnexpectedCardForSection):
        nec.NecFile.from_text("GW 1 3 0 0 0 0 0 1 0.01\nGE\n")        # geometry card inside the comments section
    with pytest.raises(nec.InvalidParameter):
        nec.NecFile.from_text("CE\nXX 1\nGE\n")                        # unknown identifier fails CardType::from_str
    with pytest.raises(nec.InvalidParameter):
        nec.NecFile.from_text("CE\nGW 1 x 0 0 0 0 0 1 0.01\nGE\n")
    with pytest.raises(nec.UnexpectedEnd):
        nec.NecFile.from_text("CE\nGW 1 3 0 0 0\nGE\n")


def test_ce_comment_and_ground_plane_flag():
    d = nec.NecFile.from_text("CM a\nCM  b\nCE last\nGW 1 1 0 0 0 1 0 0 0.1\nGE 1\n")
    assert d.comments == [" a", "  b", " last"] and d.ground_plane is True
    assert nec.NecFile.from_text("CE\nGW 1 1 0 0 0 1 0 0 0.1\nGE -1\n").ground_plane is False
    assert nec.NecFile.from_text("CE\nGW 1 1 0 0 0 1 0 0 0.1\nGE 7\n").ground_plane is None    # unparsable → default


def test_wire_segments_fixed_and_tapered():
    fixed = list(nec.WireSegments("fixed", radius=F(0.01)).dimensions(4, F(2.0)))
    assert fixed == [("flat", F(0.5), F(0.01))] * 4
    # GW with radius 0 defers to the next card (read as taper parameters straight after its identifier, parser.rs:249-262)
    d = nec.NecFile.from_text("CE\nGW 2 3 0 0 0 0 0 7 0\nGC 2.0 0.01 0.04\nGE\n")
    g = d.geometry[0][1]
    assert g.segments.kind == "tapered"
    dims = list(g.segments.dimensions(g.num_segments, g.length))
    lengths = [float(x[1]) for x in dims]
    assert lengths == pytest.approx([1.0, 2.0, 4.0], rel=1e-6)          # 7·(1−2)/(1−2³) = 1, ratio 2
    assert [float(x[2]) for x in dims] == pytest.approx([0.01, 0.02, 0.04], rel=1e-5)   # radius ratio (0.04/0.01)^(1/2) = 2
    assert float(dims[-1][3]) == pytest.approx(0.08, rel=1e-5)
    with pytest.raises(NotImplementedError):
        nec.wire_cylinders(d)                                             # `todo!("truncated cone shape")`


def test_gm_moves_copies_and_retags():
    base = "CE\nGW 1 2 0 0 0 0 0 1 0.01\n"
    moved = nec.NecFile.from_text(base + "GM 0 0 0 0 0 1 2 3\nGE\n").geometry
    assert [t for t, _ in moved] == [1] and np.allclose(moved[0][1].transform[:3, 3], [1, 2, 3])
    retag = nec.NecFile.from_text(base + "GM 5 0 0 0 0 1 0 0\nGE\n").geometry
    assert [t for t, _ in retag] == [6]
    copies = nec.NecFile.from_text(base + "GM 1 2 0 0 90 1 0 0\nGE\n").geometry
    assert [t for t, _ in copies] == [1, 2, 3]
    # copy i carries base^(i): rotate 90° about z then translate (1,0,0), applied cumulatively
    assert np.allclose(copies[1][1].transform[:3, 3], [1, 0, 0], atol=1e-6)
    assert np.allclose(copies[2][1].transform[:3, 3], [1, 1, 0], atol=1e-6)
    # tag_start filters which structures are touched
    two = "CE\nGW 1 2 0 0 0 0 0 1 0.01\nGW 4 2 5 0 0 5 0 1 0.01\n"
    part = nec.NecFile.from_text(two + "GM 0 0 0 0 0 0 0 9 4\nGE\n").geometry
    assert np.allclose(part[0][1].transform[:3, 3], [0, 0, 0]) and np.allclose(part[1][1].transform[:3, 3], [5, 0, 9])


def test_gs_scales_and_gx_is_deferred_until_the_next_gm():
    d = nec.NecFile.from_text("CE\nGW 1 2 1 0 0 1 0 1 0.01\nGS 2\nGE\n").geometry[0][1]
    assert d.length == F(2.0) and d.segments.radius == F(0.02) and np.allclose(d.transform[:3, 3], [2, 0, 0])
    # GX only queues the mirrored copies (interpreter.rs:338-371 never applies them) …
    lone = nec.NecFile.from_text("CE\nGW 1 2 1 0 0 1 0 1 0.01\nGX 10 100\nGE\n")
    assert [t for t, _ in lone.geometry] == [1] and lone.symmetry == ("planar", 0b001)
    # … a later GM flushes the queue (after moving the structures that existed before)
    both = nec.NecFile.from_text("CE\nGW 1 2 1 0 0 1 0 1 0.01\nGX 10 100\nGM 0 0 0 0 0 0 0 0\nGE\n").geometry
    assert [t for t, _ in both] == [1, 11] and np.allclose(both[1][1].transform[:3, 3], [-1, 0, 0])
    with pytest.raises(nec.InvalidParameter):
        nec.NecFile.from_text("CE\nGW 1 2 1 0 0 1 0 1 0.01\nGX 10 12\nGE\n")


def test_wire_cylinders_follow_the_applications_layout():
    cyl = nec.wire_cylinders(nec.NecFile.from_text(configs.NEC_DIPOLE_DECK))
    assert len(cyl) == 21
    seg = F(F(0.5) / F(21))
    for i, c in enumerate(cyl):
        # strung along +Y from the first end, whatever the wire's direction (nec.rs:56-69); cylinder i is centred at i·len
        assert c.centre == (F(0), F(F(i) * seg), F(-0.25)) and c.half_height == F(F(0.5) * seg) and c.radius == F(0.005)


def test_scene_ref_first_match_wins_and_primitives():
    from oracle import scene_ref

    res = Resolution((0.5, 0.5, 0.5), 0.1)
    tf = CoordinateTransformations.for_fdtd(res, (16, 12, 10), aabb_mins=(-4.0, -3.0, -2.5))
    a, b, c = Material(relative_permittivity=2.0), Material(eletrical_conductivity=3.0), Material(relative_permeability=4.0)
    scene = SceneDomain([Shape.ball(1.0, a, (0.0, 0.0, 0.0)), Shape.cuboid((2.0, 0.5, 0.5), b, (0.0, 0.0, 0.0)),
                         Shape.cylinder(1.0, 0.6, c, (2.0, 1.0, 1.0))], tf)
    assert scene.palette == [Material.VACUUM, a, b, c] and scene.labels == [1, 2, 3]
    ids = scene_ref.rasterize_scene_domain(scene).reshape(10, 12, 16)

    def at(p):
        q = tf.transform_point_from_world_to_solver(p)
        return int(ids[q[2], q[1], q[0]])

    assert at((0, 0, 0)) == 1 and at((1.0, 0, 0)) == 1          # ball boundary is inclusive
    assert at((1.5, 0, 0)) == 2 and at((2.0, 0.5, 0.5)) == 2     # cuboid where the ball does not reach; faces inclusive
    assert at((2.0, 2.0, 1.0)) == 3 and at((2.5, 1.0, 1.0)) == 3 and at((2.0, 2.5, 1.0)) == 0
    assert at((-3.5, -2.5, -2.0)) == 0
    # slab rasterisation = the same planes of the whole lattice
    assert np.array_equal(scene_ref.rasterize_scene_domain(scene, 3, 7), ids[3:7].reshape(-1))
    assert tf.transform_point_from_world_to_solver((100.0, 0, 0)) is None
    assert np.array_equal(tf.transform_point_from_solver_to_world((8, 6, 5)), np.zeros(3, dtype=F))


def test_config5_workload_shape():
    from oracle import scene_ref

    w = configs.config5_nec_dipole(64, 10)
    ids = scene_ref.rasterize_scene_domain(w.scene).reshape(64, 64, 64)
    zz, yy, xx = np.nonzero(ids)
    assert set(xx) == {32} and set(zz) == {32}                    # radius < dx: a one-cell-thick wire
    assert 32 not in set(yy) and len(yy) == 14                    # the feed segment stays vacuum
    assert len(w.probes) == 18 and w.sources[0][0] == (32, 32, 32)
    assert len(w.palette) == 2 and w.palette[1].eletrical_conductivity == 1.0e4
