"""The reference's own unit tests (tests/unit_tests.cpp, GoogleTest -- unavailable here) re-hosted as
known-answer tests of the oracle restatement, and, when oracle/_ref is present, of the shim-built
reference too (which also validates the shim's cross/dot).  Line numbers refer to unit_tests.cpp."""
import math

import numpy as np
import pytest

from oracle_lib import Oracle, have_reference

KINDS = ["port"] + (["reference"] if have_reference() else [])


@pytest.fixture(scope="module", params=KINDS)
def o(request):
    return Oracle(request.param)


def test_collinear_points(o):  # :13-31
    assert o.are_collinear((1, 0), (2, 1), (3, 2), 1e-9)
    assert not o.are_collinear((1, 0), (2, 1), (1, 4), 1e-9)


def test_collinear_signed_distance_quirk(o):
    # S7: the distance is signed (math_utils.hpp:150-154): a far point on the negative side "is collinear"
    assert o.are_collinear((1, 0), (2, 1), (4, -100), 1.0)
    assert not o.are_collinear((1, 0), (2, 1), (-100, 4), 1.0)


def test_convex_hull_standard(o):  # :37-67
    pts = [(0, 3), (2, 2), (1, 1), (2, 1), (3, 0), (0, 0), (3, 3)]
    hull = o.convex_hull(pts)
    assert sorted(map(tuple, hull)) == sorted([(0, 0), (3, 0), (3, 3), (0, 3)])


def test_convex_hull_degenerate(o):  # :69-107
    hull = o.convex_hull([(1.45, -5.2)] * 10)
    assert len(hull) == 1 and np.allclose(hull[0], (1.45, -5.2), atol=1e-9)
    hull = o.convex_hull([(1.45, -5.2), (-3.14, -1.73)] * 5)
    assert sorted(map(tuple, hull)) == sorted([(1.45, -5.2), (-3.14, -1.73)])


def test_point_in_polygon(o):  # :113-148
    sq = [(0, 0), (3, 0), (3, 3), (0, 3)]
    assert o.point_in_convex_polygon(1, 2, sq)
    assert not o.point_in_convex_polygon(-1, 2, sq)
    assert o.point_in_convex_polygon(1.5, 1.5, sq)
    assert o.point_in_convex_polygon(3, 3, sq)
    assert o.point_in_convex_polygon(0, 1.5, sq)  # on an edge -> inside
    assert not o.point_in_convex_polygon(1, 1, sq[:2])  # < 3 vertices -> outside


def test_line_from_point_and_angle(o):  # :154-178
    l = o.line_from_point_angle(0.0, 0.0, 0.0)
    assert abs(l @ np.array([1.0, 0.0, 1.0])) < 1e-9
    assert abs(abs(l @ np.array([-1.0, 1.0, 1.0])) - 1.0) < 1e-9
    l2 = o.line_from_point_angle(1.0, 1.0, math.pi / 2)
    x = np.cross(l, l2)
    assert abs(x[2]) > 1e-9
    x /= x[2]
    assert abs(x[0] - 1.0) < 1e-9 and abs(x[1]) < 1e-9


def test_degrees_radians(o):  # :180-188
    assert abs(o.deg2rad(90.0) - math.pi / 2) < 1e-9
    assert abs(o.rad2deg(math.pi / 2) - 90.0) < 1e-9
    assert abs(o.rad2deg(o.deg2rad(90.0)) - 90.0) < 1e-9


def test_clip_angle(o):  # :190-203
    for a, e in [(-math.pi / 2, 1.5 * math.pi), (0.7861, 0.7861), (0.0, 0.0), (2.7 * math.pi, 0.7 * math.pi),
                 (-math.pi, math.pi)]:
        assert abs(o.clip_angle(a) - e) < 1e-9


def test_min_angle_diff(o):  # :205-217
    for a, b, e in [(0.0, 0.0, 0.0), (-math.pi / 2, 0.0, math.pi / 2), (-3.48 * math.pi, 7.41 * math.pi, 0.89 * math.pi),
                    (0.25 * math.pi, 1.25 * math.pi, math.pi)]:
        assert abs(o.min_angle_diff(a, b) - e) < 1e-9


def test_lines_angles_diff(o):  # :219-233
    for a, b, e in [(0.0, 0.0, 0.0), (-math.pi / 2, 0.0, math.pi / 2), (0.52 * math.pi, 1.49 * math.pi, 0.03 * math.pi),
                    (1.49 * math.pi, 0.52 * math.pi, 0.03 * math.pi), (-3.48 * math.pi, 5.60 * math.pi, 0.08 * math.pi),
                    (0.25 * math.pi, 1.25 * math.pi, 0.0)]:
        assert abs(o.lines_angles_diff(a, b) - e) < 1e-9


def test_n_choose_2(o):  # :239-243
    assert (o.nchoose2(6), o.nchoose2(1), o.nchoose2(0)) == (15, 0, 0)


def test_warp_rectify_consistency(o):  # :249-282
    h7, h8 = 0.0001, 0.0002
    udx, udy, uds, udt = 82.4, -12.3, 1.13, 0.56
    ds = o.model_map(1, h7, h8, udx, udy, uds)
    dt = o.model_map(3, h7, h8, udx, udy, udt)
    dx, dy = o.model_point(1, h7, h8, udx, udy)
    assert abs(o.model_map(0, h7, h8, dx, dy, ds) - uds) < 1e-12
    assert abs(o.model_map(2, h7, h8, dx, dy, dt) - udt) < 1e-12
    x, y = o.model_point(0, h7, h8, dx, dy)
    assert abs(x - udx) < 1e-12 and abs(y - udy) < 1e-12


def test_gauss3_matches_numpy(o):
    rng = np.random.default_rng(0)
    for _ in range(50):
        a = rng.normal(size=(3, 4))
        x = o.gauss3(a)
        assert np.allclose(a[:, :3] @ x, a[:, 3], atol=1e-9)


@pytest.mark.skipif(not have_reference(), reason="oracle/_ref not built (needs /root/reference)")
def test_port_helpers_bit_equal_reference():
    """Same inputs -> identical doubles from the restatement and from the reference's helpers."""
    p, r = Oracle("port"), Oracle("reference")
    rng = np.random.default_rng(1)
    for _ in range(300):
        a, b = rng.uniform(-30, 30, 2)
        assert p.clip_angle(a) == r.clip_angle(a)
        assert p.min_angle_diff(a, b) == r.min_angle_diff(a, b)
        assert p.lines_angles_diff(a, b) == r.lines_angles_diff(a, b)
        x, y, t = rng.uniform(-4000, 4000), rng.uniform(-3000, 3000), rng.uniform(0, 7)
        assert np.array_equal(p.line_from_point_angle(x, y, t), r.line_from_point_angle(x, y, t))
        m = rng.normal(size=(3, 4)) * rng.choice([1e-3, 1.0, 1e3], size=(3, 1))
        assert np.array_equal(p.gauss3(m), r.gauss3(m), equal_nan=True)
        h7, h8 = rng.uniform(-5e-4, 5e-4, 2)
        for w in range(4):
            assert p.model_map(w, h7, h8, x, y, t + 0.1) == r.model_map(w, h7, h8, x, y, t + 0.1)
        pts = rng.integers(0, 6, size=(4, 2)).astype(float)
        hp, hr = p.convex_hull(pts), r.convex_hull(pts)
        assert np.array_equal(hp, hr)
        q = rng.integers(0, 6, 2).astype(float)
        assert p.point_in_convex_polygon(q[0], q[1], hp) == r.point_in_convex_polygon(q[0], q[1], hr)
        pp = rng.uniform(-10, 10, size=(3, 2))
        assert p.are_collinear(pp[0], pp[1], pp[2], 1.0) == r.are_collinear(pp[0], pp[1], pp[2], 1.0)
