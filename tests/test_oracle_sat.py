"""Known-answer tests of the oracle's polygon collision restatement
(PathPlanning/CollisionDetection/src/utils.jl:14-74) — derived cases; the reference records
no expected outputs (CollisionDetection/main.jl only prints)."""
import math

import numpy as np
import pytest

import oracle_lib as orc
from motionplanning_b200 import _capi as capi
from motionplanning_b200 import scenarios as sc

WALL1 = [0.0, -1.0, 0.0, 5.5 / 2 + 1.0, 1.0]  # CollisionDetection/main.jl:7, block_info[1]


def test_get_rectangle_pts_order(built):
    # utils.jl:21-24: UpLeft, LowerLeft, LowerRight, UpperRight, back to UpLeft
    pts = orc.get_rectangle_pts([1.0, 2.0, 0.0, 3.0, 0.5])
    np.testing.assert_allclose(pts, [[-2, -2, 4, 4, -2], [2.5, 1.5, 1.5, 2.5, 2.5]], atol=1e-15)
    pts = orc.get_rectangle_pts([0.0, 0.0, math.pi / 2, 2.0, 1.0])
    np.testing.assert_allclose(pts[:, 0], [-1.0, -2.0], atol=1e-15)  # (-l, +w) rotated by 90 deg


def test_sat1_wall_vs_vehicle_free(built):
    # main.jl:7-12: ConvexCollision(wall_1, veh) -> true (collision-free)
    wall = orc.get_rectangle_pts(WALL1)
    veh = orc.get_rectangle_pts([5.5, 1.0, 0, 3 / 2, 2 / 2])
    assert orc.convex_collision(wall, veh) is True


def test_sat2_diamond_inside_wall_hits(built):
    # main.jl:16-19: 4-point diamond of half-size 0.1 around (0,-1), inside the wall -> false
    c, e = np.array([0.0, -1.0]), 1e-1
    diamond = np.array([[c[0], c[1] + e], [c[0] - e, c[1]], [c[0] + e, c[1]], [c[0], c[1] + e]]).T
    wall = orc.get_rectangle_pts(WALL1)
    assert orc.convex_collision(wall, diamond) is False


def test_sat3_quirk_both_directions_required(built):
    # unit squares, second at (2.45, 0) rotated 45 deg: disjoint, an axis of the FIRST separates,
    # none of the second's does -> ConvexCollision = false (utils.jl:64-74 is stricter than SAT)
    a = orc.get_rectangle_pts([0, 0, 0, 1, 1])
    b = orc.get_rectangle_pts([2.45, 0, math.pi / 4, 1, 1])
    assert orc.sat(a, b) is True
    assert orc.sat(b, a) is False
    assert orc.convex_collision(a, b) is False


def test_sat4_overlap(built):
    a = orc.get_rectangle_pts([0, 0, 0, 1, 1])
    b = orc.get_rectangle_pts([2.3, 0, math.pi / 4, 1, 1])
    assert orc.sat(a, b) is False and orc.sat(b, a) is False
    assert orc.convex_collision(a, b) is False


def test_touching_counts_as_separated(built):
    # utils.jl:57 uses <= : shared edge is "separated"
    a = orc.get_rectangle_pts([0, 0, 0, 1, 1])
    b = orc.get_rectangle_pts([2.0, 0, 0, 1, 1])
    assert orc.convex_collision(a, b) is True


def test_mod_pi(built):
    # ReedsSheppsUtils.jl:32-46
    pi = math.pi
    assert orc.mod_pi(0.3) == 0.3
    assert orc.mod_pi(pi) == pi and orc.mod_pi(-pi) == -pi
    assert orc.mod_pi(pi + 0.5) == pytest.approx(-pi + 0.5, abs=1e-15)
    assert orc.mod_pi(-pi - 0.5) == pytest.approx(pi - 0.5, abs=1e-15)
    assert orc.mod_pi(7 * pi + 0.25) == pytest.approx(-pi + 0.25, abs=1e-14)


def _centre_radius_hits(cfg, blocks, poses):
    """numpy statement of the kernel's 2-axis centre/radius form (csrc/mppi_models.cuh block_hits)."""
    hl, hw, off = cfg.veh_L / 2, cfg.veh_W / 2, cfg.veh_center_offset
    out = np.zeros(len(poses), dtype=int)
    for i, (x, y, psi) in enumerate(poses):
        cs, sn = math.cos(psi), math.sin(psi)
        vx, vy = x + off * cs, y + off * sn
        for cx, cy, Psi, wl, ww in blocks:
            C, S = math.cos(Psi), math.sin(Psi)
            dx, dy = vx - cx, vy - cy
            ac, as_ = abs(cs * C + sn * S), abs(sn * C - cs * S)
            sepW = abs(dx * C + dy * S) >= wl + (hl * ac + hw * as_) or abs(dy * C - dx * S) >= ww + (hl * as_ + hw * ac)
            sepV = abs(dx * cs + dy * sn) >= hl + (wl * ac + ww * as_) or abs(dy * cs - dx * sn) >= hw + (wl * as_ + ww * ac)
            out[i] += 0 if (sepW and sepV) else 1
    return out


def test_centre_radius_form_equals_literal_sat(built):
    """The kernel collapses the reference's 4 edges x 10 points x 2 directions into 4 interval
    tests; it must give the same verdict as the literal restatement (outside the rounding band)."""
    cfg = capi.default_config()
    rng = np.random.RandomState(5)
    blocks = sc.PARKING_BLOCKS + sc.wall_blocks(16)
    poses = np.stack([rng.uniform(-8, 100, 20000), rng.uniform(-10, 10, 20000), rng.uniform(-7, 7, 20000)], axis=1)
    hit, margin = orc.check_poses(cfg, [], blocks, poses)
    mine = _centre_radius_hits(cfg, blocks, poses)
    ok = margin > 1e-9
    assert ok.mean() > 0.999
    assert np.array_equal(hit[ok], mine[ok])
    assert hit.max() >= 1 and (hit == 0).any()
