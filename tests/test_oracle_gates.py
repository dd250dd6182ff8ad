"""CPU checks of oracle/gate_geometry.py (the analytic stand-in for PyBullet's ray / contact queries, row A10)."""
import numpy as np

from oracle import gate_geometry as gg

I3 = np.eye(3)


def test_segment_cylinder_cases():
    r, hl = 0.06, 0.0125
    a, b = np.array([0.0, 0.0, -0.2]), np.array([0.0, 0.0, 0.2])
    assert gg.segment_hits_cylinder(a, b, r, hl)                               # through the axis
    assert not gg.segment_hits_cylinder(a + [0.07, 0, 0], b + [0.07, 0, 0], r, hl)   # just outside the radius
    assert gg.segment_hits_cylinder(a + [0.059, 0, 0], b + [0.059, 0, 0], r, hl)
    assert not gg.segment_hits_cylinder(np.array([0, 0, 0.02]), np.array([0, 0, 0.3]), r, hl)   # starts above the cap
    assert gg.segment_hits_cylinder(np.array([-1, 0, 0.0]), np.array([1, 0, 0.0]), r, hl)       # horizontal, through
    assert not gg.segment_hits_cylinder(np.array([-1, 0, 0.05]), np.array([1, 0, 0.05]), r, hl) # horizontal, above
    # the hit must start before fraction 0.9999 (quadrotor.py:1088)
    assert not gg.segment_hits_cylinder(np.array([0, 0, -1.0]), np.array([0, 0, -hl]), r, hl)


def test_gate_rays_follow_the_opening():
    gx, gy, rot, h = 0.5, -2.5, -1.57, 1.0
    assert gg.gate_rays_hit_drone(gx, gy, rot, h, [gx, gy, h], I3)
    # 0.15 m along the opening (local x = (cos rot, sin rot)): the outermost ray
    assert gg.gate_rays_hit_drone(gx, gy, rot, h, [gx + 0.15 * np.cos(rot), gy + 0.15 * np.sin(rot), h], I3)
    assert not gg.gate_rays_hit_drone(gx, gy, rot, h, [gx + 0.25 * np.cos(rot), gy + 0.25 * np.sin(rot), h], I3)
    assert not gg.gate_rays_hit_drone(gx, gy, rot, h, [gx + 0.2, gy, h], I3)       # 0.2 m before the gate plane
    assert not gg.gate_rays_hit_drone(gx, gy, rot, h, [gx, gy, h + 0.25], I3)      # above the segments
    # a drone pitched by 90 deg presents its 0.025 m thickness: the rays 0.05 m apart can miss it
    Ry = np.array([[0, 0, 1], [0, 1, 0], [-1, 0, 0]], dtype=float)
    assert gg.gate_rays_hit_drone(gx, gy, rot, h, [gx, gy, h], Ry)


def test_collisions():
    gates = [[0.5, -2.5, 1.0, 0, 0, -1.57]]
    obstacles = [[1.5, -2.5, 0.525, 0, 0, 0]]
    free = [0.5, -2.5, 1.0]
    assert not gg.drone_collides(free, I3, gates, [0], obstacles)
    c, s = np.cos(-1.57), np.sin(-1.57)
    side = [0.5 + 0.225 * c, -2.5 + 0.225 * s, 1.0]         # on the side bar
    assert gg.drone_collides(side, I3, gates, [0], obstacles)
    assert gg.drone_collides([0.5, -2.5, 1.225], I3, gates, [0], obstacles)    # top bar
    assert gg.drone_collides([0.5, -2.5, 0.5], I3, gates, [0], obstacles)      # support pole of the tall gate
    assert gg.drone_collides([0.5, -2.5, 0.3], I3, gates, [0], obstacles)      # ... which reaches down to 0.2
    assert not gg.drone_collides([0.5, -2.5, 0.3], I3, gates, [1], obstacles)  # the low type stands on a short box instead
    assert gg.drone_collides([0.5, -2.5, 0.5], I3, gates, [1], obstacles)
    assert gg.drone_collides([1.5, -2.45, 0.5], I3, gates, [0], obstacles)     # obstacle cylinder
    assert gg.drone_collides([1.5, -2.5, 0.98], I3, gates, [0], obstacles)     # just above its top: sphere still touches
    assert not gg.drone_collides([1.5, -2.5, 1.0], I3, gates, [0], obstacles)
    assert gg.drone_collides([3.0, 3.0, 0.01], I3, gates, [0], obstacles)      # ground
    assert not gg.drone_collides([3.0, 3.0, 0.03], I3, gates, [0], obstacles)
    Rx = np.array([[1, 0, 0], [0, 0, -1], [0, 1, 0]], dtype=float)             # on its side: support = radius
    assert gg.drone_collides([3.0, 3.0, 0.05], Rx, gates, [0], obstacles)
