"""Host-side logic that needs no GPU: the synthetic scene generator and the core.zig API mirror."""
import numpy as np

import oracle
from particle_simulator_b200 import core, scenes


def test_default_scene_is_main_zig():
    s = scenes.default_scene()
    assert (s.count, s.radius, s.width, s.height, s.grid_size) == (13000, 2.0, 720.0, 640.0, 4.0)   # src/main.zig:25-26,63-64,74
    assert (s.target_fps, s.sub_steps, s.g, s.boundary_damping, s.particle_damping) == (60.0, 8, 50.0, 0.1, 0.9)


def test_scaled_scenes_follow_survey_8d():
    assert scenes.scaled_scene(1 << 20).width == 6144.0
    assert scenes.scaled_scene(16 << 20).width == 24576.0
    assert scenes.scaled_scene(64 << 20).width == 49152.0
    assert scenes.scaled_scene(256 << 20).width == 98304.0


def test_spawn_is_counter_based_and_in_range():
    s = scenes.default_scene()
    pos, vel = scenes.spawn(s, seed=1)
    assert pos.shape == (13000, 2) and pos.dtype == np.float64
    assert pos[:, 0].min() >= 2.0 and pos[:, 0].max() < 720.0     # x ~ U[r, W)  (src/core.zig:404)
    assert pos[:, 1].min() >= 2.0 and pos[:, 1].max() < 640.0
    assert vel.min() >= -640.0 and vel.max() < 640.0
    # any sub-range regenerates identically (what lets ranks spawn their own shard)
    p2, v2 = scenes.spawn(s, seed=1, start=5000, count=100)
    assert np.array_equal(p2, pos[5000:5100]) and np.array_equal(v2, vel[5000:5100])
    p3, _ = scenes.spawn(s, seed=2)
    assert not np.array_equal(p3, pos)
    # known first values (frozen so that C++/CUDA generators can be checked against them)
    assert pos[0].tolist() == [266.3600722378987, 585.1463313112532]
    assert vel[0].tolist() == [304.2550434433497, 34.19903433193963]


def test_grid_shell_matches_oracle_shape():
    g = core.SpatialHashGrid.init(720, 640, 4.0, 13000)
    assert (g.rows, g.cols) == oracle.grid_shape(oracle.make_params())
    g = core.SpatialHashGrid(721.9, 643.9, 4.7, 10)
    assert (g.rows, g.cols) == (160, 180)


def test_solver_mirror_derives_dt_like_the_reference():
    s = core.Solver.init(60.0, 8, 50.0, 0.1, 0.9)
    assert s.dt == (1.0 / 60.0) / 8.0 and s.t_acc is False
    s.toggle_acc()
    assert s.t_acc is True


def test_particles_init_zero_fills():
    p = core.Particles.init(10, 2.0)
    assert not p.positions.any() and not p.velocities.any() and not p.accelerations.any()
    assert p.positions.shape == (10, 2) and p.colors.shape == (10, 4)


def test_spawner_fills_host_mirrors_and_rainbow_colors():
    p = core.Particles.init(360, 2.0)
    core.Spawner.spawnRandom(p, 7, 720, 640, -640.0, 640.0)
    assert p._dirty and p.positions.any()
    # hsvToSdlColor(hue, 1, 1, 255): hue 0 -> red, 120 -> green, 240 -> blue (src/colors.zig:15-39)
    assert p.colors[0].tolist() == [255, 0, 0, 255]
    assert p.colors[120].tolist() == [0, 255, 0, 255]
    assert p.colors[240].tolist() == [0, 0, 255, 255]
    assert p.colors[60].tolist() == [255, 255, 0, 255]
