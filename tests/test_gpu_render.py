"""GPU parity of the after-step kernels: rainbow colours, draw rectangles, headless frame.
Checked against numpy restatements of src/main.zig:102-107, src/gui.zig:83-86 written here, and the oracle's restatement of
src/core.zig:411-417 + src/colors.zig (oracle/colors_ref.py, pinned by the hand KATs of tests/test_oracle_colors.py)."""
import numpy as np
import pytest

from oracle import colors_ref
from particle_simulator_b200 import Simulation, scenes

pytestmark = pytest.mark.gpu


def rects_reference(pos, radius):
    x = pos[:, 0].astype(np.float32); y = pos[:, 1].astype(np.float32)     # @floatCast  main.zig:104-105
    r = np.float32(radius)
    dia = np.float32(2.0) * r                                               # gui.zig:84
    return np.stack([x - r, y - r, np.full_like(x, dia), np.full_like(x, dia)], axis=1)


def splat_reference(pos, radius, colors, w, h):
    img = np.zeros((h, w, 4), dtype=np.uint8)
    r = np.float32(radius)
    for i in range(len(pos)):                                               # index order: later particles over earlier ones
        x, y = np.float32(pos[i, 0]), np.float32(pos[i, 1])
        x0, x1 = max(0, int(np.floor(x - r))), min(w - 1, int(np.floor(x + r)))
        y0, y1 = max(0, int(np.floor(y - r))), min(h - 1, int(np.floor(y + r)))
        for yy in range(y0, y1 + 1):
            for xx in range(x0, x1 + 1):
                dx = (np.float32(xx) + np.float32(0.5)) - x; dy = (np.float32(yy) + np.float32(0.5)) - y
                if dx * dx + dy * dy <= r * r:
                    img[yy, xx] = colors[i]
    return img


@pytest.mark.parametrize("precision", ["f64", "f32"])
@pytest.mark.parametrize("collisions", [False, True])
def test_draw_rects_and_rainbow_colors(precision, collisions):
    scene = scenes.default_scene()
    pos0, vel0 = scenes.spawn(scene, seed=1)
    with Simulation(scene, precision=precision, collisions=collisions) as sim:
        sim.upload(pos0, vel0)
        sim.step(2)
        pos, _ = sim.download()
        rects = sim.export_rects()
        sim.rainbow_colors()
        cols = sim.get_colors()
    assert np.array_equal(rects.view(np.uint32), rects_reference(pos, scene.radius).view(np.uint32))
    assert np.array_equal(cols, colors_ref.assign_rainbow_colors(scene.count))


def test_rainbow_colors_odd_count_matches_the_oracle():
    scene = scenes.Scene(count=36_001, width=500.0, height=400.0)
    with Simulation(scene, collisions=False) as sim:
        sim.rainbow_colors()
        cols = sim.get_colors()
    assert np.array_equal(cols, colors_ref.assign_rainbow_colors(scene.count))


def test_headless_frame_matches_index_order_painter():
    scene = scenes.Scene(count=1500, width=180.0, height=120.0)
    pos0, vel0 = scenes.spawn(scene, seed=4)
    rng = np.random.default_rng(0)
    colors = rng.integers(0, 256, size=(scene.count, 4), dtype=np.uint8)
    with Simulation(scene, collisions=True) as sim:
        sim.upload(pos0, vel0)
        sim.step(1)
        pos, _ = sim.download()
        sim.set_colors(colors)
        img = sim.render(180, 120)
    want = splat_reference(pos, scene.radius, colors, 180, 120)
    assert np.array_equal(img, want)
    assert img[..., 3].any() and (img[..., 3] == 0).any()   # something was drawn, and some pixels stayed clear
