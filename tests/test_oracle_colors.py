"""Known answers for the oracle's restatement of colors.zig / assignRainbowColors (oracle/colors_ref.py), worked out by
hand from src/colors.zig:6-47, and the product's host mirror (core.py) checked against that oracle."""
import numpy as np

from oracle import colors_ref
from particle_simulator_b200 import core


def test_primary_and_secondary_hues_by_hand():
    # c = v*s = 1, m = 0; x = 1 - |mod(h/60, 2) - 1|
    assert colors_ref.hsv_to_sdl_color(0.0, 1.0, 1.0, 255) == (255, 0, 0, 255)        # x = 0, first branch (c, x, 0)
    assert colors_ref.hsv_to_sdl_color(60.0, 1.0, 1.0, 255) == (255, 255, 0, 255)     # mod(1,2)=1 -> x = 1, second branch (x, c, 0)
    assert colors_ref.hsv_to_sdl_color(120.0, 1.0, 1.0, 255) == (0, 255, 0, 255)      # mod(2,2)=0 -> x = 0, third branch (0, c, x)
    assert colors_ref.hsv_to_sdl_color(240.0, 1.0, 1.0, 7) == (0, 0, 255, 7)          # mod(4,2)=0 -> x = 0, fifth branch (x, 0, c)
    assert colors_ref.hsv_to_sdl_color(359.0, 1.0, 1.0, 255)[0] == 255                # last branch (c, 0, x)


def test_half_way_hues_truncate():
    # x = 1 - |0.5 - 1| = 0.5 -> 0.5 * 255 = 127.5 -> @intFromFloat truncates to 127
    assert colors_ref.hsv_to_sdl_color(30.0, 1.0, 1.0, 255) == (255, 127, 0, 255)
    assert colors_ref.hsv_to_sdl_color(90.0, 1.0, 1.0, 255) == (127, 255, 0, 255)
    assert colors_ref.hsv_to_sdl_color(270.0, 1.0, 1.0, 255) == (127, 0, 255, 255)


def test_saturation_and_value():
    # h = 210, s = 0.5, v = 0.8: c = 0.4, h/60 = 3.5, mod 2 = 1.5, x = 0.4 * 0.5 = 0.2, m = 0.4; branch h < 240: (0, x, c)
    # -> (0.4, 0.6, 0.8) * 255 = (102, 153, 204)
    assert colors_ref.hsv_to_sdl_color(210.0, 0.5, 0.8, 255) == (102, 153, 204, 255)


def test_rainbow_by_index_by_hand():
    # count = 4: hues 0, 90, 180, 270
    want = np.array([[255, 0, 0, 255], [127, 255, 0, 255], [0, 255, 255, 255], [127, 0, 255, 255]], dtype=np.uint8)
    assert np.array_equal(colors_ref.assign_rainbow_colors(4), want)


def test_host_mirror_agrees_with_the_oracle():
    for count in (1, 7, 360, 13_000):
        p = core.Particles.init(count, 2.0)
        core.Spawner.assignRainbowColors(p)
        assert np.array_equal(p.colors, colors_ref.assign_rainbow_colors(count)), count
    hs = np.linspace(0, 359.9, 500, dtype=np.float32)
    got = core.hsv_to_rgba(hs, 0.7, 0.9, 200)
    want = np.array([colors_ref.hsv_to_sdl_color(h, 0.7, 0.9, 200) for h in hs], dtype=np.uint8)
    assert np.array_equal(got, want)
