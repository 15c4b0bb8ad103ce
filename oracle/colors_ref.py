"""oracle.colors_ref — CPU restatement of the reference's spawn-time colouring.

TEST INFRASTRUCTURE ONLY (same rules as the rest of oracle/): imported by tests/ as the checker of
``rainbow_kernel`` (csrc/ps_render.cuh) and of the host mirror in core.py, never by the product.

Follows /root/reference/src/colors.zig:6-47 (``hsvToSdlColor``) and /root/reference/src/core.zig:411-417
(``Spawner.assignRainbowColors``) statement by statement, one particle at a time, every operation rounded
to f32 as Zig's f32 arithmetic does.  PARITY UNPINNED by reference output (no Zig toolchain here); pinned by
the hand-computed known answers in tests/test_oracle_colors.py.
"""
from __future__ import annotations

import math

import numpy as np

_f = np.float32


def hsv_to_sdl_color(h, s, v, a: int):
    """src/colors.zig:6-47.  h in degrees; returns (r, g, b, a) as ints 0..255."""
    h, s, v = _f(h), _f(s), _f(v)
    c = _f(v * s)                                                     # :7
    t = _f(h / _f(60.0))
    t = _f(math.fmod(float(t), 2.0))                                  # @mod(h / 60.0, 2.0); operands are non-negative here
    x = _f(c * _f(_f(1.0) - _f(abs(_f(t - _f(1.0))))))                # :8
    m = _f(v - c)                                                     # :9
    zero = _f(0.0)
    if h < _f(60.0):                                                  # :15-39
        r, g, b = c, x, zero
    elif h < _f(120.0):
        r, g, b = x, c, zero
    elif h < _f(180.0):
        r, g, b = zero, c, x
    elif h < _f(240.0):
        r, g, b = zero, x, c
    elif h < _f(300.0):
        r, g, b = x, zero, c
    else:
        r, g, b = c, zero, x
    k = _f(255.0)
    return (int(_f(_f(r + m) * k)), int(_f(_f(g + m) * k)), int(_f(_f(b + m) * k)), int(a))   # @intFromFloat truncates, :41-46


def assign_rainbow_colors(count: int) -> np.ndarray:
    """src/core.zig:411-417: hue_i = f32(i) / f32(count) * 360, colour = hsvToSdlColor(hue, 1, 1, 255).  (count, 4) uint8."""
    out = np.empty((count, 4), dtype=np.uint8)
    fc = _f(count)
    for i in range(count):
        hue = _f(_f(_f(i) / fc) * _f(360.0))
        out[i] = hsv_to_sdl_color(hue, 1.0, 1.0, 255)
    return out
