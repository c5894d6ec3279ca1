"""Sub-terrain generators for the int16 height field (``utils/terrain.py:109-145`` calls these
through ``isaacgym.terrain_utils``, which is closed and not vendored in the reference).

Written from the documented behaviour of those generators (pyramid slope, uniform noise on a
coarse grid, concentric stairs, random boxes, stepping stones).  Heights are integer multiples of
``vertical_scale``; ``height_field_raw`` is ``int16 [width, length]`` exactly as the reference's
``Terrain.add_terrain_to_map`` expects (``utils/terrain.py:147-164``).
"""
from __future__ import annotations

import numpy as np


class SubTerrain:
    def __init__(self, terrain_name="terrain", width=256, length=256, vertical_scale=1.0, horizontal_scale=1.0):
        self.terrain_name = terrain_name
        self.vertical_scale = vertical_scale
        self.horizontal_scale = horizontal_scale
        self.width = width
        self.length = length
        self.height_field_raw = np.zeros((self.width, self.length), dtype=np.int16)


def _flatten_platform(terrain, platform_size):
    half = int(platform_size / terrain.horizontal_scale / 2)
    x1, x2 = terrain.width // 2 - half, terrain.width // 2 + half
    y1, y2 = terrain.length // 2 - half, terrain.length // 2 + half
    return x1, x2, y1, y2


def pyramid_sloped_terrain(terrain, slope=1.0, platform_size=1.0):
    cx, cy = terrain.width // 2, terrain.length // 2
    x = np.arange(terrain.width)
    y = np.arange(terrain.length)
    xx = ((cx - np.abs(cx - x)) / cx).reshape(-1, 1)
    yy = ((cy - np.abs(cy - y)) / cy).reshape(1, -1)
    peak = int(slope * (terrain.horizontal_scale / terrain.vertical_scale) * (terrain.width / 2))
    terrain.height_field_raw += (peak * xx * yy).astype(terrain.height_field_raw.dtype)
    x1, x2, y1, y2 = _flatten_platform(terrain, platform_size)
    ref = int(terrain.height_field_raw[x1, y1])
    lo, hi = min(ref, 0), max(ref, 0)
    terrain.height_field_raw = np.clip(terrain.height_field_raw, lo, hi)
    return terrain


def random_uniform_terrain(terrain, min_height, max_height, step=1.0, downsampled_scale=None):
    if downsampled_scale is None:
        downsampled_scale = terrain.horizontal_scale
    lo = int(min_height / terrain.vertical_scale)
    hi = int(max_height / terrain.vertical_scale)
    st = max(int(step / terrain.vertical_scale), 1)
    levels = np.arange(lo, hi + st, st)
    nx = int(terrain.width * terrain.horizontal_scale / downsampled_scale)
    ny = int(terrain.length * terrain.horizontal_scale / downsampled_scale)
    coarse = np.random.choice(levels, (nx, ny)).astype(np.float64)
    # bilinear up-sampling of the coarse grid to the full resolution
    gx = np.linspace(0, nx - 1, terrain.width)
    gy = np.linspace(0, ny - 1, terrain.length)
    x0 = np.floor(gx).astype(int).clip(0, nx - 2) if nx > 1 else np.zeros_like(gx, dtype=int)
    y0 = np.floor(gy).astype(int).clip(0, ny - 2) if ny > 1 else np.zeros_like(gy, dtype=int)
    fx = (gx - x0).reshape(-1, 1)
    fy = (gy - y0).reshape(1, -1)
    x1 = np.minimum(x0 + 1, nx - 1)
    y1 = np.minimum(y0 + 1, ny - 1)
    c00 = coarse[np.ix_(x0, y0)]
    c10 = coarse[np.ix_(x1, y0)]
    c01 = coarse[np.ix_(x0, y1)]
    c11 = coarse[np.ix_(x1, y1)]
    fine = (c00 * (1 - fx) * (1 - fy) + c10 * fx * (1 - fy) + c01 * (1 - fx) * fy + c11 * fx * fy)
    terrain.height_field_raw += np.rint(fine).astype(np.int16)
    return terrain


def pyramid_stairs_terrain(terrain, step_width, step_height, platform_size=1.0):
    sw = max(int(step_width / terrain.horizontal_scale), 1)
    sh = int(step_height / terrain.vertical_scale)
    plat = int(platform_size / terrain.horizontal_scale)
    height = 0
    x0, x1, y0, y1 = 0, terrain.width, 0, terrain.length
    while (x1 - x0) > plat and (y1 - y0) > plat:
        x0 += sw
        x1 -= sw
        y0 += sw
        y1 -= sw
        height += sh
        terrain.height_field_raw[x0:x1, y0:y1] = height
    return terrain


def discrete_obstacles_terrain(terrain, max_height, min_size, max_size, num_rects, platform_size=1.0):
    mh = int(max_height / terrain.vertical_scale)
    lo = int(min_size / terrain.horizontal_scale)
    hi = int(max_size / terrain.horizontal_scale)
    plat = int(platform_size / terrain.horizontal_scale)
    heights = [-mh, -mh // 2, mh // 2, mh]
    sizes = range(lo, hi, 4)
    xs = range(0, terrain.width - hi, 4)
    ys = range(0, terrain.length - hi, 4)
    for _ in range(num_rects):
        w = np.random.choice(sizes)
        l = np.random.choice(sizes)
        i = np.random.choice(xs)
        j = np.random.choice(ys)
        terrain.height_field_raw[i:i + w, j:j + l] = np.random.choice(heights)
    x1, x2 = (terrain.width - plat) // 2, (terrain.width + plat) // 2
    y1, y2 = (terrain.length - plat) // 2, (terrain.length + plat) // 2
    terrain.height_field_raw[x1:x2, y1:y2] = 0
    return terrain


def stepping_stones_terrain(terrain, stone_size, stone_distance, max_height, platform_size=1.0, depth=-10):
    ss = max(int(stone_size / terrain.horizontal_scale), 1)
    sd = max(int(stone_distance / terrain.horizontal_scale), 1)
    mh = int(max_height / terrain.vertical_scale)
    plat = int(platform_size / terrain.horizontal_scale)
    terrain.height_field_raw[:, :] = int(depth / terrain.vertical_scale)
    hs = np.arange(-mh - 1, mh, 1)
    for i in range(0, terrain.width, ss + sd):
        for j in range(0, terrain.length, ss + sd):
            terrain.height_field_raw[i:i + ss, j:j + ss] = np.random.choice(hs)
    x1, x2 = (terrain.width - plat) // 2, (terrain.width + plat) // 2
    y1, y2 = (terrain.length - plat) // 2, (terrain.length + plat) // 2
    terrain.height_field_raw[x1:x2, y1:y2] = 0
    return terrain


def convert_heightfield_to_trimesh(height_field_raw, horizontal_scale, vertical_scale, slope_threshold=None):
    """Regular two-triangles-per-cell mesh.  The hot path never reads it; the headless backend
    only needs shapes (``legged_robot.py:633-642``).  Slope correction is not applied."""
    rows, cols = height_field_raw.shape
    xs = np.arange(rows, dtype=np.float32) * horizontal_scale
    ys = np.arange(cols, dtype=np.float32) * horizontal_scale
    xx, yy = np.meshgrid(xs, ys, indexing="ij")
    vertices = np.stack([xx.ravel(), yy.ravel(), height_field_raw.astype(np.float32).ravel() * vertical_scale], axis=1)
    idx = np.arange(rows * cols, dtype=np.uint32).reshape(rows, cols)
    a, b, c, d = idx[:-1, :-1].ravel(), idx[1:, :-1].ravel(), idx[:-1, 1:].ravel(), idx[1:, 1:].ravel()
    triangles = np.concatenate([np.stack([a, d, c], 1), np.stack([a, b, d], 1)], axis=0)
    return vertices, triangles
