"""Global int16 height field + sub-terrain origins (reference ``utils/terrain.py:38-187``).

One-time, host-side numpy setup.  Its products are hot-path inputs: ``heightsamples``
``[tot_rows, tot_cols]`` int16, ``env_origins [num_rows, num_cols, 3]`` and ``env_length``.
Row index is world x, column index is world y (``legged_robot.py:841-848``).
"""
from __future__ import annotations

import numpy as np

from legged_gym_b200.sim import terrain_utils


class Terrain:
    def __init__(self, cfg, num_robots) -> None:
        self.cfg = cfg
        self.num_robots = num_robots
        self.type = cfg.mesh_type
        if self.type in ("none", "plane"):
            return
        self.env_length = cfg.terrain_length
        self.env_width = cfg.terrain_width
        self.proportions = [float(np.sum(cfg.terrain_proportions[:i + 1])) for i in range(len(cfg.terrain_proportions))]
        self.cfg.num_sub_terrains = cfg.num_rows * cfg.num_cols
        self.env_origins = np.zeros((cfg.num_rows, cfg.num_cols, 3))
        self.width_per_env_pixels = int(self.env_width / cfg.horizontal_scale)
        self.length_per_env_pixels = int(self.env_length / cfg.horizontal_scale)
        self.border = int(cfg.border_size / cfg.horizontal_scale)
        self.tot_cols = int(cfg.num_cols * self.width_per_env_pixels) + 2 * self.border
        self.tot_rows = int(cfg.num_rows * self.length_per_env_pixels) + 2 * self.border
        self.height_field_raw = np.zeros((self.tot_rows, self.tot_cols), dtype=np.int16)
        terrain_file = getattr(cfg, "terrain_file", None)
        if terrain_file:
            self._load(terrain_file)
        elif cfg.curriculum:
            self.curiculum()
        elif cfg.selected:
            self.selected_terrain()
        else:
            self.randomized_terrain()
        self.heightsamples = self.height_field_raw
        if self.type == "trimesh":
            self.vertices, self.triangles = terrain_utils.convert_heightfield_to_trimesh(
                self.height_field_raw, cfg.horizontal_scale, cfg.vertical_scale, cfg.slope_treshold)

    # ---- on-disk format: the hot path only needs the int16 field, the sub-terrain
    # origins and the three scales, so a generated (or externally authored) terrain can be reused across runs
    # and ranks instead of being regenerated from the numpy stream: one .npz, fields named as below
    FILE_VERSION = 1

    def save(self, path):
        c = self.cfg
        np.savez_compressed(path, version=np.array(self.FILE_VERSION), height_field_raw=self.height_field_raw,
                            env_origins=self.env_origins, horizontal_scale=np.array(c.horizontal_scale),
                            vertical_scale=np.array(c.vertical_scale), border_size=np.array(c.border_size),
                            terrain_length=np.array(c.terrain_length), terrain_width=np.array(c.terrain_width),
                            num_rows=np.array(c.num_rows), num_cols=np.array(c.num_cols))

    def _load(self, path):
        d = np.load(path)
        c = self.cfg
        if int(d["version"]) != self.FILE_VERSION:
            raise ValueError(f"{path}: terrain file version {int(d['version'])} != {self.FILE_VERSION}")
        for k in ("horizontal_scale", "vertical_scale", "border_size", "terrain_length", "terrain_width", "num_rows",
                  "num_cols"):
            if float(d[k]) != float(getattr(c, k)):
                raise ValueError(f"{path}: {k}={float(d[k])} does not match cfg.terrain.{k}={getattr(c, k)}")
        hf = d["height_field_raw"]
        if hf.dtype != np.int16 or hf.shape != (self.tot_rows, self.tot_cols):
            raise ValueError(f"{path}: height field {hf.dtype}{hf.shape}, expected int16{(self.tot_rows, self.tot_cols)}")
        self.height_field_raw = hf.copy()
        self.env_origins = d["env_origins"].astype(np.float64).copy()

    # difficulty grows with the row (level), the type changes with the column
    def curiculum(self):
        for j in range(self.cfg.num_cols):
            for i in range(self.cfg.num_rows):
                self.add_terrain_to_map(self.make_terrain(j / self.cfg.num_cols + 0.001, i / self.cfg.num_rows), i, j)

    def randomized_terrain(self):
        for k in range(self.cfg.num_sub_terrains):
            i, j = np.unravel_index(k, (self.cfg.num_rows, self.cfg.num_cols))
            choice = np.random.uniform(0, 1)
            difficulty = np.random.choice([0.5, 0.75, 0.9])
            self.add_terrain_to_map(self.make_terrain(choice, difficulty), i, j)

    def selected_terrain(self):
        kwargs = dict(self.cfg.terrain_kwargs)
        gen = getattr(terrain_utils, kwargs.pop('type'))
        for k in range(self.cfg.num_sub_terrains):
            i, j = np.unravel_index(k, (self.cfg.num_rows, self.cfg.num_cols))
            sub = self._blank()
            gen(sub, **kwargs)
            self.add_terrain_to_map(sub, i, j)

    def _blank(self):
        return terrain_utils.SubTerrain("terrain", width=self.width_per_env_pixels, length=self.width_per_env_pixels,
                                        vertical_scale=self.cfg.vertical_scale,
                                        horizontal_scale=self.cfg.horizontal_scale)

    def make_terrain(self, choice, difficulty):
        sub = self._blank()
        slope = difficulty * 0.4
        step_height = 0.05 + 0.18 * difficulty
        obstacle_height = 0.05 + difficulty * 0.2
        stone_size = 1.5 * (1.05 - difficulty)
        stone_distance = 0.05 if difficulty == 0 else 0.1
        p = self.proportions + [np.inf] * 3
        if choice < p[0]:
            terrain_utils.pyramid_sloped_terrain(sub, slope=-slope if choice < p[0] / 2 else slope, platform_size=3.)
        elif choice < p[1]:
            terrain_utils.pyramid_sloped_terrain(sub, slope=slope, platform_size=3.)
            terrain_utils.random_uniform_terrain(sub, min_height=-0.05, max_height=0.05, step=0.005,
                                                 downsampled_scale=0.2)
        elif choice < p[3]:
            terrain_utils.pyramid_stairs_terrain(sub, step_width=0.31,
                                                 step_height=-step_height if choice < p[2] else step_height,
                                                 platform_size=3.)
        elif choice < p[4]:
            terrain_utils.discrete_obstacles_terrain(sub, obstacle_height, 1., 2., 20, platform_size=3.)
        elif choice < p[5]:
            terrain_utils.stepping_stones_terrain(sub, stone_size=stone_size, stone_distance=stone_distance,
                                                  max_height=0., platform_size=4.)
        elif choice < p[6]:
            gap_terrain(sub, gap_size=1. * difficulty, platform_size=3.)
        else:
            pit_terrain(sub, depth=1. * difficulty, platform_size=4.)
        return sub

    def add_terrain_to_map(self, terrain, row, col):
        x0 = self.border + row * self.length_per_env_pixels
        y0 = self.border + col * self.width_per_env_pixels
        self.height_field_raw[x0:x0 + self.length_per_env_pixels, y0:y0 + self.width_per_env_pixels] = \
            terrain.height_field_raw
        hs = terrain.horizontal_scale
        x1, x2 = int((self.env_length / 2. - 1) / hs), int((self.env_length / 2. + 1) / hs)
        y1, y2 = int((self.env_width / 2. - 1) / hs), int((self.env_width / 2. + 1) / hs)
        z = np.max(terrain.height_field_raw[x1:x2, y1:y2]) * terrain.vertical_scale
        self.env_origins[row, col] = [(row + 0.5) * self.env_length, (col + 0.5) * self.env_width, z]


def gap_terrain(terrain, gap_size, platform_size=1.):
    gap = int(gap_size / terrain.horizontal_scale)
    plat = int(platform_size / terrain.horizontal_scale)
    cx, cy = terrain.length // 2, terrain.width // 2
    x1 = (terrain.length - plat) // 2
    y1 = (terrain.width - plat) // 2
    x2, y2 = x1 + gap, y1 + gap
    terrain.height_field_raw[cx - x2:cx + x2, cy - y2:cy + y2] = -1000
    terrain.height_field_raw[cx - x1:cx + x1, cy - y1:cy + y1] = 0


def pit_terrain(terrain, depth, platform_size=1.):
    d = int(depth / terrain.vertical_scale)
    half = int(platform_size / terrain.horizontal_scale / 2)
    cx, cy = terrain.length // 2, terrain.width // 2
    terrain.height_field_raw[cx - half:cx + half, cy - half:cy + half] = -d
