"""Global int16 height field + sub-terrain origins (reference ``utils/terrain.py:38-187``).

One-time, host-side numpy setup.  Its products are hot-path inputs: ``heightsamples``
``[tot_rows, tot_cols]`` int16, ``env_origins [num_rows, num_cols, 3]`` and ``env_length``.
Row index is world x, column index is world y (``legged_robot.py:841-848``).
"""
from __future__ import annotations

import numpy as np

from legged_gym_b200.sim import terrain_utils


class Terrain:
    def __init__(self, cfg, num_robots, device=None) -> None:
        """device: a CUDA device string turns on height-field generation on the device when
        cfg.device_generation is set; the np.random stream stays on the host."""
        self.cfg = cfg
        self.num_robots = num_robots
        self.type = cfg.mesh_type
        self.height_field_dev = None
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
        on_device = (getattr(cfg, "device_generation", False) and device is not None and str(device).startswith("cuda")
                     and not terrain_file and (cfg.curriculum or not cfg.selected))
        if terrain_file:
            self._load(terrain_file)
        elif on_device:
            self._generate_on_device(device)
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

    # ---- generation on the device -------------------------------------------------------------------------------
    def _recipe(self, choice, difficulty, row, col, draws):
        """make_terrain (utils/terrain.py:109-145) and the generators behind it as ONE recipe for
        lgk_generate_terrain: same parameter arithmetic, same np.random calls in the same order (the resolved
        draws are appended to `draws`), no cells touched."""
        from legged_gym_b200 import _lgk
        c = self.cfg
        hs, vs = c.horizontal_scale, c.vertical_scale
        W = L = self.width_per_env_pixels          # the reference's sub-terrains are width x width (:111-115)
        t = _lgk.LgkSubTerrain()
        t.row, t.col, t.draws_off = row, col, len(draws)
        slope = difficulty * 0.4
        step_height = 0.05 + 0.18 * difficulty
        obstacle_height = 0.05 + difficulty * 0.2
        stone_size = 1.5 * (1.05 - difficulty)
        stone_distance = 0.05 if difficulty == 0 else 0.1
        p = self.proportions + [np.inf] * 3

        def platform(size):
            plat = int(size / hs)
            return (W - plat) // 2, (W + plat) // 2, (L - plat) // 2, (L + plat) // 2

        def sloped(sl):
            cx, cy = W // 2, L // 2
            peak = int(sl * (hs / vs) * (W / 2))
            half = int(3. / hs / 2)
            x1, y1 = W // 2 - half, L // 2 - half
            xx, yy = (cx - abs(cx - x1)) / cx, (cy - abs(cy - y1)) / cy
            ref = int(np.int16(peak * xx * yy))
            t.p[0], t.p[1], t.p[2] = peak, min(ref, 0), max(ref, 0)
        if choice < p[0]:
            t.type = _lgk.TER_SLOPE
            sloped(-slope if choice < p[0] / 2 else slope)
        elif choice < p[1]:
            t.type = _lgk.TER_SLOPE_NOISE
            sloped(slope)
            lo, hi, st = int(-0.05 / vs), int(0.05 / vs), max(int(0.005 / vs), 1)
            levels = np.arange(lo, hi + st, st)
            nx, ny = int(W * hs / 0.2), int(L * hs / 0.2)
            t.p[3], t.p[4] = nx, ny
            draws.extend(int(v) for v in np.random.choice(levels, (nx, ny)).ravel())
        elif choice < p[3]:
            t.type = _lgk.TER_STAIRS
            sw = max(int(0.31 / hs), 1)
            sh = int((-step_height if choice < p[2] else step_height) / vs)
            plat = int(3. / hs)
            k, x0, x1, y0, y1 = 0, 0, W, 0, L
            while (x1 - x0) > plat and (y1 - y0) > plat:
                x0 += sw; x1 -= sw; y0 += sw; y1 -= sw; k += 1
            t.p[0], t.p[1], t.p[2] = sw, sh, k
        elif choice < p[4]:
            t.type = _lgk.TER_OBSTACLES
            mh = int(obstacle_height / vs)
            lo, hi = int(1. / hs), int(2. / hs)
            heights = [-mh, -mh // 2, mh // 2, mh]
            sizes, xs, ys = range(lo, hi, 4), range(0, W - hi, 4), range(0, L - hi, 4)
            t.p[0] = 20
            t.p[1], t.p[2], t.p[3], t.p[4] = platform(3.)
            for _ in range(20):
                w = np.random.choice(sizes); l = np.random.choice(sizes)
                i = np.random.choice(xs); j = np.random.choice(ys)
                draws.extend((int(i), int(j), int(w), int(l), int(np.random.choice(heights))))
        elif choice < p[5]:
            t.type = _lgk.TER_STONES
            ss, sd = max(int(stone_size / hs), 1), max(int(stone_distance / hs), 1)
            hsr = np.arange(-0 - 1, 0, 1)
            ni, nj = len(range(0, W, ss + sd)), len(range(0, L, ss + sd))
            t.p[0], t.p[1], t.p[2], t.p[3] = ss, sd, int(-10 / vs), nj
            t.p[4], t.p[5], t.p[6], t.p[7] = platform(4.)
            for _ in range(ni * nj):
                draws.append(int(np.random.choice(hsr)))
        elif choice < p[6]:
            t.type = _lgk.TER_GAP
            gap, plat = int(1. * difficulty / hs), int(3. / hs)
            x1, y1 = (L - plat) // 2, (W - plat) // 2
            assert L // 2 - (x1 + gap) >= 0, "gap wider than the sub-terrain"
            t.p[0], t.p[1], t.p[2], t.p[3] = x1 + gap, y1 + gap, x1, y1
        else:
            t.type = _lgk.TER_PIT
            t.p[0], t.p[1] = int(1. * difficulty / vs), int(4. / hs / 2)
        return t

    def _generate_on_device(self, device):
        """The same field as curiculum() / randomized_terrain() below, bit for bit, written by lgk_generate_terrain."""
        import ctypes as C
        import torch
        from legged_gym_b200 import _lgk
        c = self.cfg
        assert self.length_per_env_pixels == self.width_per_env_pixels, "the reference's generators build squares"
        subs, draws = [], []
        if c.curriculum:
            for j in range(c.num_cols):
                for i in range(c.num_rows):
                    subs.append(self._recipe(j / c.num_cols + 0.001, i / c.num_rows, i, j, draws))
        else:
            for k in range(c.num_sub_terrains):
                i, j = np.unravel_index(k, (c.num_rows, c.num_cols))
                choice = np.random.uniform(0, 1)
                difficulty = np.random.choice([0.5, 0.75, 0.9])
                subs.append(self._recipe(choice, difficulty, int(i), int(j), draws))
        W = self.width_per_env_pixels
        nx = int(W * c.horizontal_scale / 0.2)
        lin = np.concatenate([np.linspace(0, nx - 1, W), np.linspace(0, nx - 1, W)])
        g = _lgk.LgkTerrainGrid()
        g.num_rows, g.num_cols = c.num_rows, c.num_cols
        g.sub_len, g.sub_wid, g.gen_w, g.gen_l = self.length_per_env_pixels, W, W, W
        g.border, g.tot_rows, g.tot_cols = self.border, self.tot_rows, self.tot_cols
        hs = c.horizontal_scale
        g.patch_x1, g.patch_x2 = int((self.env_length / 2. - 1) / hs), int((self.env_length / 2. + 1) / hs)
        g.patch_y1, g.patch_y2 = int((self.env_width / 2. - 1) / hs), int((self.env_width / 2. + 1) / hs)
        g.vertical_scale = c.vertical_scale
        arr = (_lgk.LgkSubTerrain * len(subs))(*subs)
        subs_dev = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8).to(device)
        draws_dev = torch.tensor(draws if draws else [0], dtype=torch.int32, device=device)
        lin_dev = torch.from_numpy(lin).to(device)
        field = torch.empty(self.tot_rows, self.tot_cols, dtype=torch.int16, device=device)
        origin_z = torch.zeros(c.num_rows * c.num_cols, dtype=torch.float64, device=device)
        lib = _lgk.load()
        _lgk.check("lgk_generate_terrain", lib.lgk_generate_terrain(
            C.byref(g), _lgk.ptr(subs_dev), len(subs), _lgk.ptr(draws_dev), _lgk.ptr(lin_dev), _lgk.ptr(field),
            _lgk.ptr(origin_z), _lgk.current_stream(device)))
        self.height_field_dev = field                       # stays on the device for the kernels
        self.height_field_raw = field.cpu().numpy()         # the simulator's mesh is built on the host
        z = origin_z.cpu().numpy().reshape(c.num_rows, c.num_cols)
        for i in range(c.num_rows):
            for j in range(c.num_cols):
                self.env_origins[i, j] = [(i + 0.5) * self.env_length, (j + 0.5) * self.env_width, z[i, j]]

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
