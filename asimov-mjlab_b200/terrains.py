"""Procedural box terrain of the velocity-rough tasks, restated against this package's `Spec`.

Follows the reference's generator for the sub-terrains that are ACTIVE in its default configuration
(src/mjlab/terrains/config.py:7-27; the heightfield ones are commented out there):
  * TerrainGenerator curriculum layout     src/mjlab/terrains/terrain_generator.py:166-196,198-213
  * BoxFlatTerrainCfg                      src/mjlab/terrains/primitive_terrains.py:53-64, utils.py:11-33
  * BoxPyramidStairsTerrainCfg             src/mjlab/terrains/primitive_terrains.py:66-222
  * BoxInvertedPyramidStairsTerrainCfg     src/mjlab/terrains/primitive_terrains.py:225-373
  * make_border                            src/mjlab/terrains/utils.py:36-107
  * outer border                           src/mjlab/terrains/terrain_generator.py:250-275
  * HfPyramidSlopedTerrainCfg              src/mjlab/terrains/heightfield_terrains.py:104-252
  * HfRandomUniformTerrainCfg              src/mjlab/terrains/heightfield_terrains.py:255-391
  * HfWaveTerrainCfg                       src/mjlab/terrains/heightfield_terrains.py:394-499
    (the height-field ones are commented out of ROUGH_TERRAINS_CFG "due to compilation issues in mujoco-warp",
    config.py:28-53; ROUGH_TERRAINS_HF_CFG below is that commented configuration, enabled)
Colours and lights are not physics and are left out.  Geoms are named terrain_<k> in creation order
(terrain_generator.py:110-113), all on one static body "terrain".
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np


@dataclass
class StairsCfg:
  proportion: float
  step_height_range: tuple = (0.0, 0.1)
  step_width: float = 0.3
  platform_width: float = 3.0
  border_width: float = 1.0
  inverted: bool = False


@dataclass
class FlatCfg:
  proportion: float


@dataclass
class HfPyramidSlopeCfg:
  proportion: float
  slope_range: tuple = (0.0, 1.0)
  platform_width: float = 2.0
  border_width: float = 0.25
  inverted: bool = False
  horizontal_scale: float = 0.1
  vertical_scale: float = 0.005
  base_thickness_ratio: float = 1.0


@dataclass
class HfRandomUniformCfg:
  proportion: float
  noise_range: tuple = (0.02, 0.10)
  noise_step: float = 0.02
  downsampled_scale: float | None = None
  border_width: float = 0.25
  horizontal_scale: float = 0.1
  vertical_scale: float = 0.005
  base_thickness_ratio: float = 1.0


@dataclass
class HfWaveCfg:
  proportion: float
  amplitude_range: tuple = (0.0, 0.2)
  num_waves: float = 4
  border_width: float = 0.25
  horizontal_scale: float = 0.1
  vertical_scale: float = 0.005
  base_thickness_ratio: float = 0.25


@dataclass
class RoughTerrainCfg:
  """ROUGH_TERRAINS_CFG (terrains/config.py:7-27) with curriculum=True (velocity_env_cfg.py:105-107)."""
  size: tuple = (8.0, 8.0)
  border_width: float = 20.0
  border_height: float = 1.0
  num_rows: int = 10
  num_cols: int = 20
  difficulty_range: tuple = (0.0, 1.0)
  sub_terrains: tuple = field(default_factory=lambda: (
    FlatCfg(0.4), StairsCfg(0.3), StairsCfg(0.3, inverted=True)))


def _box(out, size, pos):
  out.append((np.asarray(size, float) / 2.0, np.asarray(pos, float)))


def _make_border(out, size, inner, height, position):
  tx, ty = (size[0] - inner[0]) / 2.0, (size[1] - inner[1]) / 2.0
  _box(out, (size[0], ty, height), (position[0], position[1] + inner[1] / 2.0 + ty / 2.0, position[2]))
  _box(out, (size[0], ty, height), (position[0], position[1] - inner[1] / 2.0 - ty / 2.0, position[2]))
  _box(out, (tx, inner[1], height), (position[0] - inner[0] / 2.0 - tx / 2.0, position[1], position[2]))
  _box(out, (tx, inner[1], height), (position[0] + inner[0] / 2.0 + tx / 2.0, position[1], position[2]))


def _flat(size):
  out = []
  _box(out, (size[0], size[1], 1.0), (size[0] / 2.0, size[1] / 2.0, -0.5))
  return out, np.array([size[0] / 2, size[1] / 2, 0.0])


def _stairs(c: StairsCfg, size, difficulty):
  out = []
  sh = c.step_height_range[0] + difficulty * (c.step_height_range[1] - c.step_height_range[0])
  nx = (size[0] - 2 * c.border_width - c.platform_width) // (2 * c.step_width) + 1
  ny = (size[1] - 2 * c.border_width - c.platform_width) // (2 * c.step_width) + 1
  n = int(min(nx, ny))
  total = (n + 1) * sh
  if c.border_width > 0.0:
    _make_border(out, size, (size[0] - 2 * c.border_width, size[1] - 2 * c.border_width), sh,
                 (0.5 * size[0], 0.5 * size[1], -sh / 2))
  cx, cy = 0.5 * size[0], 0.5 * size[1]
  ts = (size[0] - 2 * c.border_width, size[1] - 2 * c.border_width)
  for k in range(n):
    bs = (ts[0] - 2 * k * c.step_width, ts[1] - 2 * k * c.step_width)
    if c.inverted:
      z = -total / 2 - (k + 1) * sh / 2.0
      h = total - (k + 1) * sh
    else:
      z = k * sh / 2.0
      h = (k + 2) * sh
    off = (k + 0.5) * c.step_width
    _box(out, (bs[0], c.step_width, h), (cx, cy + ts[1] / 2.0 - off, z))
    _box(out, (bs[0], c.step_width, h), (cx, cy - ts[1] / 2.0 + off, z))
    _box(out, (c.step_width, bs[1] - 2 * c.step_width, h), (cx + ts[0] / 2.0 - off, cy, z))
    _box(out, (c.step_width, bs[1] - 2 * c.step_width, h), (cx - ts[0] / 2.0 + off, cy, z))
  if c.inverted:
    _box(out, (ts[0] - 2 * n * c.step_width, ts[1] - 2 * n * c.step_width, sh), (cx, cy, -total - sh / 2))
    origin = np.array([cx, cy, -(n + 1) * sh])
  else:
    _box(out, (ts[0] - 2 * n * c.step_width, ts[1] - 2 * n * c.step_width, (n + 2) * sh), (cx, cy, n * sh / 2))
    origin = np.array([cx, cy, (n + 1) * sh])
  return out, origin


def rough_terrains_hf_cfg() -> RoughTerrainCfg:
  """ROUGH_TERRAINS_CFG with its commented-out height-field sub-terrains enabled (terrains/config.py:14-53)."""
  return RoughTerrainCfg(sub_terrains=(
    FlatCfg(0.2), StairsCfg(0.2), StairsCfg(0.2, inverted=True),
    HfPyramidSlopeCfg(0.1), HfPyramidSlopeCfg(0.1, inverted=True), HfRandomUniformCfg(0.2), HfWaveCfg(0.2)))


def _hf_finish(noise, c, size):
  """Common tail of the reference's hfield sub-terrains: int16 elevation -> normalised data + hfield size."""
  emin, emax = int(noise.min()), int(noise.max())
  erange = emax - emin if emax != emin else 1
  max_h = erange * c.vertical_scale
  base = max_h * c.base_thickness_ratio
  norm = (noise - emin) / erange
  return norm, [size[0] / 2, size[1] / 2, max_h, base], max_h, base


def _hf_inner(c, size):
  bp = int(c.border_width / c.horizontal_scale)
  wp, lp = int(size[0] / c.horizontal_scale), int(size[1] / c.horizontal_scale)
  return bp, wp, lp, wp - 2 * bp, lp - 2 * bp


def _hf_pyramid(c: HfPyramidSlopeCfg, size, difficulty, rng):
  slope = c.slope_range[0] + difficulty * (c.slope_range[1] - c.slope_range[0])
  if c.inverted:
    slope = -slope
  bp, wp, lp, iw, il = _hf_inner(c, size)
  noise = np.zeros((wp, lp), dtype=np.int16)
  w, l = (iw, il) if bp > 0 else (wp, lp)
  span = iw * c.horizontal_scale if bp > 0 else size[0]
  height_max = int(slope * span / 2 / c.vertical_scale)
  cx, cy = int(w / 2), int(l / 2)
  xx = ((cx - np.abs(cx - np.arange(w))) / cx).reshape(w, 1)
  yy = ((cy - np.abs(cy - np.arange(l))) / cy).reshape(1, l)
  raw = height_max * xx * yy
  pf = int(c.platform_width / c.horizontal_scale / 2)
  xp, yp = w // 2 - pf, l // 2 - pf
  zp = raw[xp, yp] if xp >= 0 and yp >= 0 else 0
  raw = np.clip(raw, min(0, zp), max(0, zp))
  if bp > 0:
    noise[bp:-bp, bp:-bp] = np.rint(raw).astype(np.int16)
  else:
    noise = np.rint(raw).astype(np.int16)
  norm, hsize, max_h, base = _hf_finish(noise, c, size)
  zoff = -max_h if c.inverted else 0.0
  spawn = zoff - base if c.inverted else max_h - base
  return norm, hsize, [size[0] / 2, size[1] / 2, zoff], np.array([size[0] / 2, size[1] / 2, spawn])


def _hf_random(c: HfRandomUniformCfg, size, difficulty, rng):
  from scipy import interpolate
  ds = c.horizontal_scale if c.downsampled_scale is None else c.downsampled_scale
  bp, wp, lp, iw, il = _hf_inner(c, size)
  noise = np.zeros((wp, lp), dtype=np.int16)
  inner = (iw * c.horizontal_scale, il * c.horizontal_scale) if bp > 0 else (size[0], size[1])
  w, l = (iw, il) if bp > 0 else (wp, lp)
  wd, ld = int(inner[0] / ds), int(inner[1] / ds)
  hmin, hmax, hstep = int(c.noise_range[0] / c.vertical_scale), int(c.noise_range[1] / c.vertical_scale), int(c.noise_step / c.vertical_scale)
  coarse = rng.choice(np.arange(hmin, hmax + hstep, hstep), size=(wd, ld))
  f = interpolate.RectBivariateSpline(np.linspace(0, inner[0], wd), np.linspace(0, inner[1], ld), coarse)
  z = f(np.linspace(0, inner[0], w), np.linspace(0, inner[1], l))
  if bp > 0:
    noise[bp:-bp, bp:-bp] = np.rint(z).astype(np.int16)
  else:
    noise = np.rint(z).astype(np.int16)
  norm, hsize, max_h, base = _hf_finish(noise, c, size)
  spawn = (c.noise_range[0] + c.noise_range[1]) / 2
  return norm, hsize, [size[0] / 2, size[1] / 2, 0.0], np.array([size[0] / 2, size[1] / 2, spawn])


def _hf_wave(c: HfWaveCfg, size, difficulty, rng):
  amp = c.amplitude_range[0] + difficulty * (c.amplitude_range[1] - c.amplitude_range[0])
  bp, wp, lp, iw, il = _hf_inner(c, size)
  noise = np.zeros((wp, lp), dtype=np.int16)
  w, l = (iw, il) if bp > 0 else (wp, lp)
  ap = int(0.5 * amp / c.vertical_scale)
  k = 2 * np.pi / (l / c.num_waves)
  raw = ap * (np.cos(np.arange(l).reshape(1, l) * k) + np.sin(np.arange(w).reshape(w, 1) * k))
  if bp > 0:
    noise[bp:-bp, bp:-bp] = np.rint(raw).astype(np.int16)
  else:
    noise = np.rint(raw).astype(np.int16)
  norm, hsize, max_h, base = _hf_finish(noise, c, size)
  return norm, hsize, [size[0] / 2, size[1] / 2, -max_h / 2], np.array([size[0] / 2, size[1] / 2, 0.0])


_HF = {HfPyramidSlopeCfg: _hf_pyramid, HfRandomUniformCfg: _hf_random, HfWaveCfg: _hf_wave}


def generate(cfg: RoughTerrainCfg, seed: int):
  """-> (geoms in creation order: (half_size[3], pos[3]) boxes or ("hfield", data[nrow, ncol] in [0, 1], size[4], pos[3]),
  terrain_origins[num_rows, num_cols, 3])."""
  rng = np.random.default_rng(seed)
  props = np.array([s.proportion for s in cfg.sub_terrains], float)
  props /= props.sum()
  sub_idx = [int(np.min(np.where(i / cfg.num_cols + 0.001 < np.cumsum(props))[0])) for i in range(cfg.num_cols)]
  boxes, origins = [], np.zeros((cfg.num_rows, cfg.num_cols, 3))
  for col in range(cfg.num_cols):
    for row in range(cfg.num_rows):
      lo, hi = cfg.difficulty_range
      d = lo + (hi - lo) * ((row + rng.uniform()) / cfg.num_rows)
      wp = np.array([-cfg.num_rows * cfg.size[0] * 0.5 + row * cfg.size[0], -cfg.num_cols * cfg.size[1] * 0.5 + col * cfg.size[1], 0.0])
      sc = cfg.sub_terrains[sub_idx[col]]
      if type(sc) in _HF:
        norm, hsize, gpos, org = _HF[type(sc)](sc, cfg.size, d, rng)
        boxes.append(("hfield", norm, hsize, np.asarray(gpos, float) + wp))      # kept in creation order with the boxes
      else:
        tile, org = _flat(cfg.size) if isinstance(sc, FlatCfg) else _stairs(sc, cfg.size, d)
        boxes += [(h, p + wp) for h, p in tile]
      origins[row, col] = org + wp
  bs = (cfg.num_rows * cfg.size[0] + 2 * cfg.border_width, cfg.num_cols * cfg.size[1] + 2 * cfg.border_width)
  _make_border(boxes, bs, (cfg.num_rows * cfg.size[0], cfg.num_cols * cfg.size[1]), abs(cfg.border_height),
               (0.0, 0.0, -cfg.border_height / 2))
  return boxes, origins


def add_rough_terrain(scene, seed=0, cfg: RoughTerrainCfg | None = None):
  """Add body "terrain" with the generated boxes to `scene` (a mjcf.Spec); returns terrain_origins."""
  cfg = cfg or RoughTerrainCfg()
  boxes, origins = generate(cfg, seed)
  body = scene.worldbody.add_body(name="terrain")
  for k, item in enumerate(boxes):
    if isinstance(item[0], str):
      _, norm, hsize, pos = item
      scene.add_hfield(f"hfield_{k}", norm, hsize)
      body.add_geom(name=f"terrain_{k}", type="hfield", hfield=f"hfield_{k}", pos=[float(x) for x in pos])
    else:
      half, pos = item
      body.add_geom(name=f"terrain_{k}", type="box", size=[float(x) for x in half], pos=[float(x) for x in pos])
  scene.terrain_origins = origins
  return origins


def env_origins_curriculum(num_envs, origins, seed=0, max_init_level=None):
  """TerrainImporter._compute_env_origins_curriculum (terrain_importer.py:203-223) with a numpy RNG for the levels."""
  nr, nc = origins.shape[:2]
  mx = nr - 1 if max_init_level is None else min(max_init_level, nr - 1)
  rng = np.random.default_rng(seed)
  levels = rng.integers(0, mx + 1, num_envs)
  types = np.floor(np.arange(num_envs) / (num_envs / nc)).astype(int)
  return origins[levels, types], levels, types


def surface_height(m, xy) -> np.ndarray:
  """Highest terrain surface under each (x, y): box tops, height-field surfaces (max of the cell's four corners), 0 for
  a plane.  Static, axis-aligned terrain geoms only (what `generate` makes).  Used to spawn robots ON the terrain:
  the generator's own origins of the height-field pyramids lie inside the field's slab
  (heightfield_terrains.py:243-246: spawn_height = max_height - base_thickness = 0 for base_thickness_ratio = 1)."""
  xy = np.atleast_2d(np.asarray(xy, float))
  z = np.full(len(xy), -np.inf)
  gt = np.asarray(m.geom_type); gp = np.asarray(m.geom_pos); gs = np.asarray(m.geom_size)
  stat = np.asarray(m.body_weldid)[np.asarray(m.geom_bodyid)] == 0
  if np.any(stat & (gt == 0)):
    z = np.maximum(z, 0.0)
  bx = np.where(stat & (gt == 6))[0]
  for k0 in range(0, len(bx), 512):          # boxes, in chunks (points x boxes masks)
    b = bx[k0:k0 + 512]
    inside = (np.abs(xy[:, None, 0] - gp[None, b, 0]) <= gs[None, b, 0]) & (np.abs(xy[:, None, 1] - gp[None, b, 1]) <= gs[None, b, 1])
    top = np.where(inside, (gp[b, 2] + gs[b, 2])[None, :], -np.inf).max(axis=1)
    z = np.maximum(z, top)
  for g in np.where(stat & (gt == 1))[0]:
    h = int(m.geom_dataid[g]); sx, sy, sz, _ = np.asarray(m.hfield_size)[h]
    nrow, ncol = int(m.hfield_nrow[h]), int(m.hfield_ncol[h])
    data = np.asarray(m.hfield_data)[int(m.hfield_adr[h]): int(m.hfield_adr[h]) + nrow * ncol].reshape(nrow, ncol)
    lx, ly = xy[:, 0] - gp[g, 0], xy[:, 1] - gp[g, 1]
    inside = (np.abs(lx) <= sx) & (np.abs(ly) <= sy)
    if not inside.any():
      continue
    c = np.clip(np.floor((lx[inside] + sx) / (2 * sx) * (ncol - 1)).astype(int), 0, ncol - 2)
    r = np.clip(np.floor((ly[inside] + sy) / (2 * sy) * (nrow - 1)).astype(int), 0, nrow - 2)
    top = np.maximum.reduce([data[r, c], data[r + 1, c], data[r, c + 1], data[r + 1, c + 1]]) * sz + gp[g, 2]
    z[inside] = np.maximum(z[inside], top)
  return np.where(np.isfinite(z), z, 0.0)
