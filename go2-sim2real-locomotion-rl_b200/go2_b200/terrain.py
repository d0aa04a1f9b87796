"""Stair terrain heightfield of the stair env (host side, runs once).

Restates `build_stair_terrain` (examples/locomotion/final/go2_env_stair.py:47-185): 13 difficulty rows along y,
each row = flat approach + `num_flights` x (stairs up, platform, stairs down, flat gap) + flat exit along x;
step height grows linearly from `step_height_min` to `step_height_max` over the rows.  Heights are int16 in
units of `vertical_scale`; cell (ix, iy) sits at origin + (ix, iy) * horizontal_scale with
origin = (0, -total_y / 2, 0).  Spawn point of a row = middle of its flat approach, centred in y.
"""
from __future__ import annotations

import numpy as np


def build_stair_terrain(terrain_cfg: dict):
    g = terrain_cfg.get
    h_scale, v_scale = g("horizontal_scale", 0.05), g("vertical_scale", 0.005)
    num_rows, row_width_m = g("num_difficulty_rows", 10), g("row_width_m", 6.0)
    step_depth_m, num_steps, num_flights = g("step_depth_m", 0.30), g("num_steps", 6), g("num_flights", 4)

    def cells(metres):
        return max(1, int(round(metres / h_scale)))

    depth = cells(step_depth_m)
    before, top = cells(g("flat_before_m", 2.0)), cells(g("flat_top_m", 1.5))
    gap, after = cells(g("flat_gap_m", 1.5)), cells(g("flat_after_m", 2.0))
    width = cells(row_width_m)
    stairs = num_steps * depth
    nx = before + num_flights * (2 * stairs + top + gap) + after
    ny = num_rows * width
    heights_m = np.linspace(g("step_height_min", 0.02), g("step_height_max", 0.15), num_rows)
    heights_u = np.round(heights_m / v_scale).astype(np.int16)
    hf = np.zeros((nx, ny), dtype=np.int16)
    centers = []
    for row in range(num_rows):
        y0, y1, sh = row * width, (row + 1) * width, heights_u[row]
        x = before
        for _ in range(num_flights):
            for s in range(num_steps):  # up
                hf[x:x + depth, y0:y1] = (s + 1) * sh
                x += depth
            hf[x:x + top, y0:y1] = num_steps * sh  # platform
            x += top
            for s in range(num_steps):  # down
                hf[x:x + depth, y0:y1] = num_steps * sh - (s + 1) * sh
                x += depth
            x += gap
        centers.append((before * h_scale * 0.5, (y0 + width / 2.0) * h_scale, 0.0))
    total_x_m, total_y_m = nx * h_scale, ny * h_scale
    origin = (0.0, -total_y_m / 2.0, 0.0)
    info = {
        "heightfield": hf, "horizontal_scale": h_scale, "vertical_scale": v_scale, "terrain_origin": origin,
        "num_difficulty_rows": num_rows, "row_centers": [(cx + origin[0], cy + origin[1], cz + origin[2]) for cx, cy, cz in centers],
        "step_heights_m": heights_m.tolist(), "total_x_m": total_x_m, "total_y_m": total_y_m,
        "num_flights": num_flights, "num_steps": num_steps, "step_depth_m": step_depth_m,
    }
    return hf, info


def build_stair_terrain_variable_depth(terrain_cfg: dict):
    """Stair terrain of the stair-info variant (examples/locomotion/go2_env_stair_lidar_2.py:83-203): like `build_stair_terrain`, but
    every difficulty row has its own tread depth - `num_rows` depths evenly spaced over `step_depth_range`, shuffled with numpy's
    GLOBAL generator (`np.random.shuffle`, so the layout follows `np.random.seed` like the reference's) - and every stair section
    is padded to the deepest row's length (ascent padded at platform height, descent at ground level), so all rows keep one length."""
    g = terrain_cfg.get
    h_scale, v_scale = g("horizontal_scale", 0.05), g("vertical_scale", 0.005)
    num_rows, num_steps, num_flights = g("num_difficulty_rows", 10), g("num_steps", 6), g("num_flights", 4)
    depth_range = g("step_depth_range", [0.25, 0.25])
    if isinstance(depth_range, (int, float)):
        depth_range = [float(depth_range)] * 2
    if g("step_depths_m") is not None:  # extension: explicit per-row depths (fixtures, evaluation on a known layout) instead of the shuffle
        depths_m = np.asarray(g("step_depths_m"), dtype=np.float64)
        if depths_m.shape != (num_rows,):
            raise ValueError("terrain['step_depths_m'] needs one depth per difficulty row")
    else:
        depths_m = np.linspace(depth_range[0], depth_range[1], num_rows)
        np.random.shuffle(depths_m)
    depth_cells = np.maximum(1, np.round(depths_m / h_scale).astype(int))

    def cells(metres):
        return max(1, int(round(metres / h_scale)))

    before, top = cells(g("flat_before_m", 2.0)), cells(g("flat_top_m", 1.5))
    gap, after = cells(g("flat_gap_m", 1.5)), cells(g("flat_after_m", 2.0))
    width = cells(g("row_width_m", 6.0))
    section = num_steps * int(depth_cells.max())  # cells reserved for one ascent / descent in every row
    nx = before + num_flights * (2 * section + top + gap) + after
    ny = num_rows * width
    heights_m = np.linspace(g("step_height_min", 0.02), g("step_height_max", 0.15), num_rows)
    heights_u = np.round(heights_m / v_scale).astype(np.int16)
    hf = np.zeros((nx, ny), dtype=np.int16)
    centers = []
    for row in range(num_rows):
        y0, y1, sh, dc = row * width, (row + 1) * width, heights_u[row], int(depth_cells[row])
        x = before
        for _ in range(num_flights):
            hf[x:x + section + top, y0:y1] = num_steps * sh  # platform incl. the padding behind a short ascent ...
            for s in range(num_steps):  # ... then the ascent over its front
                hf[x + s * dc:x + (s + 1) * dc, y0:y1] = (s + 1) * sh
            x += section + top
            for s in range(num_steps):  # descent; the padding behind it stays at ground level
                hf[x + s * dc:x + (s + 1) * dc, y0:y1] = num_steps * sh - (s + 1) * sh
            x += section + gap
        centers.append((before * h_scale * 0.5, (y0 + width / 2.0) * h_scale, 0.0))
    total_x_m, total_y_m = nx * h_scale, ny * h_scale
    origin = (0.0, -total_y_m / 2.0, 0.0)
    info = {
        "heightfield": hf, "horizontal_scale": h_scale, "vertical_scale": v_scale, "terrain_origin": origin,
        "num_difficulty_rows": num_rows, "row_centers": [(cx + origin[0], cy + origin[1], cz + origin[2]) for cx, cy, cz in centers],
        "step_heights_m": heights_m.tolist(), "step_depths_m": depths_m.tolist(), "total_x_m": total_x_m, "total_y_m": total_y_m,
        "num_flights": num_flights, "num_steps": num_steps,
    }
    return hf, info


# ---------------------------------------------------------------------------------------------------------------------------------
# Sub-terrain grid of examples/locomotion/terrain_manager.py's grid mode = Genesis' `gs.morphs.Terrain(n_subterrains=...,
# subterrain_types=...)`: genesis/utils/terrain.py:17-160 (parse_terrain: tile loop, shared borders keep the maximum) over the Isaac
# Gym tile generators of genesis/ext/isaacgym/terrain_utils.py.  Heights are float32 in units of `vertical_scale`, like Genesis'.
# Random tiles draw from numpy's GLOBAL generator in the reference's order; `randomize=False` seeds 0 around every tile and restores
# the state.  tests/test_terrain_grid.py pins the tiles against digests of the reference functions' own output.
# ---------------------------------------------------------------------------------------------------------------------------------
GS_EPS = np.finfo(np.float32).eps  # gs.EPS at precision "32"
TILE_DEFAULTS = {  # genesis/options/morphs.py:1313-1356
    "flat_terrain": {},
    "fractal_terrain": {"levels": 8, "scale": 5.0},
    "stepping_stones_terrain": {"stone_size": 1.0, "stone_distance": 0.25, "max_height": 0.2, "platform_size": 0.0},
    "random_uniform_terrain": {"min_height": -0.1, "max_height": 0.1, "step": 0.1, "downsampled_scale": 0.5},
    "sloped_terrain": {"slope": -0.5},
    "pyramid_sloped_terrain": {"slope": -0.1},
    "discrete_obstacles_terrain": {"max_height": 0.05, "min_size": 1.0, "max_size": 5.0, "num_rects": 20},
    "wave_terrain": {"num_waves": 2.0, "amplitude": 0.1},
    "stairs_terrain": {"step_width": 0.75, "step_height": -0.1},
    "pyramid_stairs_terrain": {"step_width": 0.75, "step_height": -0.1},
}


def not_a_knot_spline(x, y, xv):
    """Cubic spline through (x, y[:, k]) with not-a-knot ends, evaluated at xv (genesis/utils/geom.py:2151-2230 cubic_spline_1d): the
    second-derivative system is an almost tridiagonal matrix (one extra entry in the first and the last row) eliminated top-down in
    FLOAT32 like the reference's band storage, data and abscissae stay float64."""
    y = y[:, None] if y.ndim == 1 else y
    n = len(x)
    h = np.diff(x)
    f32 = np.float32
    lo2, lo, di, up, up2 = (np.zeros(n, f32) for _ in range(5))  # A[i, i-2], A[i, i-1], A[i, i], A[i, i+1], A[i, i+2]
    r = np.zeros(y.shape, f32)
    di[0], up[0], up2[0] = h[1], -(h[0] + h[1]), h[0]
    lo2[-1], lo[-1], di[-1] = h[-1], -(h[-2] + h[-1]), h[-2]
    for i in range(1, n - 1):
        lo[i], di[i], up[i] = h[i - 1], 2 * (h[i - 1] + h[i]), h[i]
        r[i] = 3 * ((y[i + 1] - y[i]) / h[i] - (y[i] - y[i - 1]) / h[i - 1])
    for k in range(n - 1):
        piv = di[k]
        f = lo[k + 1] / piv
        di[k + 1] -= f * up[k]
        if k + 2 < n:
            up[k + 1] -= f * up2[k]
        r[k + 1] -= f * r[k]
        if k == n - 3:  # the last row's extra entry sits in column n - 3
            f = lo2[-1] / piv
            lo[-1] -= f * up[k]
            di[-1] -= f * up2[k]
            r[-1] -= f * r[k]
    c = np.zeros_like(r)
    c[-1] = r[-1] / di[-1]
    c[-2] = (r[-2] - up[-2] * c[-1]) / di[-2]
    for i in range(n - 3, -1, -1):
        c[i] = r[i] - up[i] * c[i + 1]
        if up2[i] != 0:
            c[i] -= up2[i] * c[i + 2]
        c[i] /= di[i]
    b = (y[1:] - y[:-1]) / h[:, None] - h[:, None] * (2 * c[:-1] + c[1:]) / 3
    d = (c[1:] - c[:-1]) / (3 * h[:, None])
    xv = np.atleast_1d(xv)
    k = np.clip(np.searchsorted(x[1:], xv), 0, n - 2)
    t = (xv - x[k])[:, None]
    return y[k] + b[k] * t + c[k] * t**2 + d[k] * t**3


def _tile(kind, W, L, hs, vs, p):
    """One W x L tile (float32, units of vs) of the named Isaac Gym generator (genesis/ext/isaacgym/terrain_utils.py)."""
    z = np.zeros((W, L), dtype=np.float32)
    if kind == "flat_terrain":
        return z
    if kind == "random_uniform_terrain":  # :43-90: coarse random grid -> separable spline up-sampling -> rint
        ds = p["downsampled_scale"] if p.get("downsampled_scale") is not None else hs
        sw, sl = W * hs, L * hs
        lo, hi, st = int(p["min_height"] / vs), int(p["max_height"] / vs), int(p["step"] / vs)
        coarse = np.random.choice(np.arange(lo, hi + st, st), (int(sw / ds), int(sl / ds)))
        xc, yc = np.linspace(0, sw, coarse.shape[0]), np.linspace(0, sl, coarse.shape[1])
        along_y = not_a_knot_spline(yc, coarse.T, np.linspace(0, sl, L)).T
        z += np.rint(not_a_knot_spline(xc, along_y, np.linspace(0, sw, W)))
        return z
    if kind == "sloped_terrain":  # :93-110
        top = int(p["slope"] * (hs / vs) * W)
        z[:, np.arange(L)] += top * np.arange(W).reshape(W, 1) / W
        return z
    if kind == "pyramid_sloped_terrain":  # :113-145
        cx, cy = int(W / 2), int(L / 2)
        fx = ((cx - np.abs(cx - np.arange(W))) / cx).reshape(W, 1)
        fy = ((cy - np.abs(cy - np.arange(L))) / cy).reshape(1, L)
        z += int(p["slope"] * (hs / vs) * (W / 2)) * fx * fy
        half = int(p.get("platform_size", 1.0) / hs / 2)
        edge = z[W // 2 - half, L // 2 - half]
        return np.clip(z, min(edge, 0), max(edge, 0))
    if kind == "discrete_obstacles_terrain":  # :148-185
        top, smin, smax = int(p["max_height"] / vs), int(p["min_size"] / hs), int(p["max_size"] / hs)
        plat = int(p.get("platform_size", 1.0) / hs)
        ws = np.arange(min(smin, W - 1), min(smax, W - 1) + 1, 4)
        ls = np.arange(min(smin, L - 1), min(smax, L - 1) + 1, 4)
        hts = [-top, -top // 2, top // 2, top]
        for _ in range(int(p["num_rects"])):
            w, l, hgt = np.random.choice(ws), np.random.choice(ls), np.random.choice(hts)
            x0 = np.random.choice(range(0, W - w, 4))
            y0 = np.random.choice(range(0, L - l, 4))
            z[x0:x0 + w, y0:y0 + l] = hgt
        z[(W - plat) // 2:(W + plat) // 2, (L - plat) // 2:(L + plat) // 2] = 0
        return z
    if kind == "wave_terrain":  # :188-208
        amp = 0.5 * p["amplitude"] / vs
        if p["num_waves"] > 0:
            div = L / (p["num_waves"] * np.pi * 2)
            z += amp * np.cos(np.arange(L).reshape(1, L) / div) + amp * np.sin(np.arange(W).reshape(W, 1) / div)
        return z
    if kind == "stairs_terrain":  # :211-231
        sw, sh = int(p["step_width"] / hs), int(p["step_height"] / vs)
        for i in range(W // sw):
            z[i * sw:(i + 1) * sw, :] += (i + 1) * sh
        return z
    if kind == "pyramid_stairs_terrain":  # :234-263
        sw, sh, plat = int(p["step_width"] / hs), int(p["step_height"] / vs), int(p.get("platform_size", 1.0) / hs)
        x0, x1, y0, y1, hgt = 0, W, 0, L, 0
        while (x1 - x0) > plat and (y1 - y0) > plat:
            x0, x1, y0, y1, hgt = x0 + sw, x1 - sw, y0 + sw, y1 - sw, hgt + sh
            z[x0:x1, y0:y1] = hgt
        return z
    if kind == "fractal_terrain":  # genesis/utils/terrain.py:190-216: midpoint displacement, coarse to fine; a point takes the mean of
        # the coarser level's points around it (the two on its row / column, or the four diagonal ones) plus a halving random term
        levels, out = int(p["levels"]), np.zeros((W, L), dtype=np.float32)
        for level in range(1, levels + 1):
            st = 2 ** (levels - level)
            for r in range(0, W, st):
                odd_r = (r // st) % 2  # rows of the coarser level are the even multiples of st
                dr = st * odd_r
                for c in range(st * (1 - odd_r), L, st * (2 - odd_r)):
                    dc = st * ((c // st) % 2)
                    around = out[r - dr:r + dr + 1:2 * st, c - dc:c + dc + 1:2 * st]
                    out[r, c] = around.mean() + p["scale"] * (2.0 ** (-level) * np.random.uniform(-1, 1))
        out /= vs
        return out
    if kind == "stepping_stones_terrain":  # terrain_utils.py:266-322: stones in lanes along the longer side, holes of `depth` between
        size, gap = int(p["stone_size"] / hs), int(p["stone_distance"] / hs)
        top, plat = int(p["max_height"] / vs), int(p.get("platform_size", 1.0) / hs)
        levels = np.arange(-top - 1, top, step=1)
        z[:, :] = int(p.get("depth", -10) / vs)
        lanes_along_x = L >= W  # lanes are strips of constant y, stones laid along x
        n_lane, n_run = (L, W) if lanes_along_x else (W, L)
        a = 0
        while a < n_lane:
            a1 = min(n_lane, a + size)
            b = np.random.randint(0, size)
            runs = [(0, max(0, b - gap))]  # the piece before the first stone
            while b < n_run:
                runs.append((b, min(n_run, b + size)))
                b += size + gap
            for b0, b1 in runs:
                hgt = np.random.choice(levels)
                if lanes_along_x:
                    z[b0:b1, a:a1] = hgt
                else:
                    z[a:a1, b0:b1] = hgt
            a += size + gap
        z[(W - plat) // 2:(W + plat) // 2, (L - plat) // 2:(L + plat) // 2] = 0
        return z
    raise NotImplementedError(f"subterrain type '{kind}' is not built (supported: {sorted(TILE_DEFAULTS)})")


def build_subterrain_grid(subterrain_types, subterrain_size=(12.0, 12.0), horizontal_scale=0.25, vertical_scale=0.005, randomize=False,
                          subterrain_parameters=None):
    """Height field (float32, units of vertical_scale) of a grid of tiles; tile (i, j) covers rows i * (R - 1) .. (i + 1) * (R - 1)
    with R = int(size_x / horizontal_scale + eps) + 1 cells, neighbours share their border row / column (maximum of the two)."""
    types = [list(r) for r in subterrain_types]
    R = int(subterrain_size[0] / horizontal_scale + GS_EPS) + 1
    Cc = int(subterrain_size[1] / horizontal_scale + GS_EPS) + 1
    hf = np.full((len(types) * (R - 1) + 1, len(types[0]) * (Cc - 1) + 1), -np.inf, dtype=np.float32)
    for i, row in enumerate(types):
        for j, kind in enumerate(row):
            params = dict(TILE_DEFAULTS.get(kind, {}), **((subterrain_parameters or {}).get(kind, {})))
            if not randomize:
                state = np.random.get_state()
                np.random.seed(0)
            tile = _tile(kind, R, Cc, horizontal_scale, vertical_scale, params)
            if not randomize:
                np.random.set_state(state)
            view = hf[i * (R - 1):(i + 1) * (R - 1) + 1, j * (Cc - 1):(j + 1) * (Cc - 1) + 1]
            view[:] = np.maximum(view, tile)
    return hf
