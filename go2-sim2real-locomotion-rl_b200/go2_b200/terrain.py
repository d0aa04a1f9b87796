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
