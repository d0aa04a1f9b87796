"""TEST INFRASTRUCTURE - runs the REFERENCE's own tile generators (genesis/ext/isaacgym/terrain_utils.py) and the tile loop of
genesis/utils/terrain.py:parse_terrain in this container (needs /root/reference) without importing Genesis: the two files' functions are
exec'd from their source with a stand-in `gs` (np_float / EPS of precision "32") and `gu` (cubic_spline_1d taken from
genesis/utils/geom.py).  Prints the sha256 digests tests/test_terrain_grid.py holds.

    python oracle/ref_terrain_tiles.py
"""
import ast
import hashlib
import os
import types

import numpy as np

REF = "/root/reference"


def _funcs(path, names, ns):
    tree = ast.parse(open(path).read())
    body = [n for n in tree.body if isinstance(n, (ast.FunctionDef, ast.ClassDef)) and n.name in names]
    exec(compile(ast.Module(body=body, type_ignores=[]), path, "exec"), ns)
    return ns


def reference_namespace():
    gs = types.ModuleType("gs")  # a module, so that numba can resolve gs.np_float as a constant
    gs.np_float, gs.EPS = np.float32, np.finfo(np.float32).eps
    try:  # the reference compiles the spline with numba (geom.py:2150); without numba (or REF_NO_NUMBA=1) the same source runs as plain numpy
        if os.environ.get("REF_NO_NUMBA"):
            raise ImportError
        import numba

        nb = types.SimpleNamespace(jit=lambda *a, **k: numba.jit(nopython=True))  # cache=True wants a real source file location
    except ImportError:
        nb = types.SimpleNamespace(jit=lambda *a, **k: (lambda f: f))
    geom = _funcs(os.path.join(REF, "genesis/utils/geom.py"), ["cubic_spline_1d"], {"np": np, "gs": gs, "nb": nb})
    gu = types.SimpleNamespace(cubic_spline_1d=geom["cubic_spline_1d"])
    names = ["random_uniform_terrain", "sloped_terrain", "pyramid_sloped_terrain", "discrete_obstacles_terrain", "wave_terrain",
             "stairs_terrain", "pyramid_stairs_terrain", "stepping_stones_terrain", "SubTerrain"]
    ns = _funcs(os.path.join(REF, "genesis/ext/isaacgym/terrain_utils.py"), names, {"np": np, "gs": gs, "gu": gu})
    ns["fractal_terrain"] = _funcs(os.path.join(REF, "genesis/utils/terrain.py"), ["fractal_terrain"], {"np": np, "gs": gs})["fractal_terrain"]
    return ns


def reference_grid(ns, types_, size, hs, vs, randomize, params):
    """the tile loop of parse_terrain (genesis/utils/terrain.py:64-160) around the reference generators"""
    from go2_b200.terrain import TILE_DEFAULTS  # Terrain.default_params (morphs.py:1313), plain data

    eps = np.finfo(np.float32).eps
    R, C = int(size[0] / hs + eps) + 1, int(size[1] / hs + eps) + 1
    hf = np.full((len(types_) * (R - 1) + 1, len(types_[0]) * (C - 1) + 1), float("-inf"), dtype=np.float32)
    for i in range(len(types_)):
        for j in range(len(types_[0])):
            kind = types_[i][j]
            p = dict(TILE_DEFAULTS[kind], **(params or {}).get(kind, {}))
            t = ns["SubTerrain"](width=R, length=C, vertical_scale=vs, horizontal_scale=hs)
            if not randomize:
                st = np.random.get_state()
                np.random.seed(0)
            sub = None if kind == "flat_terrain" else ns[kind](t, **p)
            if not randomize:
                np.random.set_state(st)
            data = np.zeros((R, C), np.float32) if sub is None else sub.height_field_raw
            v = hf[i * (R - 1):(i + 1) * (R - 1) + 1, j * (C - 1):(j + 1) * (C - 1) + 1]
            v[:] = np.maximum(v, data)
    return hf


CASES = {
    "train_cfg_2x2": dict(types_=[["flat_terrain", "random_uniform_terrain"], ["random_uniform_terrain", "flat_terrain"]], size=(8.0, 8.0), hs=0.25, vs=0.005,
                          randomize=True, params={"random_uniform_terrain": {"min_height": -0.08, "max_height": 0.08}}),
    "all_types_fixed": dict(types_=[["flat_terrain", "random_uniform_terrain", "sloped_terrain", "pyramid_sloped_terrain"],
                                    ["discrete_obstacles_terrain", "wave_terrain", "stairs_terrain", "pyramid_stairs_terrain"]],
                            size=(12.0, 12.0), hs=0.25, vs=0.005, randomize=False, params=None),
    "fractal_stones": dict(types_=[["fractal_terrain", "stepping_stones_terrain"], ["stepping_stones_terrain", "fractal_terrain"]], size=(6.0, 8.0), hs=0.25, vs=0.005,
                           randomize=True, params={"fractal_terrain": {"levels": 5, "scale": 0.4}, "stepping_stones_terrain": {"stone_size": 0.75, "platform_size": 1.0}}),
    "stones_wide": dict(types_=[["stepping_stones_terrain", "fractal_terrain"]], size=(8.0, 5.0), hs=0.25, vs=0.005, randomize=True, params=None),
    "all_types_random": dict(types_=[["pyramid_stairs_terrain", "discrete_obstacles_terrain"], ["random_uniform_terrain", "wave_terrain"]],
                             size=(6.0, 9.0), hs=0.1, vs=0.005, randomize=True, params={"random_uniform_terrain": {"step": 0.02, "downsampled_scale": 0.3}}),
}

if __name__ == "__main__":
    import sys

    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "go2-sim2real-locomotion-rl_b200"))
    from go2_b200.terrain import build_subterrain_grid

    ns = reference_namespace()
    for name, c in CASES.items():
        np.random.seed(123)
        ref = reference_grid(ns, **c)
        after_ref = np.random.get_state()[1][:4].tolist()
        np.random.seed(123)
        mine = build_subterrain_grid(c["types_"], c["size"], c["hs"], c["vs"], c["randomize"], c["params"])
        same_rng = after_ref == np.random.get_state()[1][:4].tolist()
        print(name, ref.shape, "equal", np.array_equal(ref, mine), "rng state equal", same_rng, "range", float(ref.min()), float(ref.max()),
              hashlib.sha256(np.ascontiguousarray(ref).tobytes()).hexdigest())
