"""world_size-2 gloo checks of the two multi-process modes: independent env shards (default; no collective on the step path) and
`sync_ranks=True`, where the per-tick all-reduce of the episode statistics row keeps the curriculum / global-DR state machines of all
shards in lock-step."""
import os
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, lib, port, q, sync=True):
    sys.path[:0] = [ROOT, os.path.join(ROOT, "go2-sim2real-locomotion-rl_b200"), os.path.join(ROOT, "tests")]
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from util import load_golden, make_env

    _, cfgs = load_golden("walk_short")
    B = 4
    env = make_env("walk", cfgs, B, "hostemu", lib, env_offset=rank * B, total_envs=world * B, seed=7, sync_ranks=sync)
    env.reset()
    g = torch.Generator().manual_seed(100 + rank)
    n_resets = 0
    for _ in range(70):
        _, _, done, _ = env.step(0.8 * torch.randn(B, 16, generator=g))
        n_resets += int(done.sum())
    q.put((rank, n_resets, env._curr_ep_total, env._curr_timeout_total, env.global_dr_values()["friction_ground"], env._step_count,
           env.core.buf["commands"][0].tolist()))
    dist.destroy_process_group()


def test_two_rank_stats_allreduce(hostemu_lib):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 1000
    ps = [ctx.Process(target=_worker, args=(r, 2, hostemu_lib, port, q)) for r in range(2)]
    [p.start() for p in ps]
    res = sorted(q.get(timeout=300) for _ in ps)
    [p.join(timeout=60) for p in ps]
    (_, n0, tot0, to0, fr0, sc0, _), (_, n1, tot1, to1, fr1, sc1, _) = res
    assert n0 + n1 > 0
    # every rank accumulated the GLOBAL episode count / time-outs and holds identical global-DR scalars
    assert tot0 == tot1 == n0 + n1 + 8  # global count: the initial reset() of 2x4 envs plus every rank's step resets
    assert to0 == to1 and fr0 == fr1 and sc0 == sc1


def test_two_ranks_as_independent_shards(hostemu_lib):
    """Default mode: every rank counts only its own episodes (what N copies of the reference script would do); no collective is issued."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() + 500) % 1000
    ps = [ctx.Process(target=_worker, args=(r, 2, hostemu_lib, port, q, False)) for r in range(2)]
    [p.start() for p in ps]
    res = sorted(q.get(timeout=300) for _ in ps)
    [p.join(timeout=60) for p in ps]
    for _, n, tot, _, _, sc, _ in res:
        assert tot == n + 4 and sc == res[0][5]  # own resets + the initial reset() of its 4 envs
    assert res[0][1] != res[1][1] or res[0][6] != res[1][6]  # different action streams / env ids: the shards evolve differently
