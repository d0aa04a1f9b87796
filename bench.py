"""Benchmark of the hot path: env-steps/s of the fused Go2 env step.

    python bench.py --gpus N --steps K --warmup W [--envs B] [--workload walk|stair|crouch|jump] [--all-configs] [--impl reference]

A "step" = one `Go2Env.step(actions)` over all B envs of a rank (2 physics substeps + all env logic).
Headline workload (BASELINE.json configs[4], the one the north-star target is quoted on): Go2 walk, full
domain-randomisation cfg of go2_train_walk.get_cfgs(), 65,536 envs per GPU, random actions resident on the device.
  value     device-resident throughput through the public env API (`env.step`), CUDA events, max over ranks; measured on a
            steady-state env population (episode lengths randomised like rsl-rl does, `--settle` untimed ticks after reset()) -
            `fresh_after_reset` is the same window right after reset(), the cheaper protocol earlier rounds quoted
  e2e       the same tick through the C-ABI host-buffer call go2_step_host (pinned actions H2D, obs / rew / done D2H inside
            the timed region)
  roofline  go2_step_kernel alone (events bracketing its launch inside the timed loop), algorithmic bytes vs measured HBM peak
`--all-configs` appends one line per other BASELINE config that fits one GPU (walk 4,096 + policy, stair 8,192, crouch / jump
4,096).  `--impl reference` times the CPU oracle port of the reference's rigid step on all host cores on a bounded sample of the
same workload (the reference itself cannot run here: its solver needs the `quadrants` JIT, absent from the wheelhouse).
GO2_B200_LIB=<path> benches another build of the CUDA library (kernel-variant A/B runs).
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "go2-sim2real-locomotion-rl_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)

METRIC = "env-steps/sec (whole box) Go2 walk"
UNIT = "env-steps/s"
BYTES_PER_ENV_STEP = {"walk": 1952, "stair": 2272, "crouch": 948, "jump": 948,  # algorithmic bytes per env-step
                      "stair_lidar": 2384,  # stair + 15 obs words - 17 privileged-obs words + the kept actor scan read and written (30)
                      "stair_info": 2044}   # stair + 4 obs words - 69 privileged-obs words + the kept stair info read and written (8)
NCU_JSON = os.path.join(ROOT, "profiles", "step_kernel_ncu.json")  # written from the ncu capture of the CURRENT build (profiles/tools/ncu_to_json.py)
DEFAULT_ENVS = {"walk": 65536, "stair": 8192, "stair_lidar": 8192, "stair_info": 8192, "crouch": 4096, "jump": 4096}


def load_cfgs(name):
    with open(os.path.join(PKG, "go2_b200", "cfgs", name + ".json")) as f:
        return json.load(f)


def workload_config(workload, B):
    dr = ", full DR + curriculum" if workload in ("walk", "stair", "stair_lidar", "stair_info") else ""
    return {"workload": f"Go2 {workload} (go2_train_{workload} cfg{dr}), {B} envs/GPU, random actions on device", "envs_per_gpu": B,
            "l2": "per-tick state (2 KB/env) + thread-local working set of the resident threads exceed the 126 MB L2 at 65,536 envs: inputs larger than L2, no flush"}


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            d = json.load(f)
        return float(d["hbm_gbs"]), float(d["bf16_tflops"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, 2250.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region through NVML (5 ms period)."""

    def __init__(self, index=0):
        self.index, self.sm, self.reasons, self.mx, self._stop, self.err = index, [], set(), None, False, None

    def start(self):
        try:
            import pynvml as nv

            nv.nvmlInit()
            self.nv, self.h = nv, nv.nvmlDeviceGetHandleByIndex(self.index)
            self.mx = nv.nvmlDeviceGetMaxClockInfo(self.h, nv.NVML_CLOCK_SM)
        except Exception as e:  # noqa: BLE001
            self.err = repr(e)
            return
        self.t = threading.Thread(target=self._run, daemon=True)
        self.t.start()

    def _run(self):
        nv = self.nv
        bits = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20}
        while not self._stop:
            try:
                self.sm.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h) if hasattr(nv, "nvmlDeviceGetCurrentClocksEventReasons") else nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                self.reasons |= {k for k, b in bits.items() if r & b}
            except Exception as e:  # noqa: BLE001
                self.err = repr(e)
                return
            time.sleep(0.005)

    def stop(self):
        self._stop = True
        if self.err and not self.sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvml unavailable: " + self.err]}
        try:
            self.t.join(timeout=1)
        except Exception:
            pass
        sm = sorted(self.sm)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": self.mx, "reasons": sorted(self.reasons), "samples": len(sm)}


_CPU_KIND = None


def _best_cpu_kind(threads):
    """The -O3 -march=native oracle build with or without FMA contraction, whichever is faster on this host."""
    global _CPU_KIND
    if _CPU_KIND is None:
        n = max(64, 8 * threads)
        rates = {k: max(cpu_oracle_throughput(n, 40, threads, kind=k)[0] for _ in range(2)) for k in ("f32_fast", "f32_fast_nc")}
        _CPU_KIND = max(rates, key=rates.get)
    return _CPU_KIND


def cpu_oracle_throughput(n_envs, n_steps, threads, kind=None):
    """Oracle rigid step (2 substeps, torque control held) on `threads` host threads; env-steps/s."""
    import numpy as np
    from oracle.rigid import OracleRigid

    o = OracleRigid(n_envs, kind=kind or _best_cpu_kind(threads), n_threads=threads)
    mdi = [7, 11, 15, 6, 10, 14, 9, 13, 17, 8, 12, 16]
    q0 = np.array([0, 0.8, -1.5, 0, 0.8, -1.5, 0, 1.0, -1.5, 0, 1.0, -1.5])
    o.set_dofs_position(np.tile(q0, (n_envs, 1)), mdi)
    rng = np.random.default_rng(0)
    kp, kd = 40.0, 1.26

    def ctrl():
        a = 0.5 * rng.standard_normal((n_envs, 12))
        tq = kp * (a * 0.25 + q0 - o.get_dofs_position()[:, mdi]) - kd * o.get_qvel()[:, mdi]
        o.control_dofs_force(np.clip(tq, -23.7, 23.7), mdi)

    for _ in range(10):
        ctrl(); o.step(2)
    t0 = time.perf_counter()
    for _ in range(n_steps):
        ctrl(); o.step(2)
    dt = time.perf_counter() - t0
    return n_envs * n_steps / dt, dt


def run_reference(args):
    """Reference arm: the CPU implementation of the path on the host cores (oracle port; see module doc)."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    cores = os.cpu_count() or 1
    n_envs = max(64, 32 * cores)
    vals = []
    for _ in range(max(1, args.warmup // 3)):
        cpu_oracle_throughput(n_envs, 5, cores)
    per_step = 20
    for _ in range(args.steps):
        v, _dt = cpu_oracle_throughput(n_envs, per_step, cores)
        vals.append(v)
    vals.sort()
    v = vals[len(vals) // 2]
    sample = (f"oracle port of the reference's rigid step (2 substeps, PD torque control; the env-level torch ops of the reference env are "
              f"NOT included, which favours this arm), {n_envs} envs x {per_step} ticks per bench step, {cores} threads")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * n_envs * per_step / v, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic", "config": workload_config(args.workload, args.envs or DEFAULT_ENVS[args.workload]),
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def make_policy(env, dev, seed=188000):
    """The 49-512-256-128-16 ELU actor and the critic of go2_train_walk.py's train_cfg, fixed-seed random init (the shipped
    model_188000.pt blob is absent from the reference checkout)."""
    import torch

    torch.manual_seed(seed)

    def mlp(dims):
        layers = []
        for i in range(4):
            layers.append(torch.nn.Linear(dims[i], dims[i + 1]))
            if i < 3:
                layers.append(torch.nn.ELU())
        return torch.nn.Sequential(*layers).to(dev).eval()

    po = env.get_privileged_observations()
    n_priv = po.shape[1] if po is not None else env.num_obs
    return mlp([env.num_obs, 512, 256, 128, env.num_actions]), mlp([n_priv, 512, 256, 128, 1])


def bench_env(args, workload, B, rank, world, dev, headline):
    import torch
    import torch.distributed as dist
    from go2_b200.envs import Go2EnvBaseFlavour, Go2EnvStair, Go2EnvStairInfo, Go2EnvStairLidar, Go2EnvWalk

    env_cls = {"walk": Go2EnvWalk, "stair": Go2EnvStair, "stair_lidar": Go2EnvStairLidar, "stair_info": Go2EnvStairInfo}.get(workload, Go2EnvBaseFlavour)
    env = env_cls(B, *load_cfgs(workload), device=dev, seed=1, env_offset=rank * B, total_envs=world * B,
                  stats_sync_interval=args.stats_interval, lib_path=os.environ.get("GO2_B200_LIB") or None, fast_math=(args.math == "fast"), sync_ranks=args.rank_sync)
    env.reset()
    g = torch.Generator(device=dev).manual_seed(1234 + rank)
    n_act = 8
    acts = [0.5 * torch.randn(B, env.num_actions, device=dev, generator=g) for _ in range(n_act)]
    K, W = args.steps, max(3, args.warmup)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, k=K):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        a.record()
        for i in range(k):
            fn(i)
        b.record()
        barrier()
        return a.elapsed_time(b)

    # ---- population like a running training job, not like the tick after reset(): episode lengths randomised the way rsl-rl's
    # OnPolicyRunner.learn(init_at_random_ep_len=True) does (go2_train_walk.py:476) - otherwise all envs time out in the same tick and
    # no reset ever falls into the window - and `--settle` untimed ticks (default 240 = BASELINE C5's 10 warm-up rollouts of 24) so
    # that the poses have left the all-standing start (step time rises ~6 % over the first 300 ticks, profiles/tools/tick_time_drift.py)
    if headline and args.settle > 0:  # the window right after reset() that rounds 1 / early 2 quoted, kept for continuity
        for i in range(W):
            env.step(acts[i % n_act])
        ms_fresh = timed(lambda i: env.step(acts[i % n_act]))
    else:
        ms_fresh = None
    env.episode_length_buf = torch.randint(0, int(env.max_episode_length), (B,), device=dev, generator=g, dtype=torch.int32)
    for i in range(args.settle):
        env.step(acts[i % n_act])
    for i in range(W):
        env.step(acts[i % n_act])
    kev = []

    def hook():
        kev.append((torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)))
        return kev[-1]

    sampler = ClockSampler(dev.index or 0)
    if rank == 0 and headline:
        sampler.start()
    l0 = env.core.launch_count()
    env._kernel_event_hook = hook
    ms_total = timed(lambda i: env.step(acts[i % n_act]))
    env._kernel_event_hook = None
    launches = env.core.launch_count() - l0
    clocks = sampler.stop() if (rank == 0 and headline) else None
    kern_ms = sorted(a.elapsed_time(b) for a, b in kev) or [ms_total / K]
    kern_avg = sum(kern_ms) / len(kern_ms)
    res = {"env": env, "ms_fresh": ms_fresh, "ms_total": ms_total, "kern_avg": kern_avg, "kern_med": kern_ms[len(kern_ms) // 2], "launches": launches, "clocks": clocks,
           "shape": env.core.launch_shape()}

    # ---- end to end through the C ABI with pinned host buffers
    h_act = [a.cpu().pin_memory() for a in acts]
    h_obs = torch.empty((B, env.num_obs), dtype=torch.float32).pin_memory()
    h_rew = torch.empty((B,), dtype=torch.float32).pin_memory()
    h_done = torch.empty((B,), dtype=torch.uint8).pin_memory()
    for i in range(W):
        env.step_host(h_act[i % n_act], h_obs, h_rew, h_done)
    res["ms_e2e"] = timed(lambda i: env.step_host(h_act[i % n_act], h_obs, h_rew, h_done))

    # ---- closed loop with the policy in the timed region (torch fp32 actor, then the fused tcgen05 actor), rollout collection
    if args.policy or headline:
        from go2_b200.rollout import FusedActorCritic, RolloutCollector

        actor, critic = make_policy(env, dev)
        state = {"obs": env.obs_buf}

        def torch_loop(i):
            state["obs"] = env.step(actor(state["obs"]))[0]

        with torch.no_grad():
            for i in range(W):
                torch_loop(i)
            res["ms_pol"] = timed(torch_loop)
        ac = FusedActorCritic(actor, critic, 0.5 * torch.ones(env.num_actions), dev, seed=188000, env_offset=rank * B)
        act_buf = torch.empty((B, env.num_actions), device=dev)

        def fused_loop(i):
            state["obs"] = env.step(ac.act_inference(state["obs"], act_buf))[0]

        for i in range(W):
            fused_loop(i)
        res["ms_fused"] = timed(fused_loop)
        x = torch.randn(B, env.num_obs, device=dev)
        with torch.no_grad():
            res["ms_k_fused"] = timed(lambda i: ac.act_inference(x, act_buf), 20) / 20
            res["ms_k_torch"] = timed(lambda i: actor(x), 20) / 20
        col = RolloutCollector(env, ac, num_steps=24)
        st = {"oc": col.collect()}
        n_roll = max(1, K // 24)

        def roll(i):
            st["oc"] = col.collect(*st["oc"])

        res["ms_roll"], res["n_roll"] = timed(roll, n_roll), n_roll
        res["actor_macs"] = sum(a * b for a, b in zip([env.num_obs, 512, 256, 128], [512, 256, 128, env.num_actions]))
    env.flush_stats()
    if headline and env.fast_math and not os.environ.get("GO2_B200_LIB"):
        # the same window on the IEEE-division build (the one the bit-faithful parity claim is stated for), for reference
        env2 = env_cls(B, *load_cfgs(workload), device=dev, seed=1, env_offset=rank * B, total_envs=world * B, stats_sync_interval=args.stats_interval,
                       fast_math=False)
        env2.reset()
        env2.episode_length_buf = torch.randint(0, int(env2.max_episode_length), (B,), device=dev, generator=g, dtype=torch.int32)
        for i in range(args.settle + W):
            env2.step(acts[i % n_act])
        res["ms_precise"] = timed(lambda i: env2.step(acts[i % n_act]))
        env2.flush_stats()
        env2.close()
    return res


def emit(args, workload, B, world, res, headline, cpu):
    import torch

    K = args.steps
    env = res["env"]
    hbm_peak, tensor_peak, which = peaks()
    bpe = BYTES_PER_ENV_STEP[workload]
    value = world * B * K / (res["ms_total"] * 1e-3)
    ach = B * bpe / (res["kern_avg"] * 1e-3) / 1e9
    ncu = None
    if workload == "walk" and B == 65536 and os.path.exists(NCU_JSON):
        with open(NCU_JSON) as f:
            ncu = json.load(f)
    sm_mhz = ((res["clocks"] or {}).get("sm_mhz") or 1965)
    n_sm = torch.cuda.get_device_properties(env.device).multi_processor_count
    issue_frac = (ncu["warp_instr_per_launch"] / (res["kern_avg"] * 1e-3)) / (4.0 * n_sm * sm_mhz * 1e6) if ncu else None
    b = env.core.buf
    out = {
        "metric": METRIC if workload == "walk" else METRIC.replace("walk", workload), "value": value, "unit": UNIT, "n_gpus": world, "steps": K,
        "warmup": max(3, args.warmup), "ms_per_step": res["ms_total"] / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": dict(workload_config(workload, B), stats_sync_interval=env.stats_sync_interval, settle_ticks=args.settle, rank_sync=bool(args.rank_sync),
                                                       math=("fast (approximate division / sqrt like the reference's fast_math; libgo2_b200_fast.so)" if env.fast_math else "precise (IEEE division / sqrt; libgo2_b200.so)"),
                                                       population="episode lengths randomised (rsl-rl init_at_random_ep_len), settle_ticks untimed ticks before the warm-up: steady-state poses and a steady stream of time-out resets"),
        "e2e": {"value": world * B * K / (res["ms_e2e"] * 1e-3), "unit": UNIT, "h2d_bytes_per_step": B * env.num_actions * 4,
                "d2h_bytes_per_step": B * (env.num_obs * 4 + 4 + 1), "ms_per_step": res["ms_e2e"] / K,
                "what": "go2_step_host: pinned actions H2D (copy stream, beside the kernel's opening phase), step + post kernels, rew / done / obs written to the pinned host buffers by the kernel, stream sync - every tick"},
        "gpu_launches": int(res["launches"]),
        "launch_shape": res["shape"],
        "roofline": {"bound": "hbm", "achieved": ach, "peak": hbm_peak, "unit": "GB/s", "frac": ach / hbm_peak,
                     "traffic": ncu["dram_bytes_per_launch"] if ncu else None,
                     "traffic_source": (ncu or {}).get("source"), "algorithmic_bytes_per_launch": B * bpe, "bytes_per_env_step": bpe,
                     "issue_frac": issue_frac, "peak_source": which, "kernel": "go2_step_kernel_small" if res["shape"]["small_kernel"] else "go2_step_kernel",
                     "kernel_ms_avg": res["kern_avg"], "kernel_ms_median": res["kern_med"],
                     "note": "latency bound on thread-local memory, not HBM-bandwidth shaped (DESIGN.md 4): the HBM fraction is the designated accounting; "
                             "issue_frac = ncu warp-instructions per launch / kernel time / (4 schedulers x SMs x clock)"},
        "cpu_baseline": cpu, "clocks": res["clocks"],
        "diag": {"mean_contacts": float(b["ncontact"].float().mean()), "mean_newton_iters": float(b["nnewton"].float().mean()),
                 "errno_envs": int((b["errno"] != 0).sum()),
                 "note": "throughput depends on the action distribution: 0.5*randn actions here; a policy's smaller outputs give fewer contacts / Newton iterations (with_policy_fused)"},
    }
    if res.get("ms_fresh"):
        out["fresh_after_reset"] = {"value": world * B * K / (res["ms_fresh"] * 1e-3), "unit": UNIT, "ms_per_step": res["ms_fresh"] / K,
                                    "what": "the same K ticks timed right after reset() + W warm-up ticks (the protocol of BENCH_r01): every env still near the standing start pose, no resets in the window - cheaper than a running job, reported for continuity only"}
    if res.get("ms_precise"):
        out["precise_math"] = {"value": world * B * K / (res["ms_precise"] * 1e-3), "unit": UNIT, "ms_per_step": res["ms_precise"] / K,
                               "what": "the same steady-state window on libgo2_b200.so (IEEE division / sqrt): the build whose results are bit-faithful to the CPU oracle's operation order"}
    if "ms_pol" in res:
        out["with_policy"] = {"value": world * B * K / (res["ms_pol"] * 1e-3), "unit": UNIT, "ms_per_step": res["ms_pol"] / K,
                              "policy": f"{env.num_obs}-512-256-128-{env.num_actions} ELU actor, fixed-seed random init (shipped model_188000.pt blob absent), torch fp32, inside the timed loop"}
        out["with_policy_fused"] = {"value": world * B * K / (res["ms_fused"] * 1e-3), "unit": UNIT, "ms_per_step": res["ms_fused"] / K,
                                    "policy": "same actor on the fused tcgen05 kernel (bf16 operands, fp32 accumulate), one launch per tick"}
        out["rollout"] = {"value": world * B * 24 * res["n_roll"] / (res["ms_roll"] * 1e-3), "unit": UNIT, "ms_per_step": res["ms_roll"] / (24 * res["n_roll"]),
                          "rollouts": res["n_roll"], "what": "24-step collection: fused actor (sample + log-prob) + fused critic + env step + rsl-rl storage rows + GAE"}
        out["policy_kernel"] = {"kernel": "go2_mlp_kernel", "bound": "tensor", "ms_fused": res["ms_k_fused"], "ms_torch_fp32": res["ms_k_torch"],
                                "achieved": 2.0 * B * res["actor_macs"] / (res["ms_k_fused"] * 1e-3) / 1e12, "peak": tensor_peak, "unit": "TFLOP/s"}
    print(json.dumps(out), flush=True)


def run_gpu(args):
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    jobs = [(args.workload, args.envs or DEFAULT_ENVS[args.workload], True)]
    if args.all_configs:
        jobs += [("walk", 4096, False), ("stair", 8192, False), ("stair_lidar", 8192, False), ("stair_info", 8192, False), ("crouch", 4096, False), ("jump", 4096, False)]
    for workload, B, headline in jobs:
        res = bench_env(args, workload, B, rank, world, dev, headline)
        keys = ["ms_total", "ms_e2e", "kern_avg", "kern_med"] + [k for k in ("ms_pol", "ms_fused", "ms_roll", "ms_k_fused", "ms_k_torch", "ms_fresh", "ms_precise") if res.get(k) is not None]
        t = torch.tensor([res[k] for k in keys], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        res.update(dict(zip(keys, [float(x) for x in t.tolist()])))
        if rank == 0:
            cpu = None
            if world == 1 and headline and not args.no_cpu_baseline:
                cores = os.cpu_count() or 1
                n_envs, cpu_ticks = max(64, 64 * cores), 2500  # ~10 s of host work on this pool's 16-core boxes
                v, dt_cpu = cpu_oracle_throughput(n_envs, cpu_ticks, cores)
                cpu = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                       "sample": f"oracle rigid step (2 substeps, torque control; env-level torch ops excluded), {n_envs} envs x {cpu_ticks} ticks, {dt_cpu:.1f} s"}
            emit(args, workload, B, world, res, headline, cpu)
        res["env"].close()
        del res
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=48)
    ap.add_argument("--warmup", type=int, default=12)
    ap.add_argument("--math", default="fast", choices=["fast", "precise"],
                    help="fast = the library build with approximate division / sqrt, the arithmetic class of the reference (its solver is JIT-compiled "
                         "with fast_math on, genesis/__init__.py:289) and what the drop-in genesis.init() selects; precise = IEEE division / sqrt, the "
                         "build the bit-faithful parity against the CPU oracle is stated for")
    ap.add_argument("--rank-sync", action="store_true",
                    help="N > 1: all-reduce the statistics row every tick so that all ranks run ONE curriculum / global-DR state machine (default: "
                         "independent env shards, no collective on the step path - what N copies of the reference script would be)")
    ap.add_argument("--settle", type=int, default=240, help="untimed ticks after reset() before the warm-up (steady-state population); 0 = time right after reset()")
    ap.add_argument("--envs", type=int, default=0, help="envs per GPU (default: 65,536 walk, 8,192 stair, 4,096 crouch / jump)")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--policy", action="store_true", help="policy / rollout legs for the non-headline configs too")
    ap.add_argument("--all-configs", action="store_true", help="one more JSON line per other BASELINE config (walk 4,096, stair 8,192, crouch / jump 4,096)")
    ap.add_argument("--stats-interval", type=int, default=24, help="ticks between asynchronous reads of the device-side episode statistics (1 = blocking read every tick)")
    ap.add_argument("--workload", default="walk", choices=["walk", "stair", "stair_lidar", "stair_info", "crouch", "jump"],
                    help="walk = headline metric; stair = BASELINE configs[2]; crouch / jump = configs[3] (base env)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
