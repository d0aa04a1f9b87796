"""Benchmark of the hot path: env-steps/s of the fused Go2 walk step.

    python bench.py --gpus N --steps K --warmup W [--envs B] [--impl reference]

A "step" = one `Go2Env.step(actions)` over all B envs of a rank (2 physics substeps + all env logic).
Workload (BASELINE.json configs[4], the one the north-star target is quoted on): Go2 walk, full
domain-randomisation cfg of go2_train_walk.get_cfgs(), 65,536 envs per GPU, random actions resident on the
device.  `value` = device-resident throughput through the Python env API (incl. its per-step host sync for
episode statistics, as the reference env does); `e2e` = the same through the C-ABI host-buffer call
(pinned actions H2D + obs/rew/done D2H inside the timed region).
`--impl reference` times the CPU oracle (restatement of the reference's rigid step; the reference itself
cannot be installed here: its solver needs the `quadrants` JIT which is not in the wheelhouse) on all host
cores on a bounded sample of the same workload.
"""
import argparse
import json
import os
import subprocess
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
BYTES_PER_ENV_STEP = {"walk": 1952, "stair": 2272, "crouch": 948, "jump": 948}  # algorithmic bytes per env-step
# from profiles/r01_v4_step_kernel_ncu_summary.txt (ncu --set full, walk, 65,536 envs, one launch)
NCU_PROFILE = "profiles/r01_v4_step_kernel_ncu_summary.txt"
NCU_DRAM_BYTES_PER_LAUNCH_65536 = 2.516372e9 + 2.011230e9
NCU_WARP_INSTR_PER_ENV_STEP = 495122640 / 65536


def load_cfgs(name):
    with open(os.path.join(PKG, "go2_b200", "cfgs", name + ".json")) as f:
        return json.load(f)


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


def tensor_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["bf16_tflops"])
    except Exception:
        return 2250.0  # nominal dense bf16 (B200_PROFILING.md fallback)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region (B200_PROFILING.md)."""

    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index=0):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.t.join(timeout=2)
        except Exception:
            pass
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 6 for i in range(4) if r[2 + i].lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons, "samples": len(sm)}


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
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
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
    sample = f"oracle rigid step (2 substeps, torque control; env-level torch ops excluded) on {n_envs} envs x {per_step} ticks per bench step"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * n_envs * per_step / v, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic", "config": {"workload": f"Go2 walk, CPU oracle, {n_envs} envs (bounded sample of the 65536-env/GPU workload)"},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def run_gpu(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    from go2_b200.envs import Go2EnvBaseFlavour, Go2EnvStair, Go2EnvWalk

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    B = args.envs
    env_cls = {"walk": Go2EnvWalk, "stair": Go2EnvStair}.get(args.workload, Go2EnvBaseFlavour)
    env = env_cls(B, *load_cfgs(args.workload), device=dev, seed=1, env_offset=rank * B, total_envs=world * B)
    env.reset()
    g = torch.Generator(device=dev).manual_seed(1234 + rank)
    n_act = 8
    acts = [0.5 * torch.randn(B, env.num_actions, device=dev, generator=g) for _ in range(n_act)]

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident throughput through the env API
    for i in range(args.warmup):
        env.step(acts[i % n_act])
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = env.core.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    kev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    barrier()
    ev0.record()
    for i in range(args.steps):
        # kernel-only events bracket just the launch of the fused step (same stream)
        sp = env._make_params(physics=True)
        kev[i][0].record()
        env.core.step(acts[i % n_act], sp)
        kev[i][1].record()
        if env.stats_sync_interval > 0 and env._step_count % env.stats_sync_interval == 0:
            env._after_resets(env._read_stats())
    ev1.record()
    barrier()
    launches = env.core.launch_count() - l0
    ms_total = ev0.elapsed_time(ev1)
    kern_ms = sorted(a.elapsed_time(b) for a, b in kev)
    kern_avg = sum(kern_ms) / len(kern_ms)
    clocks = sampler.stop() if rank == 0 else None

    # ---- end to end through the C ABI with pinned host buffers
    h_act = [a.cpu().pin_memory() for a in acts]
    h_obs = torch.empty((B, env.num_obs), dtype=torch.float32).pin_memory()
    h_rew = torch.empty((B,), dtype=torch.float32).pin_memory()
    h_done = torch.empty((B,), dtype=torch.uint8).pin_memory()
    for i in range(max(3, args.warmup // 2)):
        env.step_host(h_act[i % n_act], h_obs, h_rew, h_done)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for i in range(args.steps):
        env.step_host(h_act[i % n_act], h_obs, h_rew, h_done)
    e1.record()
    barrier()
    ms_e2e = e0.elapsed_time(e1)

    # ---- closed loop with a policy in the timed region (BASELINE north star: "with the shipped go2-walk checkpoint
    # policy"; the blob is missing from the reference checkout, so: the same 49-512-256-128-16 ELU actor of
    # go2_train_walk.py's train_cfg, random-initialised with a fixed seed, evaluated by torch in fp32)
    torch.manual_seed(188000)
    dims = [env.num_obs, 512, 256, 128, env.num_actions]
    layers = []
    for i in range(4):
        layers.append(torch.nn.Linear(dims[i], dims[i + 1]))
        if i < 3:
            layers.append(torch.nn.ELU())
    actor = torch.nn.Sequential(*layers).to(dev).eval()
    obs = env.obs_buf
    with torch.no_grad():
        for i in range(max(3, args.warmup // 2)):
            obs, _, _, _ = env.step(actor(obs))
        p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        p0.record()
        for i in range(args.steps):
            obs, _, _, _ = env.step(actor(obs))
        p1.record()
        barrier()
    ms_pol = p0.elapsed_time(p1)

    # ---- the same closed loop on the fused tcgen05 actor (go2_b200.rollout, csrc/go2_policy.cu), the kernel timed alone
    # against torch fp32 on the same weights, and 24-step rollout collection (actor sample + log-prob, critic, env step,
    # storage rows, GAE) as rsl-rl's OnPolicyRunner does it (BASELINE.json configs[4])
    from go2_b200.rollout import FusedActorCritic, RolloutCollector
    _po = env.get_privileged_observations()
    n_priv = _po.shape[1] if _po is not None else env.num_obs
    cdims = [n_priv, 512, 256, 128, 1]
    clayers = []
    for i in range(4):
        clayers.append(torch.nn.Linear(cdims[i], cdims[i + 1]))
        if i < 3:
            clayers.append(torch.nn.ELU())
    critic = torch.nn.Sequential(*clayers).to(dev).eval()
    ac = FusedActorCritic(actor, critic, 0.5 * torch.ones(env.num_actions), dev, seed=188000, env_offset=rank * B)
    act_buf = torch.empty((B, env.num_actions), device=dev)
    obs = env.obs_buf
    for i in range(max(3, args.warmup // 2)):
        obs, _, _, _ = env.step(ac.act_inference(obs, act_buf))
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    f0.record()
    for i in range(args.steps):
        obs, _, _, _ = env.step(ac.act_inference(obs, act_buf))
    f1.record()
    barrier()
    ms_fused = f0.elapsed_time(f1)
    x = torch.randn(B, env.num_obs, device=dev)
    def _time(fn, n=20):
        for _ in range(3):
            fn()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record()
        for _ in range(n):
            fn()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / n
    with torch.no_grad():
        ms_k_fused = _time(lambda: ac.act_inference(x, act_buf))
        ms_k_torch = _time(lambda: actor(x))
    col = RolloutCollector(env, ac, num_steps=24)
    o_, c_ = col.collect()
    r0, r1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n_roll = max(1, args.steps // 24)
    barrier()
    r0.record()
    for _ in range(n_roll):
        o_, c_ = col.collect(o_, c_)
    r1.record()
    barrier()
    ms_roll = r0.elapsed_time(r1)

    t = torch.tensor([ms_total, ms_e2e, kern_avg, ms_pol, ms_fused, ms_roll, ms_k_fused, ms_k_torch], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total, ms_e2e, kern_avg, ms_pol, ms_fused, ms_roll, ms_k_fused, ms_k_torch = [float(x) for x in t.tolist()]
    if rank == 0:
        value = world * B * args.steps / (ms_total * 1e-3)
        e2e = world * B * args.steps / (ms_e2e * 1e-3)
        peak, which = peaks()
        bpe = BYTES_PER_ENV_STEP[args.workload]
        ach = B * bpe / (kern_avg * 1e-3) / 1e9
        walk64k = args.workload == "walk" and B == 65536
        sm_mhz = (clocks or {}).get("sm_mhz") or 1965
        n_sm = torch.cuda.get_device_properties(dev).multi_processor_count
        issue_frac = (NCU_WARP_INSTR_PER_ENV_STEP * B / (kern_avg * 1e-3)) / (4.0 * n_sm * sm_mhz * 1e6) if walk64k else None
        cores = os.cpu_count() or 1
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            n_envs = max(64, 64 * cores)
            cpu_ticks = 2500  # ~10 s of host work on this pool's 16-core boxes
            v, dt_cpu = cpu_oracle_throughput(n_envs, cpu_ticks, cores)
            cpu = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                   "sample": f"oracle rigid step (2 substeps, torque control; env-level torch ops excluded), {n_envs} envs x {cpu_ticks} ticks, {dt_cpu:.1f} s"}
        stats = env.core.buf
        print(json.dumps({
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": f"Go2 {args.workload} (go2_train_{args.workload} cfg{', full DR + curriculum' if args.workload in ('walk', 'stair') else ''}), {B} envs/GPU, random actions on device",
                       "envs_per_gpu": B, "l2": "per-tick state (2 KB/env) + thread-local working set of the resident threads (~300 MB at 65,536 envs) exceed the 126 MB L2: inputs larger than L2, no flush needed",
                       "stats_sync_interval": env.stats_sync_interval},
            "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": B * env.num_actions * 4, "d2h_bytes_per_step": B * (env.num_obs * 4 + 4 + 1),
                    "ms_per_step": ms_e2e / args.steps},
            "with_policy": {"value": world * B * args.steps / (ms_pol * 1e-3), "unit": UNIT, "ms_per_step": ms_pol / args.steps,
                            "policy": f"{env.num_obs}-512-256-128-{env.num_actions} ELU actor, fixed-seed random init (shipped model_188000.pt blob absent), torch fp32, inside the timed loop"},
            "with_policy_fused": {"value": world * B * args.steps / (ms_fused * 1e-3), "unit": UNIT, "ms_per_step": ms_fused / args.steps,
                                  "policy": "same actor on the fused tcgen05 kernel (bf16 operands, fp32 accumulate), one launch per tick"},
            "rollout": {"value": world * B * 24 * n_roll / (ms_roll * 1e-3), "unit": UNIT, "ms_per_step": ms_roll / (24 * n_roll), "rollouts": n_roll,
                        "what": "24-step collection: fused actor (sample + log-prob) + fused critic + env step + rsl-rl storage rows + GAE"},
            "policy_kernel": {"kernel": "go2_mlp_kernel", "bound": "tensor", "ms_fused": ms_k_fused, "ms_torch_fp32": ms_k_torch,
                              "achieved": 2.0 * B * sum(dims[i] * dims[i + 1] for i in range(4)) / (ms_k_fused * 1e-3) / 1e12,
                              "peak": tensor_peak(), "unit": "TFLOP/s"},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": NCU_DRAM_BYTES_PER_LAUNCH_65536 if walk64k else None,
                         "traffic_unit": f"B per launch (dram read+write, ncu --set full, {NCU_PROFILE})",
                         "algorithmic_bytes_per_launch": B * bpe, "issue_frac": issue_frac,
                         "peak_source": which, "kernel": "go2_step_kernel", "kernel_ms_avg": kern_avg, "kernel_ms_median": kern_ms[len(kern_ms) // 2],
                         "bytes_per_env_step": bpe,
                         "note": "thread-local-memory latency bound, not HBM-bandwidth shaped (DESIGN.md 4 item 9): HBM fraction is the designated accounting; issue_frac = issued warp-instr/s (ncu instr count per env-step x measured rate) / (4 x SMs x clock)"},
            "cpu_baseline": cpu, "clocks": clocks,
            "diag": {"mean_contacts": float(stats["ncontact"].float().mean()), "mean_newton_iters": float(stats["nnewton"].float().mean()),
                     "errno_envs": int((stats["errno"] != 0).sum())},
        }))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=48)
    ap.add_argument("--warmup", type=int, default=12)
    ap.add_argument("--envs", type=int, default=65536, help="envs per GPU")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--workload", default="walk", choices=["walk", "stair", "crouch", "jump"],
                    help="walk = headline metric; stair = BASELINE configs[2]; crouch / jump = configs[3] (base env)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
