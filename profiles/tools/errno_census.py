import json, os, sys, torch
ROOT="/root/repo"
sys.path[:0]=[ROOT, ROOT+"/go2-sim2real-locomotion-rl_b200"]
from go2_b200.envs import Go2EnvWalk, Go2EnvStair, Go2EnvBaseFlavour
for wl,B,cls in (("walk",65536,Go2EnvWalk),("stair",8192,Go2EnvStair),("crouch",4096,Go2EnvBaseFlavour),("jump",4096,Go2EnvBaseFlavour)):
    cfgs=json.load(open(f"{ROOT}/go2-sim2real-locomotion-rl_b200/go2_b200/cfgs/{wl}.json"))
    env=cls(B,*cfgs,device="cuda:0",seed=1,stats_sync_interval=24)
    env.reset()
    env.episode_length_buf = torch.randint(0, int(env.max_episode_length), (B,), device="cuda", dtype=torch.int32)
    g=torch.Generator(device="cuda").manual_seed(1)
    try:
        for i in range(600):
            env.step((0.6 if wl in ("crouch","jump") else 0.5)*torch.randn(B,env.num_actions,device="cuda",generator=g))
        env.flush_stats()
        err=None
    except RuntimeError as e:
        err=str(e)
    u,c=torch.unique(env.core.buf["errno"],return_counts=True)
    print(wl, dict(zip(u.tolist(), c.tolist())), "max contacts", int(env.core.buf["ncontact"].max()), "raised:", err, env.extras.get("errno_env_ticks"))
