import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from asimov_mjlab_b200.rl.velocity_env import G1VelocityFlatEnv
from asimov_mjlab_b200.rl.ppo import PPO
env = G1VelocityFlatEnv(256, "cuda:0", seed=0)
algo = PPO(env, torch.device("cuda:0"), steps_per_env=8, use_graph=False)
ac = algo.ac
def piece_norm(): ac.obs_norm.update(env.obs); ac.critic_norm.update(env.critic_obs)
def piece_dist():
  d = ac.dist(env.obs); algo._a = d.mean + d.stddev * torch.randn_like(d.mean); algo._d = d
def piece_val(): algo._v = ac.value(env.critic_obs)
def piece_put(): algo.b_obs.index_copy_(0, algo.t_idx, env.obs.unsqueeze(0))
def piece_env(): env._step_impl()
def piece_book():
  donef = env.dones.float(); algo.ep_ret += env.rew; algo.ep_len += 1
  algo.stats["ret_sum"] += (algo.ep_ret * donef).sum(); algo.ep_ret *= 1 - donef; algo.t_idx += 1
for name, fn in [("norm", piece_norm), ("dist", piece_dist), ("val", piece_val), ("put", piece_put), ("env", piece_env), ("book", piece_book)]:
  try:
    with torch.no_grad():
      s = torch.cuda.Stream(); s.wait_stream(torch.cuda.current_stream())
      with torch.cuda.stream(s):
        for _ in range(3): fn()
      torch.cuda.current_stream().wait_stream(s); torch.cuda.synchronize()
      g = torch.cuda.CUDAGraph()
      with torch.cuda.graph(g):
        fn()
      g.replay(); torch.cuda.synchronize()
    print(name, "ok")
  except Exception as e:
    print(name, "FAILED", str(e).splitlines()[0])
    break
