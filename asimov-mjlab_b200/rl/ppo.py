"""PPO as rsl_rl runs it for the reference's G1 velocity task, with the multi-GPU plumbing the reference lacks.

Hyper-parameters and structure follow src/mjlab/tasks/velocity/config/g1/rl_cfg.py:1-41 (actor / critic MLP 512-256-128
ELU, init_noise_std 1.0, empirical observation normalisation, clip 0.2, entropy 0.01, 5 epochs x 4 mini-batches,
lr 1e-3 adaptive on KL 0.01, gamma 0.99, lambda 0.95, max_grad_norm 1.0, 24 steps per env) and the on-policy loop of
scripts/train.py:36-124.  One process per GPU, each with its own env shard; the only collectives are
  * the gradient all-reduce (average) after every mini-batch backward,
  * the merge of the observation normalisers' (count, mean, M2) once per iteration,
  * the KL average that drives the adaptive learning rate, and the logged scalars
(none of this exists in the reference, train.py:29).  The MLP GEMMs are plain library calls (cuBLAS
through torch): the policy is not the hot path this repository accelerates.
"""
from __future__ import annotations

import torch
import torch.distributed as dist
import torch.nn as nn


def _world():
  return dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1


def all_reduce_mean_(t: torch.Tensor) -> torch.Tensor:
  if _world() > 1:
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    t /= _world()
  return t


class EmpiricalNormalization(nn.Module):
  """Running mean / variance (Welford, batched); `sync()` merges the replicas' statistics so all ranks stay identical."""

  def __init__(self, dim: int, eps: float = 1e-2):
    super().__init__()
    self.eps = eps
    self.register_buffer("mean", torch.zeros(dim))
    self.register_buffer("m2", torch.zeros(dim))
    self.register_buffer("count", torch.zeros(()))
    # statistics gathered since the last sync (local to this rank)
    self.register_buffer("d_mean", torch.zeros(dim))
    self.register_buffer("d_m2", torch.zeros(dim))
    self.register_buffer("d_count", torch.zeros(()))

  @staticmethod
  def _merge(mean_a, m2_a, n_a, mean_b, m2_b, n_b):
    n = n_a + n_b
    delta = mean_b - mean_a
    w = n_b / torch.clamp(n, min=1.0)
    return mean_a + delta * w, m2_a + m2_b + delta * delta * n_a * w, n

  @torch.no_grad()
  def update(self, x: torch.Tensor) -> None:
    n = torch.tensor(float(x.shape[0]), device=x.device)
    mean = x.mean(dim=0)
    m2 = ((x - mean) ** 2).sum(dim=0)
    self.d_mean, self.d_m2, self.d_count = self._merge(self.d_mean, self.d_m2, self.d_count, mean, m2, n)

  @torch.no_grad()
  def sync(self) -> None:
    """Fold the pending local statistics of ALL ranks into the shared ones (Chan et al. pairwise merge via sums)."""
    n, mean, m2 = self.d_count.clone(), self.d_mean.clone(), self.d_m2.clone()
    if _world() > 1:
      tot = n.clone(); dist.all_reduce(tot)
      gmean = mean * n; dist.all_reduce(gmean); gmean /= torch.clamp(tot, min=1.0)
      gm2 = m2 + n * (mean - gmean) ** 2; dist.all_reduce(gm2)
      n, mean, m2 = tot, gmean, gm2
    self.mean, self.m2, self.count = self._merge(self.mean, self.m2, self.count, mean, m2, n)
    self.d_mean.zero_(); self.d_m2.zero_(); self.d_count.zero_()

  def forward(self, x: torch.Tensor) -> torch.Tensor:
    # unit variance until the first statistics arrive (rsl_rl's normaliser starts from mean 0, var 1)
    var = torch.where(self.count > 0, self.m2 / torch.clamp(self.count, min=1.0), torch.ones_like(self.m2))
    return (x - self.mean) / (torch.sqrt(var) + self.eps)


def mlp(inp: int, hidden, out: int) -> nn.Sequential:
  layers, d = [], inp
  for h in hidden:
    layers += [nn.Linear(d, h), nn.ELU()]
    d = h
  layers.append(nn.Linear(d, out))
  return nn.Sequential(*layers)


class ActorCritic(nn.Module):
  def __init__(self, num_obs: int, num_critic_obs: int, num_actions: int, hidden=(512, 256, 128), init_noise_std: float = 1.0):
    super().__init__()
    self.actor = mlp(num_obs, hidden, num_actions)
    self.critic = mlp(num_critic_obs, hidden, 1)
    self.std = nn.Parameter(init_noise_std * torch.ones(num_actions))
    self.obs_norm = EmpiricalNormalization(num_obs)
    self.critic_norm = EmpiricalNormalization(num_critic_obs)

  def dist(self, obs):
    mean = self.actor(self.obs_norm(obs))
    return torch.distributions.Normal(mean, self.std.expand_as(mean).clamp(min=1e-6))

  def value(self, critic_obs):
    return self.critic(self.critic_norm(critic_obs)).squeeze(-1)


class PPO:
  def __init__(self, env, device, steps_per_env: int = 24, lr: float = 1e-3, gamma: float = 0.99, lam: float = 0.95, clip: float = 0.2,
               entropy_coef: float = 0.01, value_coef: float = 1.0, epochs: int = 5, mini_batches: int = 4, desired_kl: float = 0.01,
               max_grad_norm: float = 1.0, seed: int = 0):
    self.env, self.device, self.T = env, device, steps_per_env
    self.gamma, self.lam, self.clip, self.entropy_coef, self.value_coef = gamma, lam, clip, entropy_coef, value_coef
    self.epochs, self.mini_batches, self.desired_kl, self.max_grad_norm, self.lr = epochs, mini_batches, desired_kl, max_grad_norm, lr
    torch.manual_seed(seed)            # same seed on every rank: identical initial weights, no broadcast needed
    self.ac = ActorCritic(env.num_obs, env.num_privileged_obs, env.num_actions).to(device)
    self.opt = torch.optim.Adam(self.ac.parameters(), lr=lr)
    W, T = env.num_envs, steps_per_env
    z = lambda *s: torch.zeros(*s, device=device)      # noqa: E731
    self.b_obs, self.b_cobs = z(T, W, env.num_obs), z(T, W, env.num_privileged_obs)
    self.b_act, self.b_mu, self.b_sigma = z(T, W, env.num_actions), z(T, W, env.num_actions), z(T, W, env.num_actions)
    self.b_logp, self.b_val, self.b_rew, self.b_done = z(T, W), z(T, W), z(T, W), z(T, W)
    self.ep_ret, self.ep_len = z(W), z(W)
    self.stats = dict(ret_sum=z(()), len_sum=z(()), episodes=z(()))

  @torch.no_grad()
  def rollout(self):
    env, ac = self.env, self.ac
    obs, extras = env.get_observations()
    cobs = extras["observations"]["critic"]
    for t in range(self.T):
      ac.obs_norm.update(obs); ac.critic_norm.update(cobs)
      d = ac.dist(obs)
      act = d.sample()
      self.b_obs[t].copy_(obs); self.b_cobs[t].copy_(cobs); self.b_act[t].copy_(act)
      self.b_mu[t].copy_(d.mean); self.b_sigma[t].copy_(d.stddev)
      self.b_logp[t].copy_(d.log_prob(act).sum(-1)); self.b_val[t].copy_(ac.value(cobs))
      obs, rew, dones, extras = env.step(act)
      cobs = extras["observations"]["critic"]
      donef = dones.float()
      # bootstrap through time-outs (rsl_rl: rewards += gamma * value * time_outs)
      self.b_rew[t].copy_(rew + self.gamma * self.b_val[t] * extras["time_outs"].float())
      self.b_done[t].copy_(donef)
      self.ep_ret += rew; self.ep_len += 1
      self.stats["ret_sum"] += (self.ep_ret * donef).sum(); self.stats["len_sum"] += (self.ep_len * donef).sum()
      self.stats["episodes"] += donef.sum()
      self.ep_ret *= 1 - donef; self.ep_len *= 1 - donef
    last_val = ac.value(cobs)
    adv = torch.zeros_like(self.b_rew)
    gae = torch.zeros_like(last_val)
    for t in reversed(range(self.T)):
      nv = last_val if t == self.T - 1 else self.b_val[t + 1]
      nonterm = 1.0 - self.b_done[t]
      delta = self.b_rew[t] + self.gamma * nv * nonterm - self.b_val[t]
      gae = delta + self.gamma * self.lam * nonterm * gae
      adv[t] = gae
    self.b_ret = adv + self.b_val
    # advantage normalisation over the GLOBAL batch
    s = torch.stack([adv.sum(), (adv * adv).sum(), torch.tensor(float(adv.numel()), device=adv.device)])
    if _world() > 1:
      dist.all_reduce(s)
    mean = s[0] / s[2]
    std = torch.sqrt(torch.clamp(s[1] / s[2] - mean * mean, min=0.0))
    self.b_adv = (adv - mean) / (std + 1e-8)
    ac.obs_norm.sync(); ac.critic_norm.sync()

  def update(self):
    ac = self.ac
    N = self.T * self.env.num_envs
    flat = lambda t: t.reshape(N, *t.shape[2:])      # noqa: E731
    obs, cobs, act, mu0, sig0 = flat(self.b_obs), flat(self.b_cobs), flat(self.b_act), flat(self.b_mu), flat(self.b_sigma)
    logp0, val0, ret, adv = flat(self.b_logp), flat(self.b_val), flat(self.b_ret), flat(self.b_adv)
    mb = N // self.mini_batches
    loss_v = loss_s = torch.zeros((), device=self.device)
    for _ in range(self.epochs):
      perm = torch.randperm(N, device=self.device)
      for k in range(self.mini_batches):
        idx = perm[k * mb:(k + 1) * mb]
        d = ac.dist(obs[idx])
        logp = d.log_prob(act[idx]).sum(-1)
        val = ac.value(cobs[idx])
        with torch.no_grad():      # adaptive learning rate on the (all-rank) mean KL, rsl_rl schedule="adaptive"
          kl = (torch.log(d.stddev / sig0[idx] + 1e-5) + (sig0[idx] ** 2 + (mu0[idx] - d.mean) ** 2) / (2.0 * d.stddev ** 2) - 0.5).sum(-1).mean()
          kl = all_reduce_mean_(kl.clone())
          up = (kl > 2.0 * self.desired_kl).float(); down = ((kl < 0.5 * self.desired_kl) & (kl > 0.0)).float()
          self.lr = float(min(max(self.lr * (1.0 + 0.5 * down.item() - (1.0 / 3.0) * up.item()), 1e-5), 1e-2))
          for g in self.opt.param_groups:
            g["lr"] = self.lr
        ratio = torch.exp(logp - logp0[idx])
        loss_s = torch.max(-adv[idx] * ratio, -adv[idx] * ratio.clamp(1.0 - self.clip, 1.0 + self.clip)).mean()
        vclip = val0[idx] + (val - val0[idx]).clamp(-self.clip, self.clip)
        loss_v = torch.max((val - ret[idx]) ** 2, (vclip - ret[idx]) ** 2).mean()
        loss = loss_s + self.value_coef * loss_v - self.entropy_coef * d.entropy().sum(-1).mean()
        self.opt.zero_grad(set_to_none=True)
        loss.backward()
        if _world() > 1:       # one bucket: ~0.4 M parameters
          grads = [p.grad for p in ac.parameters() if p.grad is not None]
          flat_g = torch.cat([g.reshape(-1) for g in grads])
          all_reduce_mean_(flat_g)
          o = 0
          for g in grads:
            g.copy_(flat_g[o:o + g.numel()].view_as(g)); o += g.numel()
        nn.utils.clip_grad_norm_(ac.parameters(), self.max_grad_norm)
        self.opt.step()
    return dict(value_loss=float(loss_v.detach()), surrogate_loss=float(loss_s.detach()), lr=self.lr, noise_std=float(ac.std.detach().mean()))

  def episode_stats(self):
    s = torch.stack([self.stats["ret_sum"], self.stats["len_sum"], self.stats["episodes"]])
    if _world() > 1:
      dist.all_reduce(s)
    for v in self.stats.values():
      v.zero_()
    n = max(float(s[2]), 1.0)
    return dict(mean_return=float(s[0]) / n, mean_length=float(s[1]) / n, episodes=float(s[2]))
