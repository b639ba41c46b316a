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
    """In place and free of host values, so that it can sit inside a CUDA graph (the graph keeps the buffers' storage)."""
    mean = x.mean(dim=0)
    m2 = ((x - mean) ** 2).sum(dim=0)
    new = self._merge(self.d_mean, self.d_m2, self.d_count, mean, m2, float(x.shape[0]))
    self.d_mean.copy_(new[0]); self.d_m2.copy_(new[1]); self.d_count.copy_(new[2])

  @torch.no_grad()
  def sync(self) -> None:
    """Fold the pending local statistics of ALL ranks into the shared ones (Chan et al. pairwise merge via sums)."""
    n, mean, m2 = self.d_count.clone(), self.d_mean.clone(), self.d_m2.clone()
    if _world() > 1:
      tot = n.clone(); dist.all_reduce(tot)
      gmean = mean * n; dist.all_reduce(gmean); gmean /= torch.clamp(tot, min=1.0)
      gm2 = m2 + n * (mean - gmean) ** 2; dist.all_reduce(gm2)
      n, mean, m2 = tot, gmean, gm2
    new = self._merge(self.mean, self.m2, self.count, mean, m2, n)
    self.mean.copy_(new[0]); self.m2.copy_(new[1]); self.count.copy_(new[2])
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
    return torch.distributions.Normal(mean, self.std.expand_as(mean).clamp(min=1e-6), validate_args=False)

  def value(self, critic_obs):
    return self.critic(self.critic_norm(critic_obs)).squeeze(-1)


def init_policy(seed: int, num_obs: int, num_critic_obs: int, num_actions: int, device) -> ActorCritic:
  """Identical initial weights on every rank (no broadcast needed) WITHOUT touching the process's random streams: the
  env seeded them with seed + rank (reference utils/random.py:9-26), and commands, reset poses, observation noise, pushes
  and the action noise must keep differing between ranks.  The network is initialised under a forked generator state."""
  device = torch.device(device)
  with torch.random.fork_rng(devices=[device] if device.type == "cuda" else []):
    torch.manual_seed(seed)
    ac = ActorCritic(num_obs, num_critic_obs, num_actions)
  return ac.to(device)


class PPO:
  def __init__(self, env, device, steps_per_env: int = 24, lr: float = 1e-3, gamma: float = 0.99, lam: float = 0.95, clip: float = 0.2,
               entropy_coef: float = 0.01, value_coef: float = 1.0, epochs: int = 5, mini_batches: int = 4, desired_kl: float = 0.01,
               max_grad_norm: float = 1.0, seed: int = 0, use_graph: bool = True, tf32: bool = True):
    self.env, self.device, self.T = env, device, steps_per_env
    if tf32:      # the policy MLPs are library GEMMs; TF32 tensor cores for them (the physics stays fp32)
      torch.backends.cuda.matmul.allow_tf32 = True
      torch.backends.cudnn.allow_tf32 = True
    self.gamma, self.lam, self.clip, self.entropy_coef, self.value_coef = gamma, lam, clip, entropy_coef, value_coef
    self.epochs, self.mini_batches, self.desired_kl, self.max_grad_norm, self.lr = epochs, mini_batches, desired_kl, max_grad_norm, lr
    self.ac = init_policy(seed, env.num_obs, env.num_privileged_obs, env.num_actions, device)
    # learning rate lives on the device: the adaptive-KL schedule then needs no host synchronisation per mini-batch
    self.lr_t = torch.tensor(float(lr), device=device)
    self.opt = torch.optim.Adam(self.ac.parameters(), lr=self.lr_t, capturable=True)
    W, T = env.num_envs, steps_per_env
    z = lambda *s: torch.zeros(*s, device=device)      # noqa: E731
    self.b_obs, self.b_cobs = z(T, W, env.num_obs), z(T, W, env.num_privileged_obs)
    self.b_act, self.b_mu, self.b_sigma = z(T, W, env.num_actions), z(T, W, env.num_actions), z(T, W, env.num_actions)
    self.b_logp, self.b_val, self.b_rew, self.b_done = z(T, W), z(T, W), z(T, W), z(T, W)
    self.ep_ret, self.ep_len = z(W), z(W)
    self.stats = dict(ret_sum=z(()), len_sum=z(()), episodes=z(()))
    self.b_ret, self.b_adv = z(T, W), z(T, W)
    self.loss_s, self.loss_v, self.kl = z(()), z(()), z(())
    self._params = [p for p in self.ac.parameters()]
    self._flat, self.flat_g = None, None
    self.use_graph = bool(use_graph and device.type == "cuda")
    self.t_idx = torch.zeros(1, dtype=torch.long, device=device)
    self.step_graph = None
    if self.use_graph:
      self._capture()

  @torch.no_grad()
  def _rollout_step(self):
    """One policy step + env step + bookkeeping, all in place on static tensors (row `t_idx` of the rollout buffers)."""
    env, ac, t = self.env, self.ac, self.t_idx
    obs, cobs = env.obs, env.critic_obs
    ac.obs_norm.update(obs); ac.critic_norm.update(cobs)
    d = ac.dist(obs)
    act = d.mean + d.stddev * torch.randn_like(d.mean)      # Normal.sample() checks std on the host: not capturable
    val = ac.value(cobs)
    put = lambda buf, x: buf.index_copy_(0, t, x.unsqueeze(0))      # noqa: E731
    put(self.b_obs, obs); put(self.b_cobs, cobs); put(self.b_act, act); put(self.b_mu, d.mean); put(self.b_sigma, d.stddev)
    put(self.b_logp, d.log_prob(act).sum(-1)); put(self.b_val, val)
    env.actions_in.copy_(act)
    env._step_impl()
    rew, donef = env.rew, env.dones.float()
    # bootstrap through time-outs (rsl_rl: rewards += gamma * value * time_outs)
    put(self.b_rew, rew + self.gamma * val * env.time_outs.float()); put(self.b_done, donef)
    self.ep_ret += rew; self.ep_len += 1
    self.stats["ret_sum"] += (self.ep_ret * donef).sum(); self.stats["len_sum"] += (self.ep_len * donef).sum()
    self.stats["episodes"] += donef.sum()
    self.ep_ret *= 1 - donef; self.ep_len *= 1 - donef
    t += 1

  def _capture(self):
    """Policy inference, the whole env step and the rollout bookkeeping as ONE CUDA graph per rollout step."""
    dev = self.device
    snap = [b.clone() for b in self.ac.buffers()]
    side = torch.cuda.Stream(device=dev)
    side.wait_stream(torch.cuda.current_stream(dev))
    with torch.cuda.stream(side):
      for _ in range(3):
        self.t_idx.zero_()
        self._rollout_step()
    torch.cuda.current_stream(dev).wait_stream(side)
    torch.cuda.synchronize(dev)
    self.step_graph = torch.cuda.CUDAGraph()
    self.t_idx.zero_()
    with torch.cuda.graph(self.step_graph):
      self._rollout_step()
    # undo the warm-up's side effects: normaliser statistics, episode statistics, env state
    for b, v in zip(self.ac.buffers(), snap):
      b.copy_(v)
    for v in (self.ep_ret, self.ep_len, *self.stats.values()):
      v.zero_()
    self.env._reset_all()
    for v in self.env.term_sums.values():
      v.zero_()

  @torch.no_grad()
  def rollout(self):
    env, ac = self.env, self.ac
    self.t_idx.zero_()
    for _ in range(self.T):
      if self.step_graph is not None:
        self.step_graph.replay()
      else:
        self._rollout_step()
    cobs = env.critic_obs
    last_val = ac.value(cobs)
    adv = torch.zeros_like(self.b_rew)
    gae = torch.zeros_like(last_val)
    for t in reversed(range(self.T)):
      nv = last_val if t == self.T - 1 else self.b_val[t + 1]
      nonterm = 1.0 - self.b_done[t]
      delta = self.b_rew[t] + self.gamma * nv * nonterm - self.b_val[t]
      gae = delta + self.gamma * self.lam * nonterm * gae
      adv[t] = gae
    self.b_ret.copy_(adv + self.b_val)
    # advantage normalisation over the GLOBAL batch
    s = torch.stack([adv.sum(), (adv * adv).sum(), torch.tensor(float(adv.numel()), device=adv.device)])
    if _world() > 1:
      dist.all_reduce(s)
    mean = s[0] / s[2]
    std = torch.sqrt(torch.clamp(s[1] / s[2] - mean * mean, min=0.0))
    self.b_adv.copy_((adv - mean) / (std + 1e-8))
    ac.obs_norm.sync(); ac.critic_norm.sync()

  # ---- update: one mini-batch = [gather, forward, losses, backward, KL, (flatten)] -> (all-reduce) -> [lr, clip, Adam]
  def _mb_backward(self):
    """Loss and gradients of the mini-batch whose sample indices sit in `self.mb_idx`; with several ranks the gradients and
    the KL are also packed into one bucket for the single collective of the mini-batch."""
    ac, idx = self.ac, self.mb_idx
    obs, cobs, act, mu0, sig0, logp0, val0, ret, adv = (t[idx] for t in self._flat)
    d = ac.dist(obs)
    logp = d.log_prob(act).sum(-1)
    val = ac.value(cobs)
    with torch.no_grad():
      kl = (torch.log(d.stddev / sig0 + 1e-5) + (sig0 ** 2 + (mu0 - d.mean) ** 2) / (2.0 * d.stddev ** 2) - 0.5).sum(-1).mean()
    ratio = torch.exp(logp - logp0)
    loss_s = torch.max(-adv * ratio, -adv * ratio.clamp(1.0 - self.clip, 1.0 + self.clip)).mean()
    vclip = val0 + (val - val0).clamp(-self.clip, self.clip)
    loss_v = torch.max((val - ret) ** 2, (vclip - ret) ** 2).mean()
    loss = loss_s + self.value_coef * loss_v - self.entropy_coef * d.entropy().sum(-1).mean()
    self.opt.zero_grad(set_to_none=True)
    loss.backward()
    with torch.no_grad():
      self.loss_s.copy_(loss_s); self.loss_v.copy_(loss_v); self.kl.copy_(kl)
      if _world() > 1:
        self.flat_g = torch.cat([p.grad.reshape(-1) for p in self._params] + [kl.reshape(1)])

  @torch.no_grad()
  def _mb_apply(self):
    """Adaptive learning rate on the (all-rank) mean KL (rsl_rl schedule="adaptive"), gradient clipping, Adam step."""
    kl = self.kl
    if _world() > 1:
      o = 0
      for p in self._params:
        p.grad.copy_(self.flat_g[o:o + p.numel()].view_as(p)); o += p.numel()
      kl = self.flat_g[o]
    up = (kl > 2.0 * self.desired_kl).float(); down = ((kl < 0.5 * self.desired_kl) & (kl > 0.0)).float()
    self.lr_t.copy_((self.lr_t * (1.0 + 0.5 * down - (1.0 / 3.0) * up)).clamp(1e-5, 1e-2))
    nn.utils.clip_grad_norm_(self._params, self.max_grad_norm)
    self.opt.step()

  def _capture_update(self):
    """The two halves of a mini-batch as CUDA graphs sharing one memory pool (one graph when there is a single rank);
    the NCCL all-reduce between them stays an ordinary stream-ordered call."""
    dev = self.device
    state = [p.detach().clone() for p in self._params]
    lr0 = self.lr_t.clone()
    side = torch.cuda.Stream(device=dev)
    side.wait_stream(torch.cuda.current_stream(dev))
    with torch.cuda.stream(side):
      for _ in range(3):
        self._mb_backward()
        if _world() > 1:
          all_reduce_mean_(self.flat_g)
        self._mb_apply()
    torch.cuda.current_stream(dev).wait_stream(side)
    torch.cuda.synchronize(dev)
    self.g_back = torch.cuda.CUDAGraph()
    if _world() > 1:
      with torch.cuda.graph(self.g_back):
        self._mb_backward()
      self.g_apply = torch.cuda.CUDAGraph()
      with torch.cuda.graph(self.g_apply, pool=self.g_back.pool()):
        self._mb_apply()
    else:
      with torch.cuda.graph(self.g_back):
        self._mb_backward()
        self._mb_apply()
      self.g_apply = None
    # undo the warm-up steps: weights, optimiser moments, learning rate
    with torch.no_grad():
      for p, v in zip(self._params, state):
        p.copy_(v)
      for st in self.opt.state.values():
        for v in st.values():
          if torch.is_tensor(v):
            v.zero_()
      self.lr_t.copy_(lr0)

  def update(self):
    ac = self.ac
    N = self.T * self.env.num_envs
    flat = lambda t: t.reshape(N, *t.shape[2:])      # noqa: E731
    mb = N // self.mini_batches
    if self._flat is None:
      # static views / buffers the mini-batch graphs read (b_ret / b_adv are refilled in place by rollout())
      self._flat = tuple(flat(t) for t in (self.b_obs, self.b_cobs, self.b_act, self.b_mu, self.b_sigma, self.b_logp, self.b_val, self.b_ret, self.b_adv))
      self.mb_idx = torch.zeros(mb, dtype=torch.long, device=self.device)
      if self.use_graph:
        self._capture_update()
    for _ in range(self.epochs):
      perm = torch.randperm(N, device=self.device)
      for k in range(self.mini_batches):
        self.mb_idx.copy_(perm[k * mb:(k + 1) * mb])
        if self.use_graph:
          self.g_back.replay()
          if self.g_apply is not None:
            all_reduce_mean_(self.flat_g)      # ONE collective per mini-batch: gradients (~0.4 M parameters) + KL
            self.g_apply.replay()
        else:
          self._mb_backward()
          if _world() > 1:
            all_reduce_mean_(self.flat_g)
          self._mb_apply()
    self.lr = float(self.lr_t)
    return dict(value_loss=float(self.loss_v), surrogate_loss=float(self.loss_s), lr=self.lr, noise_std=float(ac.std.detach().mean()))

  def episode_stats(self):
    s = torch.stack([self.stats["ret_sum"], self.stats["len_sum"], self.stats["episodes"]])
    if _world() > 1:
      dist.all_reduce(s)
    for v in self.stats.values():
      v.zero_()
    n = max(float(s[2]), 1.0)
    return dict(mean_return=float(s[0]) / n, mean_length=float(s[1]) / n, episodes=float(s[2]))
