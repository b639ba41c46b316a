"""Mjlab-Tracking-Flat-Unitree-G1 (BASELINE.json configs[3]) as a vectorised env running directly on the engine.

Restates `create_tracking_env_cfg` (src/mjlab/tasks/tracking/tracking_env_cfg.py:66-332) with the G1 parameters of
tasks/tracking/config/g1/env_cfgs.py:19-82: the motion command with adaptive / uniform / start sampling
(tasks/tracking/mdp/commands.py:69-420), its rewards (mdp/rewards.py), terminations (mdp/terminations.py) and
observations (mdp/observations.py), in the step order of ManagerBasedRlEnv.step (envs/manager_based_rl_env.py:111-152).
The whole env step -- 4 physics sub-steps, manager terms, masked resets, forward() -- is captured in ONE CUDA graph:
host branches of the reference (`if torch.any(...)`, `env_ids = nonzero()`, `torch.multinomial`) are restated as masked,
fixed-shape tensor ops (failure histogram by scatter-add, inverse-CDF sampling), so there is no host synchronisation.

The reference loads the motion from a W&B artifact; `synthetic_motion()` builds the stand-in SURVEY.md §8d prescribes
(T = 500 frames at 50 fps of the knees-bent pose + 0.1 sin(2 pi 0.5 t + phi_j) per joint, body arrays from the engine's
own forward() at every frame, exactly how scripts/csv_to_npz.py:247-276 produces them) with the npz keys of
scripts/csv_to_npz.py:206-211.
"""
from __future__ import annotations

import math

import numpy as np
import torch

from .. import engine, scenes
from ..hostmodel import HostData
from ..sim import MujocoCfg, Simulation, SimulationCfg
from . import fused
from . import velocity_mdp as mdp

BODY_NAMES = ("pelvis", "left_hip_roll_link", "left_knee_link", "left_ankle_roll_link", "right_hip_roll_link", "right_knee_link",
              "right_ankle_roll_link", "torso_link", "left_shoulder_roll_link", "left_elbow_link", "left_wrist_yaw_link",
              "right_shoulder_roll_link", "right_elbow_link", "right_wrist_yaw_link")
EE_BODY_NAMES = ("left_ankle_roll_link", "right_ankle_roll_link", "left_wrist_yaw_link", "right_wrist_yaw_link")
ANCHOR = "torso_link"
POSE_RANGE = dict(x=(-0.05, 0.05), y=(-0.05, 0.05), z=(-0.01, 0.01), roll=(-0.1, 0.1), pitch=(-0.1, 0.1), yaw=(-0.2, 0.2))
VELOCITY_RANGE = dict(x=(-0.5, 0.5), y=(-0.5, 0.5), z=(-0.2, 0.2), roll=(-0.52, 0.52), pitch=(-0.52, 0.52), yaw=(-0.78, 0.78))
JOINT_POSITION_RANGE = (-0.1, 0.1)
REWARD_WEIGHTS = dict(motion_global_root_pos=0.5, motion_global_root_ori=0.5, motion_body_pos=1.0, motion_body_ori=1.0,
                      motion_body_lin_vel=1.0, motion_body_ang_vel=1.0, action_rate_l2=-1e-1, joint_limit=-10.0, self_collisions=-10.0)
REWARD_STD = dict(motion_global_root_pos=0.3, motion_global_root_ori=0.4, motion_body_pos=0.3, motion_body_ori=0.4,
                  motion_body_lin_vel=1.0, motion_body_ang_vel=3.14)
_AXES = ("x", "y", "z", "roll", "pitch", "yaw")


# ------------------------------------------------------------------------------------------------ quaternion helpers
def quat_conj(q):
  return torch.cat([q[..., :1], -q[..., 1:]], dim=-1)


def quat_inv(q):
  return quat_conj(q) / q.pow(2).sum(dim=-1, keepdim=True).clamp(min=1e-9)


def yaw_quat(q):
  w, x, y, z = q.unbind(-1)
  yaw = torch.atan2(2 * (w * z + x * y), 1 - 2 * (y * y + z * z))
  o = torch.zeros_like(q)
  o[..., 0] = torch.cos(yaw / 2); o[..., 3] = torch.sin(yaw / 2)
  return o / o.norm(dim=-1, keepdim=True).clamp(min=1e-9)


def quat_from_euler_xyz(roll, pitch, yaw):
  cy, sy, cr, sr, cp, sp = torch.cos(yaw * 0.5), torch.sin(yaw * 0.5), torch.cos(roll * 0.5), torch.sin(roll * 0.5), torch.cos(pitch * 0.5), torch.sin(pitch * 0.5)
  return torch.stack([cy * cr * cp + sy * sr * sp, cy * sr * cp - sy * cr * sp, cy * cr * sp + sy * sr * cp, sy * cr * cp - cy * sr * sp], dim=-1)


def quat_error_magnitude(q1, q2):
  """|axis-angle of q1 * conj(q2)| (isaaclab utils/math.py quat_error_magnitude -> quat_box_minus -> axis_angle_from_quat)."""
  q = mdp.quat_mul(q1, quat_conj(q2))
  q = q * (1.0 - 2.0 * (q[..., 0:1] < 0.0))
  mag = torch.linalg.norm(q[..., 1:], dim=-1)
  half = torch.atan2(mag, q[..., 0])
  angle = 2.0 * half
  s = torch.where(angle.abs() > 1e-6, torch.sin(half) / angle, 0.5 - angle * angle / 48)
  return torch.norm(q[..., 1:4] / s.unsqueeze(-1), dim=-1)


def matrix_from_quat(q):
  r, i, j, k = q.unbind(-1)
  two_s = 2.0 / (q * q).sum(-1)
  o = torch.stack([1 - two_s * (j * j + k * k), two_s * (i * j - k * r), two_s * (i * k + j * r),
                   two_s * (i * j + k * r), 1 - two_s * (i * i + k * k), two_s * (j * k - i * r),
                   two_s * (i * k - j * r), two_s * (j * k + i * r), 1 - two_s * (i * i + j * j)], dim=-1)
  return o.reshape(q.shape[:-1] + (3, 3))


def subtract_frame_transforms(t01, q01, t02, q02):
  q10 = quat_inv(q01)
  return mdp.quat_apply(q10, t02 - t01), mdp.quat_mul(q10, q02)


# ------------------------------------------------------------------------------------------------ synthetic motion
def synthetic_motion(m=None, T: int = 500, fps: float = 50.0, device: str = "cuda:0", seed: int = 0) -> dict:
  """npz-shaped dict (numpy): fps [1], joint_pos/joint_vel [T, nu], body_pos_w/body_quat_w/body_lin_vel_w/body_ang_vel_w
  [T, nrobotbody, 3|4] -- robot bodies only, in model order (scripts/csv_to_npz.py:206-211, :267-276)."""
  m = m or scenes.load_asset("g1_tracking")
  key = m.keys["robot/init_state"]
  rng = np.random.default_rng(seed)
  phi = rng.uniform(0.0, 2.0 * np.pi, size=m.nu)
  t = np.arange(T)[:, None] / fps
  q0 = np.asarray(key["qpos"], np.float64)
  jp = q0[None, 7:] + 0.1 * np.sin(2.0 * np.pi * 0.5 * t + phi[None, :])
  jv = 0.1 * 2.0 * np.pi * 0.5 * np.cos(2.0 * np.pi * 0.5 * t + phi[None, :])
  lo, hi = np.asarray(m.jnt_range)[1:, 0], np.asarray(m.jnt_range)[1:, 1]
  jp = np.clip(jp, lo, hi)
  M = engine.put_model(m, device)
  hd = HostData(m)
  hd.qpos[:] = q0
  D = engine.put_data(m, hd, T, 35, 250, model=M)
  D.qpos[:, 7:] = torch.as_tensor(jp, dtype=torch.float32, device=device)
  D.qvel[:, 6:] = torch.as_tensor(jv, dtype=torch.float32, device=device)
  engine.forward(M, D)
  robot = [i for i, n in enumerate(m.names["body"]) if n.startswith("robot/")]
  xpos, xquat, cvel, com = D.xpos[:, robot], D.xquat[:, robot], D.cvel[:, robot], D.subtree_com[:, [int(m.body_rootid[b]) for b in robot]]
  lin = cvel[..., 3:6] - torch.cross(cvel[..., 0:3], com - xpos, dim=-1)        # entity/data.py:20-31
  out = dict(fps=np.array([fps]), joint_pos=jp.astype(np.float32), joint_vel=jv.astype(np.float32),
             body_pos_w=xpos.cpu().numpy(), body_quat_w=xquat.cpu().numpy(), body_lin_vel_w=lin.cpu().numpy(),
             body_ang_vel_w=cvel[..., 0:3].cpu().numpy())
  del D, M
  return out


class G1TrackingFlatEnv:
  decimation = 4
  episode_length_s = 10.0

  def __init__(self, num_envs: int, device: str = "cuda:0", motion: dict | str | None = None, seed: int = 0, env_spacing: float = 2.5,
               use_graph: bool = True, use_fused: bool = True, sampling_mode: str = "adaptive", adaptive_kernel_size: int = 1, adaptive_lambda: float = 0.8,
               adaptive_uniform_ratio: float = 0.1, adaptive_alpha: float = 0.001):
    assert sampling_mode in ("adaptive", "uniform", "start")
    self.num_envs, self.device, self.sampling_mode = num_envs, torch.device(device), sampling_mode
    self.use_fused = use_fused
    self.gen = None              # tests: a per-env torch.Generator so that two envs draw the same numbers (eager mode only)
    self.external_U = False      # tests fill U (observation noise) themselves
    torch.cuda.set_device(self.device)
    m = scenes.load_asset("g1_tracking")
    self.m = m
    cfg = SimulationCfg(nconmax=35, njmax=250, ls_parallel=True, mujoco=MujocoCfg(timestep=0.005, iterations=10, ls_iterations=20))
    self.sim = Simulation(num_envs, cfg, m, str(self.device))
    self.M, self.D = self.sim.model, self.sim.data
    self.step_dt = 0.005 * self.decimation
    self.max_episode_length = math.ceil(self.episode_length_s / self.step_dt)
    dev, W = self.device, num_envs
    f32 = dict(dtype=torch.float32, device=dev)
    g = torch.Generator(device="cpu").manual_seed(seed)
    # ---- indices
    bn = m.names["body"]
    robot = [i for i, n in enumerate(bn) if n.startswith("robot/")]
    self.body_ids = torch.tensor([bn.index("robot/" + n) for n in BODY_NAMES], device=dev)             # model body ids
    self.body_root = torch.tensor([int(m.body_rootid[int(b)]) for b in self.body_ids], device=dev)
    self.motion_body_ids = torch.tensor([robot.index(bn.index("robot/" + n)) for n in BODY_NAMES], device=dev)   # rows of the npz body arrays
    self.anchor_k = BODY_NAMES.index(ANCHOR)
    self.ee_k = torch.tensor([BODY_NAMES.index(n) for n in EE_BODY_NAMES], device=dev)
    self.nb = len(BODY_NAMES)
    sa = {n: int(a) for n, a in zip(m.names["sensor"], m.sensor_adr)}
    self.s_gyro, self.s_linvel, self.s_self = sa["robot/imu_ang_vel"], sa["robot/imu_lin_vel"], sa["self_collision_pelvis_found"]
    jn = [n.split("/")[-1] for n in m.names["joint"][1:]]
    self.num_actions = m.nu
    key = m.keys["robot/init_state"]
    self.default_joint_pos = torch.tensor(key["qpos"][7:], **f32)
    self.action_scale = torch.tensor(scenes.g1_action_scale(jn), **f32)
    rng = torch.tensor(np.asarray(m.jnt_range)[1:], **f32)
    mid, half = 0.5 * (rng[:, 0] + rng[:, 1]), 0.5 * (rng[:, 1] - rng[:, 0])
    self.soft_limits = torch.stack([mid - 0.9 * half, mid + 0.9 * half], dim=-1)
    self.env_origins = torch.tensor(scenes.env_origins_grid(W, env_spacing), **f32)
    self.gravity_dir = torch.tensor([0.0, 0.0, -1.0], **f32)
    # ---- startup domain randomisation (tracking_env_cfg.py:222-262): torso COM, joint zero offsets, foot friction
    self.sim.expand_model_fields(("geom_friction", "body_ipos", "qpos0"))
    foot = [i for i, n in enumerate(m.names["geom"]) if "foot" in n and n.endswith("_collision")]
    self.M.geom_friction[:, foot, 0] = (0.3 + 0.9 * torch.rand(W, len(foot), generator=g)).to(dev)
    torso = bn.index("robot/torso_link")
    span = torch.tensor([0.025, 0.05, 0.05])
    self.M.body_ipos[:, torso] += ((2.0 * torch.rand(W, 3, generator=g) - 1.0) * span).to(dev)
    self.M.qpos0[:, 7:] += (0.01 * (2.0 * torch.rand(W, m.nq - 7, generator=g) - 1.0)).to(dev)
    # ---- motion
    if motion is None:
      motion = synthetic_motion(m, device=str(dev))
    elif isinstance(motion, str):
      motion = dict(np.load(motion))
    mt = {k: torch.as_tensor(np.asarray(v), **f32) for k, v in motion.items() if k != "fps"}
    self.mo_jp, self.mo_jv = mt["joint_pos"], mt["joint_vel"]
    self.mo_bpos, self.mo_bquat = mt["body_pos_w"][:, self.motion_body_ids].contiguous(), mt["body_quat_w"][:, self.motion_body_ids].contiguous()
    self.mo_blin, self.mo_bang = mt["body_lin_vel_w"][:, self.motion_body_ids].contiguous(), mt["body_ang_vel_w"][:, self.motion_body_ids].contiguous()
    self.T = int(self.mo_jp.shape[0])
    # ---- command state
    self.time_steps = torch.zeros(W, dtype=torch.long, device=dev)
    self.body_pos_relative_w = torch.zeros(W, self.nb, 3, **f32)
    self.body_quat_relative_w = torch.zeros(W, self.nb, 4, **f32); self.body_quat_relative_w[..., 0] = 1.0
    self.bin_count = int(self.T // (1 / self.step_dt)) + 1
    self.bin_failed_count = torch.zeros(self.bin_count, **f32)
    self.current_bin_failed = torch.zeros(self.bin_count, **f32)
    kern = torch.tensor([adaptive_lambda ** i for i in range(adaptive_kernel_size)], **f32)
    self.kernel = kern / kern.sum()
    self.adaptive_uniform_ratio, self.adaptive_alpha, self.adaptive_kernel_size = adaptive_uniform_ratio, adaptive_alpha, adaptive_kernel_size
    self.pose_lo = torch.tensor([POSE_RANGE[k][0] for k in _AXES], **f32); self.pose_hi = torch.tensor([POSE_RANGE[k][1] for k in _AXES], **f32)
    self.vel_lo = torch.tensor([VELOCITY_RANGE[k][0] for k in _AXES], **f32); self.vel_hi = torch.tensor([VELOCITY_RANGE[k][1] for k in _AXES], **f32)
    # ---- buffers
    z = lambda *s: torch.zeros(*s, **f32)      # noqa: E731
    self.actions_in, self.action, self.prev_action = z(W, m.nu), z(W, m.nu), z(W, m.nu)
    self.push_time_left = z(W)
    self.episode_length_buf = torch.zeros(W, dtype=torch.long, device=dev)
    self.num_obs = 2 * m.nu + 3 + 6 + 3 + 3 + 3 * m.nu
    self.num_privileged_obs = self.num_obs + 3 * self.nb + 6 * self.nb
    self.obs, self.critic_obs = z(W, self.num_obs), z(W, self.num_privileged_obs)
    self.rew = z(W)
    self.dones, self.time_outs, self.terminated = (torch.zeros(W, dtype=torch.bool, device=dev) for _ in range(3))
    self.term_sums_buf = z(len(REWARD_WEIGHTS), W)
    self.term_sums = {k: self.term_sums_buf[i] for i, k in enumerate(REWARD_WEIGHTS)}          # B2S_TRK_TERM_* order
    self.metrics_buf = z(6, W)
    self.metrics = {k: self.metrics_buf[i] for i, k in enumerate(("error_anchor_pos", "error_anchor_rot", "error_body_pos", "error_body_rot",
                                                                   "error_joint_pos", "error_joint_vel"))}   # B2S_TRK_MET_* order
    # observation noise: anchor pos 3 | anchor ori 6 | base lin vel 3 | base ang vel 3 | joint pos nu | joint vel nu
    self.nrand = 15 + 2 * m.nu
    self.U = z(W, self.nrand)
    self.ee_mask = torch.tensor([n in EE_BODY_NAMES for n in BODY_NAMES], dtype=torch.bool, device=dev)
    self._params = self._make_params() if use_fused else None
    torch.manual_seed(seed)
    self._reset_all()
    self.graph = None
    if use_graph:
      self._capture()

  # ------------------------------------------------------------------------------------------------ command views
  def _motion(self):
    """Motion frame of every env at its time step: joint pos/vel, body pos (+ env origin) / quat / lin vel / ang vel."""
    ts = self.time_steps
    return (self.mo_jp[ts], self.mo_jv[ts], self.mo_bpos[ts] + self.env_origins[:, None, :], self.mo_bquat[ts], self.mo_blin[ts], self.mo_bang[ts])

  def _robot_bodies(self):
    D = self.D
    pos, quat, cvel = D.xpos[:, self.body_ids], D.xquat[:, self.body_ids], D.cvel[:, self.body_ids]
    lin = cvel[..., 3:6] - torch.cross(cvel[..., 0:3], D.subtree_com[:, self.body_root] - pos, dim=-1)      # entity/data.py:20-31
    return pos, quat, lin, cvel[..., 0:3]

  def _rand(self, *shape):
    return torch.rand(*shape, device=self.device, generator=self.gen)

  def _u(self, lo, hi, *shape):
    return lo + (hi - lo) * self._rand(*shape)

  # ------------------------------------------------------------------------------------------------ command
  def _sample_time_steps(self, mask):
    """MotionCommand._adaptive_sampling / _uniform_sampling / start (commands.py:261-318), for the envs in `mask`."""
    W, dev = self.num_envs, self.device
    if self.sampling_mode == "start":
      new = torch.zeros_like(self.time_steps)
    elif self.sampling_mode == "uniform":
      new = (self._rand(W) * self.T).long().clamp(max=self.T - 1)
    else:
      failed = mask & self.terminated
      cur_bin = torch.clamp((self.time_steps * self.bin_count) // max(self.T, 1), 0, self.bin_count - 1)
      hist = torch.zeros_like(self.current_bin_failed).scatter_add_(0, cur_bin, failed.float())
      # the reference overwrites _current_bin_failed only when some episode failed in this step
      self.current_bin_failed.copy_(torch.where(failed.any(), hist, self.current_bin_failed))
      p = self.bin_failed_count + self.adaptive_uniform_ratio / float(self.bin_count)
      p = torch.nn.functional.pad(p.view(1, 1, -1), (0, self.adaptive_kernel_size - 1), mode="replicate")
      p = torch.nn.functional.conv1d(p, self.kernel.view(1, 1, -1)).view(-1)
      p = p / p.sum()
      cdf = torch.cumsum(p, 0)
      bins = torch.searchsorted(cdf, self._rand(W) * cdf[-1]).clamp(max=self.bin_count - 1)      # multinomial, replacement=True
      new = ((bins + self._rand(W)) / self.bin_count * (self.T - 1)).long()
    self.time_steps.copy_(torch.where(mask, new, self.time_steps))

  def _resample_command(self, mask):
    """MotionCommand._resample_command (commands.py:320-385): new phase, then the robot is written onto the motion frame
    with pose / velocity / joint noise."""
    W, D = self.num_envs, self.D
    self._sample_time_steps(mask)
    jp, jv, bpos, bquat, blin, bang = self._motion()
    pr = self._u(self.pose_lo, self.pose_hi, W, 6)
    root_pos = bpos[:, 0] + pr[:, 0:3]
    root_ori = mdp.quat_mul(quat_from_euler_xyz(pr[:, 3], pr[:, 4], pr[:, 5]), bquat[:, 0])
    vr = self._u(self.vel_lo, self.vel_hi, W, 6)
    root_lin, root_ang = blin[:, 0] + vr[:, :3], bang[:, 0] + vr[:, 3:]
    jp = jp + self._u(JOINT_POSITION_RANGE[0], JOINT_POSITION_RANGE[1], W, self.num_actions)
    jp = torch.clip(jp, self.soft_limits[:, 0], self.soft_limits[:, 1])
    qpos = torch.cat([root_pos, root_ori, jp], dim=1)
    qvel = torch.cat([root_lin, mdp.quat_apply_inverse(root_ori, root_ang), jv], dim=1)       # entity/data.py:95-106
    m1 = mask.unsqueeze(1)
    D.qpos.copy_(torch.where(m1, qpos, D.qpos))
    D.qvel.copy_(torch.where(m1, qvel, D.qvel))
    D.qacc_warmstart.copy_(torch.where(m1, torch.zeros_like(D.qacc_warmstart), D.qacc_warmstart))

  def _update_command(self):
    """MotionCommand._update_command (commands.py:387-420): advance the phase, wrap finished motions, and re-anchor the
    motion's body poses at the robot's anchor (yaw and xy from the robot, height from the motion)."""
    self._advance_command()
    self._reanchor()
    self._adaptive_ema()

  def _advance_command(self):
    self.time_steps += 1
    self._resample_command(self.time_steps >= self.T)

  def _reanchor(self):
    _, _, bpos, bquat, _, _ = self._motion()
    rpos, rquat, _, _ = self._robot_bodies()
    a = self.anchor_k
    anchor_pos, anchor_quat = bpos[:, a:a + 1], bquat[:, a:a + 1]
    delta_pos = torch.cat([rpos[:, a:a + 1, 0:2], anchor_pos[..., 2:3]], dim=-1).expand(-1, self.nb, -1)
    delta_ori = yaw_quat(mdp.quat_mul(rquat[:, a:a + 1], quat_inv(anchor_quat))).expand(-1, self.nb, -1)
    self.body_quat_relative_w.copy_(mdp.quat_mul(delta_ori, bquat))
    self.body_pos_relative_w.copy_(delta_pos + mdp.quat_apply(delta_ori, bpos - anchor_pos))

  def _adaptive_ema(self):
    if self.sampling_mode == "adaptive":
      self.bin_failed_count.copy_(self.adaptive_alpha * self.current_bin_failed + (1 - self.adaptive_alpha) * self.bin_failed_count)
      self.current_bin_failed.zero_()

  # ------------------------------------------------------------------------------------------------ observations
  def _compute_obs(self):
    D, W, nu = self.D, self.num_envs, self.num_actions
    if not self.external_U:
      self.U.uniform_()
    un = lambda t, a, s0: t + (2.0 * self.U[:, s0:s0 + t.shape[1]] - 1.0) * a      # noqa: E731
    jp_m, jv_m, bpos, bquat, _, _ = self._motion()
    rpos, rquat, _, _ = self._robot_bodies()
    a = self.anchor_k
    pos_b, ori_b = subtract_frame_transforms(rpos[:, a], rquat[:, a], bpos[:, a], bquat[:, a])
    ori6 = matrix_from_quat(ori_b)[..., :2].reshape(W, -1)
    bp_b, bo_b = subtract_frame_transforms(rpos[:, a:a + 1].expand(-1, self.nb, -1), rquat[:, a:a + 1].expand(-1, self.nb, -1), rpos, rquat)
    bo6 = matrix_from_quat(bo_b)[..., :2].reshape(W, -1)
    sd = D.sensordata
    lin, gyro = sd[:, self.s_linvel:self.s_linvel + 3], sd[:, self.s_gyro:self.s_gyro + 3]
    jp, jv = D.qpos[:, 7:] - self.default_joint_pos, D.qvel[:, 6:]
    cmd = torch.cat([jp_m, jv_m], dim=1)
    self.obs.copy_(torch.cat([cmd, un(pos_b, 0.25, 0), un(ori6, 0.05, 3), un(lin, 0.5, 9), un(gyro, 0.2, 12), un(jp, 0.01, 15), un(jv, 0.5, 15 + nu), self.action], dim=1))
    self.critic_obs.copy_(torch.cat([cmd, pos_b, ori6, bp_b.reshape(W, -1), bo6, lin, gyro, jp, jv, self.action], dim=1))

  # ------------------------------------------------------------------------------------------------ step
  def _reset_masked(self, mask):
    m1 = mask.unsqueeze(1)
    self._resample_command(mask)                                    # command_manager.reset -> _resample_command
    for t in (self.action, self.prev_action):
      t.copy_(torch.where(m1, torch.zeros_like(t), t))
    self.episode_length_buf.copy_(torch.where(mask, torch.zeros_like(self.episode_length_buf), self.episode_length_buf))
    self.push_time_left.copy_(torch.where(mask, self._u(1.0, 3.0, self.num_envs), self.push_time_left))

  def _reset_all(self):
    self.terminated.zero_()
    self._reset_masked(torch.ones(self.num_envs, dtype=torch.bool, device=self.device))
    engine.forward(self.M, self.D)
    self.time_steps -= 1
    self._update_command()
    self._compute_obs()

  def _make_params(self):
    m, D, W = self.m, self.D, self.num_envs
    p = fused.TrackingMdp()
    p.nworld, p.nq, p.nv, p.nu, p.nbody, p.nsensordata, p.nrand = W, m.nq, m.nv, m.nu, m.nbody, m.nsensordata, self.nrand
    p.nb, p.anchor_k, p.ntime = self.nb, self.anchor_k, self.T
    self.soft_lo, self.soft_hi = self.soft_limits[:, 0].contiguous(), self.soft_limits[:, 1].contiguous()
    self.body_ids32, self.body_root32 = self.body_ids.int().contiguous(), self.body_root.int().contiguous()
    for n in ("qpos", "qvel", "sensordata", "xpos", "xquat", "cvel", "subtree_com"):
      setattr(p, n, getattr(D, n).data_ptr())
    for n in ("mo_jp", "mo_jv", "mo_bpos", "mo_bquat", "mo_blin", "mo_bang", "env_origins", "time_steps", "ee_mask", "default_joint_pos",
              "soft_lo", "soft_hi", "action", "prev_action", "body_pos_relative_w", "body_quat_relative_w", "episode_length_buf",
              "dones", "time_outs", "terminated", "rew", "obs", "critic_obs", "U"):
      t = getattr(self, n)
      assert t.is_contiguous(), n
      setattr(p, n, t.data_ptr())
    p.body_ids, p.body_root = self.body_ids32.data_ptr(), self.body_root32.data_ptr()
    p.term_sums, p.metrics = self.term_sums_buf.data_ptr(), self.metrics_buf.data_ptr()
    p.s_gyro, p.s_linvel, p.s_self, p.max_episode_length = self.s_gyro, self.s_linvel, self.s_self, self.max_episode_length
    for i, k in enumerate(REWARD_WEIGHTS):
      p.weight[i] = REWARD_WEIGHTS[k] * self.step_dt
    for i, k in enumerate(REWARD_STD):
      p.inv_std2[i] = 1.0 / REWARD_STD[k] ** 2
    return p

  def _push(self):
    """interval event push_by_setting_velocity (tracking_env_cfg.py:216-221, envs/mdp/events.py:173-187)"""
    D, W = self.D, self.num_envs
    self.push_time_left -= self.step_dt
    push = (self.push_time_left <= 0.0).unsqueeze(1)
    quat = D.qpos[:, 3:7]
    vel_w = torch.cat([D.qvel[:, 0:3], mdp.quat_apply(quat, D.qvel[:, 3:6])], dim=1) + self._u(self.vel_lo, self.vel_hi, W, 6)
    new_qvel = torch.cat([vel_w[:, :3], mdp.quat_apply_inverse(quat, vel_w[:, 3:])], dim=1)
    D.qvel[:, 0:6].copy_(torch.where(push, new_qvel, D.qvel[:, 0:6]))
    self.push_time_left.copy_(torch.where(push.squeeze(1), self._u(1.0, 3.0, W), self.push_time_left))

  def _step_fused(self):
    """Same step with the terms / re-anchoring / observations in csrc/k_mdp_tracking.cu; sampling, resets and the push event
    stay masked tensor ops."""
    D, M, p, dev = self.D, self.M, self._params, self.device
    self.prev_action.copy_(self.action)
    self.action.copy_(self.actions_in)
    D.ctrl.copy_(self.default_joint_pos + self.action * self.action_scale)
    engine.step(M, D, self.decimation)
    fused.trk_terms(p, dev)
    self._reset_masked(self.dones)
    engine.forward(M, D)
    self._advance_command()
    self._adaptive_ema()
    self._push()
    if not self.external_U:
      self.U.uniform_()
    fused.trk_observe(p, dev)

  def _step_impl(self):
    if self.use_fused:
      return self._step_fused()
    D, M, W = self.D, self.M, self.num_envs
    self.prev_action.copy_(self.action)
    self.action.copy_(self.actions_in)
    D.ctrl.copy_(self.default_joint_pos + self.action * self.action_scale)
    engine.step(M, D, self.decimation)
    self.episode_length_buf += 1
    # ---- terminations (mdp/terminations.py)
    jp_m, jv_m, bpos, bquat, blin, bang = self._motion()
    rpos, rquat, rlin, rang = self._robot_bodies()
    a = self.anchor_k
    self.time_outs.copy_(self.episode_length_buf >= self.max_episode_length)
    bad_anchor_z = (bpos[:, a, 2] - rpos[:, a, 2]).abs() > 0.25
    g = self.gravity_dir.expand(W, 3)
    bad_anchor_ori = (mdp.quat_apply_inverse(bquat[:, a], g)[:, 2] - mdp.quat_apply_inverse(rquat[:, a], g)[:, 2]).abs() > 0.8
    bad_ee = ((self.body_pos_relative_w[:, self.ee_k, 2] - rpos[:, self.ee_k, 2]).abs() > 0.25).any(dim=-1)
    self.terminated.copy_(bad_anchor_z | bad_anchor_ori | bad_ee)
    self.dones.copy_(self.terminated | self.time_outs)
    # ---- rewards (mdp/rewards.py)
    e = lambda err, k: torch.exp(-err / REWARD_STD[k] ** 2)      # noqa: E731
    terms = dict(
      motion_global_root_pos=e(torch.sum(torch.square(bpos[:, a] - rpos[:, a]), dim=-1), "motion_global_root_pos"),
      motion_global_root_ori=e(quat_error_magnitude(bquat[:, a], rquat[:, a]) ** 2, "motion_global_root_ori"),
      motion_body_pos=e(torch.sum(torch.square(self.body_pos_relative_w - rpos), dim=-1).mean(-1), "motion_body_pos"),
      motion_body_ori=e((quat_error_magnitude(self.body_quat_relative_w, rquat) ** 2).mean(-1), "motion_body_ori"),
      motion_body_lin_vel=e(torch.sum(torch.square(blin - rlin), dim=-1).mean(-1), "motion_body_lin_vel"),
      motion_body_ang_vel=e(torch.sum(torch.square(bang - rang), dim=-1).mean(-1), "motion_body_ang_vel"),
      action_rate_l2=mdp.action_rate_l2(self.action, self.prev_action),
      joint_limit=mdp.joint_pos_limits(D.qpos[:, 7:], self.soft_limits),
      self_collisions=D.sensordata[:, self.s_self])
    rew = torch.zeros_like(self.rew)
    for k, v in terms.items():
      val = v * (REWARD_WEIGHTS[k] * self.step_dt)
      rew += val
      self.term_sums[k] += val
    self.rew.copy_(rew)
    # ---- metrics (commands.py:219-259, the ones a training log reads)
    self.metrics["error_anchor_pos"].copy_(torch.norm(bpos[:, a] - rpos[:, a], dim=-1))
    self.metrics["error_anchor_rot"].copy_(quat_error_magnitude(bquat[:, a], rquat[:, a]))
    self.metrics["error_body_pos"].copy_(torch.norm(self.body_pos_relative_w - rpos, dim=-1).mean(-1))
    self.metrics["error_body_rot"].copy_(quat_error_magnitude(self.body_quat_relative_w, rquat).mean(-1))
    self.metrics["error_joint_pos"].copy_(torch.norm(jp_m - D.qpos[:, 7:], dim=-1))
    self.metrics["error_joint_vel"].copy_(torch.norm(jv_m - D.qvel[:, 6:], dim=-1))
    # ---- reset finished envs, then forward() (manager_based_rl_env.py:132-141)
    self._reset_masked(self.dones)
    engine.forward(M, D)
    # ---- command manager, interval push event (tracking_env_cfg.py:216-221), observations
    self._update_command()
    self._push()
    self._compute_obs()

  def _capture(self):
    side = torch.cuda.Stream(device=self.device)
    side.wait_stream(torch.cuda.current_stream(self.device))
    with torch.cuda.stream(side):
      for _ in range(3):
        self._step_impl()
    torch.cuda.current_stream(self.device).wait_stream(side)
    torch.cuda.synchronize(self.device)
    self.graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(self.graph):
      self._step_impl()
    self.bin_failed_count.zero_(); self.current_bin_failed.zero_()
    self._reset_all()
    for v in self.term_sums.values():
      v.zero_()

  # ------------------------------------------------------------------------------------------------ VecEnv surface
  def get_observations(self):
    return self.obs, {"observations": {"critic": self.critic_obs}}

  def step(self, actions: torch.Tensor):
    self.actions_in.copy_(actions)
    if self.graph is not None:
      self.graph.replay()
    else:
      self._step_impl()
    return self.obs, self.rew, self.dones, {"observations": {"critic": self.critic_obs}, "time_outs": self.time_outs}
