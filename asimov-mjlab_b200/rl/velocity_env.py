"""Mjlab-Velocity-Flat-Unitree-G1 as an rsl_rl-style vectorised env running directly on the engine.

Restates `create_velocity_env_cfg` (src/mjlab/tasks/velocity/velocity_env_cfg.py:64-384) with the G1 parameters of
tasks/velocity/config/g1/env_cfgs.py:16-113 and the step order of ManagerBasedRlEnv.step
(envs/manager_based_rl_env.py:111-152): actions -> decimation x (ctrl write, physics step, sensor update) ->
terminations -> rewards -> reset of finished envs (+ forward) -> commands -> interval events -> observations.
The WHOLE env step -- 4 physics sub-steps with the contact air-time tracker after each, the MDP terms, the masked
resets and forward() -- is captured in ONE CUDA graph (torch ops and the engine's kernels share torch's stream), so
a policy step costs one graph launch and no host synchronisation (N2).  Interface follows the reference's
RslRlVecEnvWrapper (rl/vecenv_wrapper.py:11-111): num_envs, num_actions, num_obs, num_privileged_obs,
max_episode_length, episode_length_buf, get_observations(), step(actions) -> (obs, rew, dones, extras).

Two implementations of the manager side share every buffer: `use_fused=True` (default) launches the four fused CUDA
kernels of csrc/k_mdp.cu (7 launches per env step next to the physics), `use_fused=False` evaluates the same terms as
torch ops (~490 small kernels) and is the oracle the fused kernels are tested against.  Both take their randomness from
one uniform buffer `U` (W, nrand) refilled once per env step and read at fixed slots (include/b200sim.h).
"""
from __future__ import annotations

import math

import numpy as np
import torch

from .. import engine, scenes
from ..mjcf import resolve_expr
from ..sim import MujocoCfg, Simulation, SimulationCfg
from . import fused
from . import velocity_mdp as mdp

G1_POSTURE_WALKING = {r".*hip_pitch.*": 0.3, r".*hip_roll.*": 0.15, r".*hip_yaw.*": 0.15, r".*knee.*": 0.35, r".*ankle_pitch.*": 0.25,
                      r".*ankle_roll.*": 0.1, r".*waist_yaw.*": 0.2, r".*waist_roll.*": 0.08, r".*waist_pitch.*": 0.1,
                      r".*shoulder_pitch.*": 0.15, r".*shoulder_roll.*": 0.15, r".*shoulder_yaw.*": 0.1, r".*elbow.*": 0.15, r".*wrist.*": 0.3}
G1_POSTURE_RUNNING = {r".*hip_pitch.*": 0.5, r".*hip_roll.*": 0.2, r".*hip_yaw.*": 0.2, r".*knee.*": 0.6, r".*ankle_pitch.*": 0.35,
                      r".*ankle_roll.*": 0.15, r".*waist_yaw.*": 0.3, r".*waist_roll.*": 0.08, r".*waist_pitch.*": 0.2,
                      r".*shoulder_pitch.*": 0.5, r".*shoulder_roll.*": 0.2, r".*shoulder_yaw.*": 0.15, r".*elbow.*": 0.35, r".*wrist.*": 0.3}
REWARD_WEIGHTS = dict(track_linear_velocity=2.0, track_angular_velocity=2.0, upright=1.0, pose=1.0, body_ang_vel=-0.05,
                      angular_momentum=-0.02, dof_pos_limits=-1.0, action_rate_l2=-0.01, self_collisions=-1.0, air_time=0.0,
                      foot_clearance=-0.5, foot_swing_height=-0.1, foot_slip=-0.1, soft_landing=-1e-5)


class G1VelocityFlatEnv:
  decimation = 4
  episode_length_s = 20.0

  def __init__(self, num_envs: int, device: str = "cuda:0", seed: int = 0, env_spacing: float = 2.5, use_graph: bool = True,
               use_fused: bool = True, origin_sites: bool = False):
    self.num_envs, self.device = num_envs, torch.device(device)
    self.use_fused = use_fused
    torch.cuda.set_device(self.device)
    m = scenes.load_asset("g1_flat")
    if origin_sites:      # the scene as mjlab compiles it: one world-body site per env (terrain_importer.py:98-120)
      m = scenes.with_env_origin_sites(m, num_envs, env_spacing)
    self.m = m
    cfg = SimulationCfg(nconmax=50, njmax=300, ls_parallel=True, mujoco=MujocoCfg(timestep=0.005, iterations=10, ls_iterations=20))
    self.sim = Simulation(num_envs, cfg, m, str(self.device))
    self.M, self.D = self.sim.model, self.sim.data
    self.physics_dt = 0.005
    self.step_dt = self.physics_dt * self.decimation
    self.max_episode_length = math.ceil(self.episode_length_s / self.step_dt)
    g = torch.Generator(device="cpu").manual_seed(seed)
    dev = self.device
    W = num_envs
    # startup event: foot friction domain randomisation (velocity_env_cfg.py:230-240)
    self.sim.expand_model_fields(("geom_friction",))
    foot = [i for i, n in enumerate(m.names["geom"]) if "foot" in n and n.endswith("_collision")]
    self.M.geom_friction[:, foot, 0] = (0.3 + 0.9 * torch.rand(W, len(foot), generator=g)).to(dev)
    # names / indices
    jn = [n.split("/")[-1] for n in m.names["joint"][1:]]
    self.joint_names = jn
    self.num_actions = m.nu
    bn = m.names["body"]
    self.torso = bn.index("robot/torso_link")
    self.root = int(m.jnt_bodyid[0])
    self.feet_sites = [m.names["site"].index("robot/left_foot"), m.names["site"].index("robot/right_foot")]
    self.feet_site_body = [int(m.site_bodyid[s]) for s in self.feet_sites]
    sa = {n: (int(a), int(d)) for n, a, d in zip(m.names["sensor"], m.sensor_adr, m.sensor_dim)}
    self.s_gyro, self.s_linvel, self.s_angmom = sa["robot/imu_ang_vel"][0], sa["robot/imu_lin_vel"][0], sa["robot/root_angmom"][0]
    self.s_found = [sa["feet_ground_contact_left_ankle_roll_link_found"][0], sa["feet_ground_contact_right_ankle_roll_link_found"][0]]
    self.s_force = [sa["feet_ground_contact_left_ankle_roll_link_force"][0], sa["feet_ground_contact_right_ankle_roll_link_force"][0]]
    self.s_self = sa["self_collision_pelvis_found"][0]
    key = m.keys["robot/init_state"]
    f32 = dict(dtype=torch.float32, device=dev)
    self.default_qpos = torch.tensor(key["qpos"], **f32)
    self.default_joint_pos = self.default_qpos[7:].clone()
    self.action_scale = torch.tensor(scenes.g1_action_scale(jn), **f32)
    rng = torch.tensor(np.asarray(m.jnt_range)[1:], **f32)
    mid, half = 0.5 * (rng[:, 0] + rng[:, 1]), 0.5 * (rng[:, 1] - rng[:, 0])
    self.soft_limits = torch.stack([mid - 0.9 * half, mid + 0.9 * half], dim=-1)      # soft_joint_pos_limit_factor = 0.9
    self.std_standing = torch.tensor(resolve_expr({".*": 0.05}, jn), **f32)
    self.std_walking = torch.tensor(resolve_expr(G1_POSTURE_WALKING, jn, 0.3), **f32)
    self.std_running = torch.tensor(resolve_expr(G1_POSTURE_RUNNING, jn, 0.3), **f32)
    self.env_origins = torch.tensor(scenes.env_origins_grid(W, env_spacing), **f32)
    self.gravity_dir = torch.tensor([0.0, 0.0, -1.0], **f32)
    # buffers
    z = lambda *s: torch.zeros(*s, **f32)      # noqa: E731
    self.actions_in = z(W, self.num_actions)
    self.action, self.prev_action = z(W, self.num_actions), z(W, self.num_actions)
    self.command, self.heading_target = z(W, 3), z(W)
    self.is_heading = torch.zeros(W, dtype=torch.bool, device=dev)
    self.is_standing = torch.zeros(W, dtype=torch.bool, device=dev)
    self.cmd_time_left = z(W)
    self.push_time_left = z(W)
    self.episode_length_buf = torch.zeros(W, dtype=torch.long, device=dev)
    self.air = dict(current_air_time=z(W, 2), last_air_time=z(W, 2), current_contact_time=z(W, 2), last_contact_time=z(W, 2), last_time=z(W))
    self.peak_heights = z(W, 2)
    self.num_obs, self.num_privileged_obs = 3 + 3 + 3 + 3 * self.num_actions + 3, 3 + 3 + 3 + 3 * self.num_actions + 3 + 2 + 2 + 2 + 6
    self.obs, self.critic_obs = z(W, self.num_obs), z(W, self.num_privileged_obs)
    self.rew, self.dones, self.time_outs = z(W), torch.zeros(W, dtype=torch.bool, device=dev), torch.zeros(W, dtype=torch.bool, device=dev)
    self.term_sums_buf = z(fused.NTERM, W)
    self.term_sums = {k: self.term_sums_buf[i] for i, k in enumerate(fused.TERMS)}
    assert list(REWARD_WEIGHTS) == list(fused.TERMS)
    # uniform draws of one env step: observation noise first, then the slots of include/b200sim.h
    self.u0 = 9 + 2 * self.num_actions
    self.nrand = self.u0 + 21
    self.U = z(W, self.nrand)
    self.external_U = False      # tests fill U themselves to drive both implementations with the same draws
    self.cmd_ranges = dict(lin_vel_x=(-1.0, 1.0), lin_vel_y=(-1.0, 1.0), ang_vel_z=(-0.5, 0.5), heading=(-math.pi, math.pi))
    torch.manual_seed(seed)
    self._params = self._make_params() if use_fused else None
    self._reset_all()
    self.graph = None
    if use_graph:
      self._capture()

  # ------------------------------------------------------------------------------------------------ state views
  def _root_quat(self):
    return self.D.qpos[:, 3:7]

  def _projected_gravity(self, quat):
    return mdp.quat_apply_inverse(quat, self.gravity_dir.expand(self.num_envs, 3))

  def _site_lin_vel_w(self, k):
    """Entity data site_lin_vel_w: velocity of the site point from the body's spatial velocity about the subtree COM."""
    b = self.feet_site_body[k]
    cv = self.D.cvel[:, b]
    off = self.D.site_xpos[:, self.feet_sites[k]] - self.D.subtree_com[:, int(self.m.body_rootid[b])]
    return cv[:, 3:6] + torch.cross(cv[:, 0:3], off, dim=-1)

  # ------------------------------------------------------------------------------------------------ pieces of the step
  def _u(self, slot, lo, hi):
    """lo + (hi - lo) * U[:, slot]"""
    return lo + (hi - lo) * self.U[:, slot]

  def _resample_commands(self, mask, s0):
    """s0: first of the 7 slots of U this resample consumes (reset: u0 + 4, timer: u0 + 11)."""
    r = self.cmd_ranges
    new = torch.stack([self._u(s0, *r["lin_vel_x"]), self._u(s0 + 1, *r["lin_vel_y"]), self._u(s0 + 2, *r["ang_vel_z"])], dim=1)
    self.command.copy_(torch.where(mask.unsqueeze(1), new, self.command))
    self.heading_target.copy_(torch.where(mask, self._u(s0 + 3, *r["heading"]), self.heading_target))
    self.is_heading.copy_(torch.where(mask, self.U[:, s0 + 4] <= 0.3, self.is_heading))
    self.is_standing.copy_(torch.where(mask, self.U[:, s0 + 5] <= 0.1, self.is_standing))
    self.cmd_time_left.copy_(torch.where(mask, self._u(s0 + 6, 3.0, 8.0), self.cmd_time_left))

  def _update_commands(self):
    err = mdp.wrap_to_pi(self.heading_target - mdp.heading_w(self._root_quat()))
    wz = torch.clip(0.5 * err, self.cmd_ranges["ang_vel_z"][0], self.cmd_ranges["ang_vel_z"][1])
    self.command[:, 2].copy_(torch.where(self.is_heading, wz, self.command[:, 2]))
    self.command.copy_(torch.where(self.is_standing.unsqueeze(1), torch.zeros_like(self.command), self.command))

  def _reset_masked(self, mask):
    """reset_root_state_uniform + reset_joints_by_offset (velocity_env_cfg.py:206-223), written through a mask."""
    W, D, u0 = self.num_envs, self.D, self.u0
    qpos = self.default_qpos.expand(W, -1).clone()
    qpos[:, 0] += self._u(u0, -0.5, 0.5) + self.env_origins[:, 0]
    qpos[:, 1] += self._u(u0 + 1, -0.5, 0.5) + self.env_origins[:, 1]
    qpos[:, 2] += self.env_origins[:, 2]
    qpos[:, 3:7] = mdp.quat_mul(qpos[:, 3:7], mdp.quat_from_yaw(self._u(u0 + 2, -3.14, 3.14)))
    m1 = mask.unsqueeze(1)
    D.qpos.copy_(torch.where(m1, qpos, D.qpos))
    D.qvel.copy_(torch.where(m1, torch.zeros_like(D.qvel), D.qvel))
    D.qacc_warmstart.copy_(torch.where(m1, torch.zeros_like(D.qacc_warmstart), D.qacc_warmstart))
    D.ctrl.copy_(torch.where(m1, self.default_joint_pos.expand(W, -1), D.ctrl))
    for t in (self.action, self.prev_action):
      t.copy_(torch.where(m1, torch.zeros_like(t), t))
    for k in ("current_air_time", "last_air_time", "current_contact_time", "last_contact_time"):
      self.air[k].copy_(torch.where(m1, torch.zeros_like(self.air[k]), self.air[k]))
    self.peak_heights.copy_(torch.where(m1, torch.zeros_like(self.peak_heights), self.peak_heights))
    self.episode_length_buf.copy_(torch.where(mask, torch.zeros_like(self.episode_length_buf), self.episode_length_buf))
    self.push_time_left.copy_(torch.where(mask, self._u(u0 + 3, 1.0, 3.0), self.push_time_left))
    self._resample_commands(mask, u0 + 4)

  def _reset_all(self):
    self.U.uniform_()
    self._reset_masked(torch.ones(self.num_envs, dtype=torch.bool, device=self.device))
    engine.forward(self.M, self.D)
    self.air["last_time"].copy_(self.D.time)
    self._update_commands()
    self._compute_obs()

  def _compute_obs(self):
    D, nu = self.D, self.num_actions
    un = lambda t, a, s: t + (2.0 * self.U[:, s:s + t.shape[1]] - 1.0) * a      # noqa: E731   Unoise(n_min=-a, n_max=a)
    sd = D.sensordata
    lin, gyro = sd[:, self.s_linvel:self.s_linvel + 3], sd[:, self.s_gyro:self.s_gyro + 3]
    pg = self._projected_gravity(self._root_quat())
    jp = D.qpos[:, 7:] - self.default_joint_pos
    jv = D.qvel[:, 6:]
    found = torch.stack([sd[:, a] for a in self.s_found], dim=1)
    force = torch.stack([sd[:, a:a + 3] for a in self.s_force], dim=1)
    clean = [lin, gyro * 0.25, pg, jp, jv * 0.05, self.action, self.command]
    noisy = [un(lin, 0.5, 0), un(gyro, 0.2, 3) * 0.25, un(pg, 0.05, 6), un(jp, 0.01, 9), un(jv, 1.5, 9 + nu) * 0.05, self.action, self.command]
    foot_z = torch.stack([D.site_xpos[:, s, 2] for s in self.feet_sites], dim=1)
    self.obs.copy_(torch.cat(noisy, dim=1))
    self.critic_obs.copy_(torch.cat(clean + [foot_z, self.air["current_air_time"], (found > 0).float(), mdp.contact_force_obs(force)], dim=1))

  def _step_impl(self):
    if not self.external_U:
      self.U.uniform_()
    if self.use_fused:
      return self._step_fused()
    D, M = self.D, self.M
    # action manager: process + apply (the same joint targets on every sub-step)
    self.prev_action.copy_(self.action)
    self.action.copy_(self.actions_in)
    D.ctrl.copy_(self.default_joint_pos + self.action * self.action_scale)
    for _ in range(self.decimation):
      engine.step(M, D)
      found = torch.stack([D.sensordata[:, a] for a in self.s_found], dim=1)
      mdp.update_air_time(self.air, found, D.time)                       # scene.update(dt) per physics step
    self.episode_length_buf += 1
    sd = D.sensordata
    quat = self._root_quat()
    pg = self._projected_gravity(quat)
    # terminations
    self.time_outs.copy_(self.episode_length_buf >= self.max_episode_length)
    fell = mdp.bad_orientation(pg, math.radians(70.0))
    self.dones.copy_(fell | self.time_outs)
    # rewards
    cmd = self.command
    lin_b = mdp.quat_apply_inverse(quat, D.qvel[:, 0:3])
    ang_b = D.qvel[:, 3:6]
    found = torch.stack([sd[:, a] for a in self.s_found], dim=1)
    force = torch.stack([sd[:, a:a + 3] for a in self.s_force], dim=1)
    foot_z = torch.stack([D.site_xpos[:, s, 2] for s in self.feet_sites], dim=1)
    foot_vxy = torch.stack([self._site_lin_vel_w(k)[:, :2] for k in range(2)], dim=1)
    landed = mdp.first_contact(self.air, self.step_dt)
    swing_cost, peaks = mdp.feet_swing_height(self.peak_heights, foot_z, found, landed, cmd, 0.1, 0.05)
    self.peak_heights.copy_(peaks)
    terms = dict(
      track_linear_velocity=mdp.track_linear_velocity(cmd, lin_b, math.sqrt(0.25)),
      track_angular_velocity=mdp.track_angular_velocity(cmd, ang_b, math.sqrt(0.5)),
      upright=mdp.flat_orientation(self._projected_gravity(D.xquat[:, self.torso]), math.sqrt(0.2)),
      pose=mdp.variable_posture(cmd, D.qpos[:, 7:], self.default_joint_pos, self.std_standing, self.std_walking, self.std_running, 0.05, 1.5),
      body_ang_vel=mdp.body_angular_velocity_penalty(D.cvel[:, self.torso, 0:3]),
      angular_momentum=mdp.angular_momentum_penalty(sd[:, self.s_angmom:self.s_angmom + 3]),
      dof_pos_limits=mdp.joint_pos_limits(D.qpos[:, 7:], self.soft_limits),
      action_rate_l2=mdp.action_rate_l2(self.action, self.prev_action),
      self_collisions=sd[:, self.s_self],
      air_time=mdp.feet_air_time(self.air["current_air_time"], cmd, 0.05, 0.5, 0.5),
      foot_clearance=mdp.feet_clearance(foot_z, foot_vxy, cmd, 0.1, 0.05),
      foot_swing_height=swing_cost,
      foot_slip=mdp.feet_slip(found, foot_vxy, cmd, 0.05),
      soft_landing=mdp.soft_landing(force, landed, cmd, 0.05))
    rew = torch.zeros_like(self.rew)
    for k, v in terms.items():
      if REWARD_WEIGHTS[k] != 0.0:
        val = v * (REWARD_WEIGHTS[k] * self.step_dt)
        rew += val
        self.term_sums[k] += val
    self.rew.copy_(rew)
    # reset finished envs, then forward() so that their derived quantities are fresh (manager_based_rl_env.py:132-141)
    self._reset_masked(self.dones)
    engine.forward(M, D)
    self.air["last_time"].copy_(torch.where(self.dones, D.time, self.air["last_time"]))
    # command manager: resample on timers, heading control
    self.cmd_time_left -= self.step_dt
    self._resample_commands(self.cmd_time_left <= 0.0, self.u0 + 11)
    self._update_commands()
    # interval event: push_by_setting_velocity every U(1, 3) s (velocity_env_cfg.py:224-229)
    self.push_time_left -= self.step_dt
    push = self.push_time_left <= 0.0
    kick = self.U[:, self.u0 + 18:self.u0 + 20] - 0.5
    D.qvel[:, 0:2].copy_(torch.where(push.unsqueeze(1), D.qvel[:, 0:2] + kick, D.qvel[:, 0:2]))
    self.push_time_left.copy_(torch.where(push, self._u(self.u0 + 20, 1.0, 3.0), self.push_time_left))
    self._compute_obs()

  # ------------------------------------------------------------------------------------------------ fused path
  def _make_params(self):
    m, D, W = self.m, self.D, self.num_envs
    p = fused.VelocityMdp()
    p.nworld, p.nq, p.nv, p.nu, p.nbody, p.nsite, p.nsensordata, p.nrand = W, m.nq, m.nv, m.nu, m.nbody, m.nsite, m.nsensordata, self.nrand
    self.soft_lo, self.soft_hi = self.soft_limits[:, 0].contiguous(), self.soft_limits[:, 1].contiguous()
    ptr = lambda t: t.data_ptr()      # noqa: E731
    for n in ("qpos", "qvel", "qacc_warmstart", "ctrl", "time", "sensordata", "xquat", "cvel", "subtree_com"):
      setattr(p, n, ptr(getattr(D, n)))
    # site frames: dense (nworld, nsite, 3), or -- with many static world-body sites (engine.SplitField) -- the per-world block
    # (nworld, nsite_world, 3) addressed by slot
    split = not torch.is_tensor(D.site_xpos)
    site_w = D.site_xpos.world if split else D.site_xpos
    p.site_xpos, p.nsite = ptr(site_w), int(site_w.shape[1])
    site_slot = (lambda k: int(D.site_xpos.slot[k])) if split else (lambda k: k)
    for n in ("default_qpos", "action_scale", "soft_lo", "soft_hi", "std_standing", "std_walking", "std_running", "env_origins",
              "actions_in", "action", "prev_action", "command", "heading_target", "cmd_time_left", "push_time_left", "is_heading",
              "is_standing", "dones", "time_outs", "episode_length_buf", "peak_heights", "obs", "critic_obs", "rew", "U"):
      t = getattr(self, n)
      assert t.is_contiguous()
      setattr(p, n, ptr(t))
    p.term_sums = ptr(self.term_sums_buf)
    p.air_cur, p.air_last = ptr(self.air["current_air_time"]), ptr(self.air["last_air_time"])
    p.con_cur, p.con_last = ptr(self.air["current_contact_time"]), ptr(self.air["last_contact_time"])
    p.air_last_time = ptr(self.air["last_time"])
    p.torso = self.torso
    for k in range(2):
      p.feet_site[k], p.feet_body[k] = site_slot(self.feet_sites[k]), self.feet_site_body[k]
      p.feet_root[k] = int(m.body_rootid[self.feet_site_body[k]])
      p.s_found[k], p.s_force[k] = self.s_found[k], self.s_force[k]
    p.s_gyro, p.s_linvel, p.s_angmom, p.s_self = self.s_gyro, self.s_linvel, self.s_angmom, self.s_self
    p.max_episode_length = self.max_episode_length
    p.step_dt, p.landed_dt, p.fall_angle = self.step_dt, self.step_dt + 1e-8, math.radians(70.0)
    for i, k in enumerate(fused.TERMS):
      p.weight[i] = REWARD_WEIGHTS[k] * self.step_dt
    for i, k in enumerate(("lin_vel_x", "lin_vel_y", "ang_vel_z", "heading")):
      lo, hi = self.cmd_ranges[k]
      p.cmd_lo[i], p.cmd_hi[i], p.cmd_span[i] = lo, hi, hi - lo
    return p

  def _step_fused(self):
    p, dev, M, D = self._params, self.device, self.M, self.D
    fused.apply_action(p, dev)
    for _ in range(self.decimation):
      engine.step(M, D)
      fused.air_time(p, dev)
    fused.post(p, dev)
    engine.forward(M, D)
    fused.observe(p, dev)

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
