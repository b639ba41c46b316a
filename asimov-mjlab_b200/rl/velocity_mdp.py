"""Velocity-task MDP terms as pure tensor functions.

Restated from the reference's manager terms so that the multi-GPU PPO harness can run the G1 velocity-flat task directly on
the engine's tensors, without mjlab's managers (which stay unchanged and out of scope for users who keep them):
  rewards        src/mjlab/tasks/velocity/mdp/rewards.py:23-359, src/mjlab/envs/mdp/rewards.py:48-70
  observations   src/mjlab/tasks/velocity/mdp/observations.py:14-44, velocity_env_cfg.py:138-193
  terminations   src/mjlab/envs/mdp/terminations.py:24-32
  command        src/mjlab/tasks/velocity/mdp/velocity_command.py:70-125
  air time       src/mjlab/sensor/contact_sensor.py:327-367
Every function takes and returns plain tensors with a leading world dimension and does no host synchronisation, so the
whole env step can sit in CUDA graphs next to the physics decimation.
"""
from __future__ import annotations

import math

import torch


# ------------------------------------------------------------------------------------------------ quaternions (w, x, y, z)
def quat_apply(q: torch.Tensor, v: torch.Tensor) -> torch.Tensor:
  w, u = q[..., :1], q[..., 1:]
  t = 2.0 * torch.cross(u, v, dim=-1)
  return v + w * t + torch.cross(u, t, dim=-1)


def quat_apply_inverse(q: torch.Tensor, v: torch.Tensor) -> torch.Tensor:
  w, u = q[..., :1], q[..., 1:]
  t = 2.0 * torch.cross(u, v, dim=-1)
  return v - w * t + torch.cross(u, t, dim=-1)


def quat_mul(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
  aw, ax, ay, az = a.unbind(-1)
  bw, bx, by, bz = b.unbind(-1)
  return torch.stack([aw * bw - ax * bx - ay * by - az * bz, aw * bx + ax * bw + ay * bz - az * by,
                      aw * by - ax * bz + ay * bw + az * bx, aw * bz + ax * by - ay * bx + az * bw], dim=-1)


def quat_from_yaw(yaw: torch.Tensor) -> torch.Tensor:
  z = torch.zeros_like(yaw)
  return torch.stack([torch.cos(0.5 * yaw), z, z, torch.sin(0.5 * yaw)], dim=-1)


def wrap_to_pi(a: torch.Tensor) -> torch.Tensor:
  return (a + math.pi) % (2.0 * math.pi) - math.pi


def heading_w(q: torch.Tensor) -> torch.Tensor:
  """Yaw of the body x axis in the world: atan2 of the first column of R(q) (no host constant: graph-capturable)."""
  w, x, y, z = q.unbind(-1)
  return torch.atan2(2.0 * (x * y + w * z), 1.0 - 2.0 * (y * y + z * z))


# ------------------------------------------------------------------------------------------------ contact air time
def update_air_time(state: dict, found: torch.Tensor, time: torch.Tensor) -> None:
  """ContactSensor._update_air_time_tracking (contact_sensor.py:327-367), in place.  state: current_air_time,
  last_air_time, current_contact_time, last_contact_time [W, N] and last_time [W]; found [W, N]; time [W]."""
  elapsed = (time - state["last_time"]).unsqueeze(-1)
  is_contact = found > 0
  first_contact = (state["current_air_time"] > 0) & is_contact
  first_detached = (state["current_contact_time"] > 0) & ~is_contact
  state["last_air_time"].copy_(torch.where(first_contact, state["current_air_time"] + elapsed, state["last_air_time"]))
  state["current_air_time"].copy_(torch.where(~is_contact, state["current_air_time"] + elapsed, torch.zeros_like(elapsed)))
  state["last_contact_time"].copy_(torch.where(first_detached, state["current_contact_time"] + elapsed, state["last_contact_time"]))
  state["current_contact_time"].copy_(torch.where(is_contact, state["current_contact_time"] + elapsed, torch.zeros_like(elapsed)))
  state["last_time"].copy_(time)


def first_contact(state: dict, dt: float) -> torch.Tensor:
  """ContactSensor.compute_first_contact: landed within the last env step."""
  return (state["current_contact_time"] > 0) & (state["current_contact_time"] < dt + 1e-8)


# ------------------------------------------------------------------------------------------------ rewards
def _cmd_speed(cmd):
  return torch.norm(cmd[:, :2], dim=1) + torch.abs(cmd[:, 2])


def track_linear_velocity(cmd, lin_vel_b, std):
  err = torch.sum(torch.square(cmd[:, :2] - lin_vel_b[:, :2]), dim=1) + torch.square(lin_vel_b[:, 2])
  return torch.exp(-err / std ** 2)


def track_angular_velocity(cmd, ang_vel_b, std):
  err = torch.square(cmd[:, 2] - ang_vel_b[:, 2]) + torch.sum(torch.square(ang_vel_b[:, :2]), dim=1)
  return torch.exp(-err / std ** 2)


def flat_orientation(projected_gravity, std):
  return torch.exp(-torch.sum(torch.square(projected_gravity[:, :2]), dim=1) / std ** 2)


def variable_posture(cmd, joint_pos, default_joint_pos, std_standing, std_walking, std_running, walking_threshold, running_threshold):
  s = _cmd_speed(cmd)
  standing = (s < walking_threshold).float().unsqueeze(1)
  walking = ((s >= walking_threshold) & (s < running_threshold)).float().unsqueeze(1)
  running = (s >= running_threshold).float().unsqueeze(1)
  std = std_standing * standing + std_walking * walking + std_running * running
  return torch.exp(-torch.mean(torch.square(joint_pos - default_joint_pos) / std ** 2, dim=1))


def body_angular_velocity_penalty(ang_vel_w):
  return torch.sum(torch.square(ang_vel_w[:, :2]), dim=1)


def angular_momentum_penalty(angmom):
  return torch.sum(torch.square(angmom), dim=-1)


def joint_pos_limits(joint_pos, soft_limits):
  out = -(joint_pos - soft_limits[..., 0]).clip(max=0.0) + (joint_pos - soft_limits[..., 1]).clip(min=0.0)
  return torch.sum(out, dim=1)


def action_rate_l2(action, prev_action):
  return torch.sum(torch.square(action - prev_action), dim=1)


def feet_air_time(current_air_time, cmd, threshold_min, threshold_max, command_threshold):
  r = torch.sum(((current_air_time > threshold_min) & (current_air_time < threshold_max)).float(), dim=1)
  return r * (_cmd_speed(cmd) > command_threshold).float()


def feet_clearance(foot_z, foot_vel_xy, cmd, target_height, command_threshold):
  cost = torch.sum(torch.abs(foot_z - target_height) * torch.norm(foot_vel_xy, dim=-1), dim=1)
  return cost * (_cmd_speed(cmd) > command_threshold).float()


def feet_swing_height(peak_heights, foot_z, found, landed, cmd, target_height, command_threshold):
  """Class term rewards.py:186-229: returns (cost, new peak_heights)."""
  peak = torch.where(found == 0, torch.maximum(peak_heights, foot_z), peak_heights)
  err = peak / target_height - 1.0
  cost = torch.sum(torch.square(err) * landed.float(), dim=1) * (_cmd_speed(cmd) > command_threshold).float()
  return cost, torch.where(landed, torch.zeros_like(peak), peak)


def feet_slip(found, foot_vel_xy, cmd, command_threshold):
  v2 = torch.square(torch.norm(foot_vel_xy, dim=-1))
  return torch.sum(v2 * (found > 0).float(), dim=1) * (_cmd_speed(cmd) > command_threshold).float()


def soft_landing(force, landed, cmd, command_threshold):
  return torch.sum(torch.norm(force, dim=-1) * landed.float(), dim=1) * (_cmd_speed(cmd) > command_threshold).float()


# ------------------------------------------------------------------------------------------------ terminations / obs helpers
def bad_orientation(projected_gravity, limit_angle):
  return torch.acos((-projected_gravity[:, 2]).clamp(-1.0, 1.0)).abs() > limit_angle


def contact_force_obs(force):
  f = force.flatten(start_dim=1)
  return torch.sign(f) * torch.log1p(torch.abs(f))
