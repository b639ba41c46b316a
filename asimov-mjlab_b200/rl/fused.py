"""ctypes mirror of `b2s_velocity_mdp` (include/b200sim.h) and the four fused manager kernels of the velocity task
(csrc/k_mdp.cu).  The struct holds raw device pointers of torch tensors owned by the env; the env
keeps those tensors alive and never rebinds them."""
from __future__ import annotations

import ctypes as C

from .. import engine

NTERM = 14
TERMS = ("track_linear_velocity", "track_angular_velocity", "upright", "pose", "body_ang_vel", "angular_momentum",
         "dof_pos_limits", "action_rate_l2", "self_collisions", "air_time", "foot_clearance", "foot_swing_height",
         "foot_slip", "soft_landing")           # B2S_VEL_TERM_* order
_P = C.c_void_p


class VelocityMdp(C.Structure):
  _fields_ = (
    [(n, C.c_int) for n in ("nworld", "nq", "nv", "nu", "nbody", "nsite", "nsensordata", "nrand")] +
    [(n, _P) for n in ("qpos", "qvel", "qacc_warmstart", "ctrl", "time", "sensordata", "xquat", "cvel", "site_xpos", "subtree_com")] +
    [(n, _P) for n in ("default_qpos", "action_scale", "soft_lo", "soft_hi", "std_standing", "std_walking", "std_running", "env_origins")] +
    [("torso", C.c_int), ("feet_site", C.c_int * 2), ("feet_body", C.c_int * 2), ("feet_root", C.c_int * 2),
     ("s_gyro", C.c_int), ("s_linvel", C.c_int), ("s_angmom", C.c_int), ("s_found", C.c_int * 2), ("s_force", C.c_int * 2), ("s_self", C.c_int),
     ("max_episode_length", C.c_int),
     ("step_dt", C.c_float), ("landed_dt", C.c_float), ("fall_angle", C.c_float),
     ("weight", C.c_float * NTERM), ("cmd_lo", C.c_float * 4), ("cmd_hi", C.c_float * 4), ("cmd_span", C.c_float * 4)] +
    [(n, _P) for n in ("actions_in", "action", "prev_action", "command", "heading_target", "cmd_time_left", "push_time_left",
                       "is_heading", "is_standing", "dones", "time_outs", "episode_length_buf",
                       "air_cur", "air_last", "con_cur", "con_last", "air_last_time", "peak_heights",
                       "obs", "critic_obs", "rew", "term_sums", "U")])


TRK_NTERM, TRK_NMET = 9, 6


class TrackingMdp(C.Structure):
  """Mirror of b2s_tracking_mdp (include/b200sim.h)."""
  _fields_ = (
    [(n, C.c_int) for n in ("nworld", "nq", "nv", "nu", "nbody", "nsensordata", "nrand", "nb", "anchor_k", "ntime")] +
    [(n, _P) for n in ("qpos", "qvel", "sensordata", "xpos", "xquat", "cvel", "subtree_com")] +
    [(n, _P) for n in ("mo_jp", "mo_jv", "mo_bpos", "mo_bquat", "mo_blin", "mo_bang", "env_origins", "time_steps", "body_ids", "body_root",
                       "ee_mask", "default_joint_pos", "soft_lo", "soft_hi", "action", "prev_action")] +
    [("s_gyro", C.c_int), ("s_linvel", C.c_int), ("s_self", C.c_int), ("max_episode_length", C.c_int),
     ("weight", C.c_float * TRK_NTERM), ("inv_std2", C.c_float * 6)] +
    [(n, _P) for n in ("body_pos_relative_w", "body_quat_relative_w", "episode_length_buf", "dones", "time_outs", "terminated",
                       "rew", "term_sums", "metrics", "obs", "critic_obs", "U")])


_bound = False


def _lib():
  global _bound
  L = engine.lib()
  if not _bound:
    for f in ("b2s_vel_apply_action", "b2s_vel_air_time", "b2s_vel_post", "b2s_vel_observe"):
      getattr(L, f).argtypes = [C.POINTER(VelocityMdp), C.c_void_p]
      getattr(L, f).restype = C.c_int
    for f in ("b2s_trk_terms", "b2s_trk_observe"):
      getattr(L, f).argtypes = [C.POINTER(TrackingMdp), C.c_void_p]
      getattr(L, f).restype = C.c_int
    _bound = True
  return L


def _stream(device):
  import torch
  return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def _call(name, p, device):
  if getattr(_lib(), name)(C.byref(p), _stream(device)) != 0:
    raise engine.BackendError(f"{name}: invalid parameter struct (nworld, nu, nb or nrand)")


def apply_action(p, device): _call("b2s_vel_apply_action", p, device)
def air_time(p, device): _call("b2s_vel_air_time", p, device)
def post(p, device): _call("b2s_vel_post", p, device)
def observe(p, device): _call("b2s_vel_observe", p, device)
def trk_terms(p, device): _call("b2s_trk_terms", p, device)
def trk_observe(p, device): _call("b2s_trk_observe", p, device)
