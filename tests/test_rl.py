"""N1/N2: velocity-task MDP terms (pure tensor functions), PPO pieces, and the multi-rank reductions (gloo, world 2)."""
import math
import os
import sys

import numpy as np
import pytest
import torch

from asimov_mjlab_b200.rl import velocity_mdp as mdp


def test_quaternion_helpers_and_heading():
  yaw = torch.tensor([0.3, -2.0, 3.0])
  q = mdp.quat_from_yaw(yaw)
  assert torch.allclose(mdp.heading_w(q), yaw, atol=1e-6)
  v = torch.randn(3, 3)
  assert torch.allclose(mdp.quat_apply_inverse(q, mdp.quat_apply(q, v)), v, atol=1e-6)
  # projected gravity of an upright base is (0, 0, -1); tilted by 80 degrees about x -> bad_orientation at 70 degrees
  g = torch.tensor([[0.0, 0.0, -1.0]])
  assert torch.allclose(mdp.quat_apply_inverse(mdp.quat_from_yaw(torch.tensor([1.0])), g), g, atol=1e-6)
  a = math.radians(80.0)
  qx = torch.tensor([[math.cos(a / 2), math.sin(a / 2), 0.0, 0.0]])
  pg = mdp.quat_apply_inverse(qx, g)
  assert bool(mdp.bad_orientation(pg, math.radians(70.0))[0]) and not bool(mdp.bad_orientation(g, math.radians(70.0))[0])
  assert torch.allclose(mdp.wrap_to_pi(torch.tensor([3.5, -3.5, 0.5])), torch.tensor([3.5 - 2 * math.pi, -3.5 + 2 * math.pi, 0.5]), atol=1e-6)


def test_air_time_state_machine_matches_reference_semantics():
  """contact_sensor.py:327-367: air time accumulates while detached, is latched into last_air_time on landing."""
  z = lambda: torch.zeros(1, 2)       # noqa: E731
  st = dict(current_air_time=z(), last_air_time=z(), current_contact_time=z(), last_contact_time=z(), last_time=torch.zeros(1))
  seq = [(1, 1), (0, 1), (0, 1), (0, 0), (1, 0), (1, 1)]
  t = 0.0
  for f in seq:
    t += 0.005
    mdp.update_air_time(st, torch.tensor([f], dtype=torch.float32), torch.tensor([t]))
  # foot 0: contact, 3 steps in the air, landed 2 steps ago; foot 1: 3 steps of contact, 2 in the air, just landed.
  # The reference latches `current + elapsed` on the transition step, so a phase of n steps reads (n + 1) * dt.
  assert st["last_air_time"][0, 0] == pytest.approx(0.020) and st["current_air_time"][0, 0] == 0
  assert st["current_contact_time"][0, 0] == pytest.approx(0.010)
  assert st["last_air_time"][0, 1] == pytest.approx(0.015) and st["last_contact_time"][0, 1] == pytest.approx(0.020)
  assert mdp.first_contact(st, 0.02).tolist() == [[True, True]] and mdp.first_contact(st, 0.004).tolist() == [[False, False]]


def test_reward_terms_closed_forms():
  cmd = torch.tensor([[1.0, 0.0, 0.5], [0.0, 0.0, 0.0]])
  lin = torch.tensor([[1.0, 0.0, 0.0], [0.5, 0.0, 0.0]])
  assert torch.allclose(mdp.track_linear_velocity(cmd, lin, 0.5), torch.tensor([1.0, math.exp(-1.0)]))
  ang = torch.tensor([[0.0, 0.0, 0.5], [0.1, 0.0, 0.0]])
  assert torch.allclose(mdp.track_angular_velocity(cmd, ang, math.sqrt(0.5)), torch.tensor([1.0, math.exp(-0.02)]))
  jp = torch.tensor([[0.0, 1.2], [-1.3, 0.0]])
  lim = torch.tensor([[-1.0, 1.0], [-1.0, 1.0]])
  assert torch.allclose(mdp.joint_pos_limits(jp, lim), torch.tensor([0.2, 0.3]))
  # posture: standing command uses the standing std
  d = torch.zeros(2)
  p = mdp.variable_posture(cmd, torch.tensor([[0.1, 0.0], [0.1, 0.0]]), d, torch.tensor([0.05, 0.05]), torch.tensor([0.3, 0.3]),
                           torch.tensor([0.5, 0.5]), 0.05, 1.5)
  assert p[0] == pytest.approx(math.exp(-0.5 * (0.1 / 0.5) ** 2)) and p[1] == pytest.approx(math.exp(-0.5 * (0.1 / 0.05) ** 2))
  # swing height: peak tracked in the air, cost on landing, then cleared
  peak = torch.tensor([[0.05, 0.0]])
  cost, new = mdp.feet_swing_height(peak, torch.tensor([[0.08, 0.02]]), torch.tensor([[0.0, 1.0]]), torch.tensor([[False, False]]), cmd[:1], 0.1, 0.05)
  assert cost.item() == 0 and new.tolist() == [[pytest.approx(0.08), 0.0]]
  cost, new = mdp.feet_swing_height(new, torch.tensor([[0.0, 0.02]]), torch.tensor([[1.0, 1.0]]), torch.tensor([[True, False]]), cmd[:1], 0.1, 0.05)
  assert cost.item() == pytest.approx((0.08 / 0.1 - 1) ** 2) and new.tolist() == [[0.0, 0.0]]
  f = torch.tensor([[[3.0, 0.0, -4.0], [0.0, 0.0, 0.0]]])
  assert mdp.soft_landing(f, torch.tensor([[True, True]]), cmd[:1], 0.05).item() == pytest.approx(5.0)
  assert torch.allclose(mdp.contact_force_obs(f)[0, :3], torch.tensor([math.log(4.0), 0.0, -math.log(5.0)]))


def test_empirical_normalization_matches_numpy():
  from asimov_mjlab_b200.rl.ppo import EmpiricalNormalization
  torch.manual_seed(0)
  n = EmpiricalNormalization(4)
  xs = [torch.randn(50, 4) * 3 + 1, torch.randn(20, 4) - 2, torch.randn(70, 4)]
  for x in xs[:2]:
    n.update(x)
  n.sync()
  n.update(xs[2]); n.sync()
  allx = torch.cat(xs).numpy()
  assert np.allclose(n.mean.numpy(), allx.mean(0), atol=1e-5) and np.allclose((n.m2 / n.count).numpy(), allx.var(0), atol=1e-4)
  assert torch.allclose(n(xs[0]), (xs[0] - n.mean) / (torch.sqrt(n.m2 / n.count) + 1e-2))


def _rank_main(rank, world, port, out):
  os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
  import torch.distributed as dist
  from asimov_mjlab_b200.rl.ppo import ActorCritic, EmpiricalNormalization, all_reduce_mean_
  dist.init_process_group("gloo", rank=rank, world_size=world)
  torch.manual_seed(100 + rank)
  x = torch.randn(40 + 10 * rank, 3) * (1 + rank) + rank
  n = EmpiricalNormalization(3)
  n.update(x); n.sync()
  # gradient all-reduce == gradient of the mean loss over the union of the shards (equal shard sizes)
  torch.manual_seed(7)
  ac = ActorCritic(5, 6, 2, hidden=(8,))
  torch.manual_seed(200 + rank)
  obs = torch.randn(16, 5)
  loss = ac.actor(obs).pow(2).mean()
  loss.backward()
  g = torch.cat([p.grad.reshape(-1) for p in ac.actor.parameters()])
  all_reduce_mean_(g)
  torch.save(dict(mean=n.mean, var=n.m2 / n.count, count=n.count, x=x, g=g, obs=obs), os.path.join(out, f"r{rank}.pt"))
  dist.destroy_process_group()


def test_two_rank_normaliser_merge_and_gradient_average(tmp_path):
  import torch.multiprocessing as mp
  port = 29600 + os.getpid() % 300
  mp.spawn(_rank_main, args=(2, port, str(tmp_path)), nprocs=2, join=True)
  r = [torch.load(os.path.join(tmp_path, f"r{k}.pt")) for k in range(2)]
  allx = torch.cat([r[0]["x"], r[1]["x"]]).numpy()
  for k in range(2):
    assert np.allclose(r[k]["mean"].numpy(), allx.mean(0), atol=1e-5) and np.allclose(r[k]["var"].numpy(), allx.var(0), atol=1e-4)
    assert float(r[k]["count"]) == len(allx)
  assert torch.allclose(r[0]["g"], r[1]["g"])
  from asimov_mjlab_b200.rl.ppo import ActorCritic
  torch.manual_seed(7)
  ac = ActorCritic(5, 6, 2, hidden=(8,))
  ac.actor(torch.cat([r[0]["obs"], r[1]["obs"]])).pow(2).mean().backward()
  g = torch.cat([p.grad.reshape(-1) for p in ac.actor.parameters()])
  assert torch.allclose(g, r[0]["g"], atol=1e-6)


@pytest.mark.gpu
def test_velocity_env_graph_step_and_ppo_iteration():
  """The whole env step replays as one CUDA graph: standing robots stay up, rewards are finite, resets restore the
  keyframe; one PPO iteration runs and changes the policy."""
  from asimov_mjlab_b200.rl.ppo import PPO
  from asimov_mjlab_b200.rl.velocity_env import G1VelocityFlatEnv
  env = G1VelocityFlatEnv(256, "cuda:0", seed=1)
  assert env.num_obs == 99 and env.num_privileged_obs == 111 and env.max_episode_length == 1000
  obs, ex = env.get_observations()
  assert obs.shape == (256, 99) and ex["observations"]["critic"].shape == (256, 111)
  tot = torch.zeros(256, device="cuda:0")
  for _ in range(50):
    obs, rew, dones, ex = env.step(torch.zeros(256, env.num_actions, device="cuda:0"))
    tot += rew
  assert torch.isfinite(obs).all() and torch.isfinite(tot).all()
  assert env.D.qpos[:, 2].mean() > 0.6 and dones.float().mean() < 0.1          # zero action = hold the keyframe: robots stand
  assert tot.mean() > 0.5 * 50 * 0.02 * 4.0                                    # tracking / posture rewards dominate while standing
  assert (env.episode_length_buf > 0).all()
  # random actions make robots fall: terminated envs are reset to the keyframe height
  for _ in range(60):
    env.step(torch.randn(256, env.num_actions, device="cuda:0") * 3.0)
  assert env.episode_length_buf.min() < 60
  fresh = env.episode_length_buf == 0
  if fresh.any():
    assert (env.D.qpos[fresh, 2] - 0.76).abs().max() < 1e-3
  algo = PPO(env, torch.device("cuda:0"), steps_per_env=8)
  w0 = algo.ac.actor[0].weight.detach().clone()      # a grad-tracking clone would pin an AccumulateGrad node made on the default stream and break the update graph capture
  algo.rollout(); info = algo.update()
  assert math.isfinite(info["value_loss"]) and not torch.equal(w0, algo.ac.actor[0].weight)


@pytest.mark.gpu
def test_fused_manager_kernels_match_torch_terms():
  """csrc/k_mdp.cu against the torch restatement of the same manager terms on identical uniform draws: physics inputs
  (ctrl, reset qpos / qvel, pushes) bit-identical, so both envs stay on one trajectory; rewards / observations /
  commands within fp32 rounding; terminations, resets and episode counters equal."""
  from asimov_mjlab_b200.rl.velocity_env import G1VelocityFlatEnv
  W = 256
  a = G1VelocityFlatEnv(W, "cuda:0", seed=3, use_graph=False, use_fused=False)
  b = G1VelocityFlatEnv(W, "cuda:0", seed=3, use_graph=False, use_fused=True)
  for e in (a, b):
    e.external_U = True
    e.max_episode_length = 37           # time-outs as well as falls
  b._params.max_episode_length = 37
  g = torch.Generator(device="cuda:0").manual_seed(11)
  # same start
  U = torch.rand(W, a.nrand, device="cuda:0", generator=g)
  for e in (a, b):
    e.U.copy_(U); e.external_U = True
    e._reset_masked(torch.ones(W, dtype=torch.bool, device="cuda:0"))
    from asimov_mjlab_b200 import engine
    engine.forward(e.M, e.D)
    e.air["last_time"].copy_(e.D.time)
    e._update_commands(); e._compute_obs()
  assert torch.equal(a.D.qpos, b.D.qpos)
  ndone = npush = 0
  for it in range(90):
    U = torch.rand(W, a.nrand, device="cuda:0", generator=g)
    act = torch.randn(W, a.num_actions, device="cuda:0", generator=g) * (3.0 if it % 3 == 0 else 0.3)
    pa, pb = a.push_time_left.clone(), b.push_time_left.clone()
    for e in (a, b):
      e.U.copy_(U)
      e.step(act)
    assert torch.equal(a.D.ctrl, b.D.ctrl) and torch.equal(a.D.qpos, b.D.qpos) and torch.equal(a.D.qvel, b.D.qvel), it
    assert torch.equal(a.dones, b.dones) and torch.equal(a.time_outs, b.time_outs)
    assert torch.equal(a.episode_length_buf, b.episode_length_buf)
    assert torch.equal(a.is_heading, b.is_heading) and torch.equal(a.is_standing, b.is_standing)
    for k in a.air:
      assert torch.allclose(a.air[k], b.air[k], atol=1e-6), k
    assert torch.allclose(a.rew, b.rew, rtol=2e-5, atol=2e-6)
    assert torch.allclose(a.obs, b.obs, rtol=1e-5, atol=2e-6) and torch.allclose(a.critic_obs, b.critic_obs, rtol=1e-5, atol=2e-6)
    assert torch.allclose(a.command, b.command, atol=2e-6) and torch.allclose(a.heading_target, b.heading_target, atol=1e-6)
    assert torch.allclose(a.cmd_time_left, b.cmd_time_left, atol=1e-6) and torch.allclose(a.push_time_left, b.push_time_left, atol=1e-6)
    assert torch.allclose(a.peak_heights, b.peak_heights, atol=1e-6)
    assert torch.equal(a.action, b.action) and torch.equal(a.prev_action, b.prev_action)
    ndone += int(a.dones.sum()); npush += int((a.push_time_left > pa).sum())
  assert ndone > W and npush > W // 2          # falls, time-outs and pushes were all exercised
  for k in a.term_sums:
    assert torch.allclose(a.term_sums[k], b.term_sums[k], rtol=1e-4, atol=1e-5), k


@pytest.mark.gpu
def test_tracking_env_follows_its_reference_motion():
  """Mjlab-Tracking-Flat-Unitree-G1 on the synthetic motion: observation sizes, a robot written onto the motion frame has
  tracking rewards near their maximum, joint targets taken from the motion keep it there better than zero actions,
  failed episodes are resampled onto the motion and feed the adaptive-sampling histogram; all inside one CUDA graph."""
  from asimov_mjlab_b200.rl import tracking_env as te
  mo = te.synthetic_motion(device="cuda:0")
  assert mo["joint_pos"].shape == (500, 29) and mo["body_pos_w"].shape == (500, 30, 3) and mo["body_quat_w"].shape == (500, 30, 4)
  assert np.allclose(np.linalg.norm(mo["body_quat_w"], axis=-1), 1.0, atol=1e-5)
  # finite-difference check of the body velocities written next to the poses (frames are 1/50 s apart)
  fd = (mo["body_pos_w"][2:] - mo["body_pos_w"][:-2]) * (50.0 / 2.0)
  assert np.abs(fd - mo["body_lin_vel_w"][1:-1]).max() < 2e-2
  env = te.G1TrackingFlatEnv(256, "cuda:0", motion=mo, seed=2)
  assert env.num_obs == 160 and env.num_privileged_obs == 286 and env.max_episode_length == 500 and env.bin_count == 11
  obs, ex = env.get_observations()
  assert obs.shape == (256, 160) and ex["observations"]["critic"].shape == (256, 286) and torch.isfinite(obs).all()
  # helper identities
  q = torch.nn.functional.normalize(torch.randn(64, 4, device="cuda:0"), dim=-1)
  assert torch.allclose(te.quat_error_magnitude(q, q), torch.zeros(64, device="cuda:0"), atol=1e-3)
  yaw = torch.rand(64, device="cuda:0") - 0.5
  assert torch.allclose(te.quat_error_magnitude(mdp_quat_yaw(yaw), mdp_quat_yaw(torch.zeros_like(yaw))), yaw.abs(), atol=1e-5)
  scale = env.action_scale
  def run(policy, n):
    tot = torch.zeros(256, device="cuda:0"); nd = 0
    for _ in range(n):
      ts = (env.time_steps + 1).clamp(max=env.T - 1)
      act = (env.mo_jp[ts] - env.default_joint_pos) / scale if policy == "motion" else torch.zeros(256, env.num_actions, device="cuda:0")
      obs, rew, dones, ex = env.step(act)
      tot += rew; nd += int(dones.sum())
      assert torch.isfinite(obs).all() and torch.isfinite(rew).all()
    return tot.mean().item() / n, nd
  track = ("motion_global_root_pos", "motion_global_root_ori", "motion_body_pos", "motion_body_ori", "motion_body_lin_vel", "motion_body_ang_vel")
  def tracking_reward(policy, n):
    for v in env.term_sums.values():
      v.zero_()
    run(policy, n)
    return sum(float(env.term_sums[k].mean()) for k in track) / n
  r_motion = tracking_reward("motion", 40)
  assert env.metrics["error_anchor_pos"].median() < 0.2 and env.metrics["error_joint_pos"].median() < 0.8
  # the six tracking terms are bounded by (0.5 + 0.5 + 4 x 1.0) x 0.02 = 0.1 per step; the motion's own joint targets keep
  # the robot well above half of that (the synthetic motion is not dynamically consistent: its root hangs still) and
  # above what zero actions reach
  r_zero = tracking_reward("zero", 40)
  assert 0.05 < r_motion <= 0.1 and r_motion > r_zero, (r_motion, r_zero)
  # tilting every robot far over terminates them (anchor orientation), resamples their phase and fills the failure histogram
  env.D.qpos[:, 3:7] = torch.tensor([0.7071, 0.7071, 0.0, 0.0], device="cuda:0")
  before = env.bin_failed_count.clone()
  obs, rew, dones, ex = env.step(torch.zeros(256, env.num_actions, device="cuda:0"))
  assert dones.float().mean() > 0.9 and (env.episode_length_buf[dones] == 0).all()
  assert env.bin_failed_count.sum() > before.sum()
  rp, rq, _, _ = env._robot_bodies()
  assert (rp[:, 0, 2] - (env.mo_bpos[env.time_steps, 0, 2])).abs().max() < 0.05      # pelvis back at the motion's height


def mdp_quat_yaw(yaw):
  from asimov_mjlab_b200.rl import velocity_mdp as mdp
  return mdp.quat_from_yaw(yaw)


@pytest.mark.gpu
def test_fused_tracking_kernels_match_torch_terms():
  """csrc/k_mdp_tracking.cu (terminations, rewards, metrics, re-anchoring, observations) against the torch restatement:
  two envs with identically seeded private generators and the same observation-noise buffer stay on ONE trajectory
  (resets, phase sampling and pushes are the same tensor ops in both), so every step compares like with like."""
  from asimov_mjlab_b200.rl import tracking_env as te
  W = 192
  mo = te.synthetic_motion(device="cuda:0")
  envs = [te.G1TrackingFlatEnv(W, "cuda:0", motion=mo, seed=4, use_graph=False, use_fused=f) for f in (False, True)]
  for e in envs:
    e.gen = torch.Generator(device="cuda:0").manual_seed(123)
    e.external_U = True
    e.bin_failed_count.zero_(); e.current_bin_failed.zero_()
    e.U.fill_(0.5)
    e._reset_all()
  a, b = envs
  assert torch.equal(a.D.qpos, b.D.qpos) and torch.equal(a.time_steps, b.time_steps)
  g = torch.Generator(device="cuda:0").manual_seed(9)
  ndone = 0
  for it in range(70):
    U = torch.rand(W, a.nrand, device="cuda:0", generator=g)
    act = torch.randn(W, a.num_actions, device="cuda:0", generator=g) * (2.0 if it % 4 == 0 else 0.2)
    for e in envs:
      e.U.copy_(U)
      e.step(act)
    assert torch.equal(a.D.qpos, b.D.qpos) and torch.equal(a.D.qvel, b.D.qvel), it
    assert torch.equal(a.time_steps, b.time_steps) and torch.equal(a.episode_length_buf, b.episode_length_buf)
    assert torch.equal(a.dones, b.dones) and torch.equal(a.terminated, b.terminated) and torch.equal(a.time_outs, b.time_outs)
    assert torch.allclose(a.rew, b.rew, rtol=1e-4, atol=2e-6)
    assert torch.allclose(a.obs, b.obs, rtol=1e-4, atol=5e-6) and torch.allclose(a.critic_obs, b.critic_obs, rtol=1e-4, atol=5e-6)
    assert torch.allclose(a.body_pos_relative_w, b.body_pos_relative_w, atol=5e-6)
    assert torch.allclose(a.body_quat_relative_w, b.body_quat_relative_w, atol=5e-6)
    assert torch.allclose(a.metrics_buf, b.metrics_buf, rtol=1e-4, atol=1e-5)
    assert torch.allclose(a.bin_failed_count, b.bin_failed_count)
    ndone += int(a.dones.sum())
  assert ndone > W // 2
  assert torch.allclose(a.term_sums_buf, b.term_sums_buf, rtol=1e-3, atol=1e-5)
