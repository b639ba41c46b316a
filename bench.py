#!/usr/bin/env python
"""Headline benchmark: G1 velocity-flat physics step with random actions (BASELINE.json configs[1]).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--envs E] [--impl reference]

One bench "step" = one control decimation = 4 physics sub-steps (dt 5 ms) of E worlds per GPU with one
fresh batch of random joint-position targets.  `value` counts physics sub-steps
(1 env-step = 1 sim.step() per world) per second over all GPUs, inputs resident in HBM; `e2e` is the same
through the host-buffer C-ABI call (H2D ctrl, D2H qpos/qvel/sensordata inside the timed region).
Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

ALG_BYTES_PER_SUBSTEP_WORLD = 5268  # 4 B x (363 in + 747 out + 207 per-world Model reads), G1-flat
WORKLOADS = {
  "g1_flat": ("g1_flat", "Mjlab-Velocity-Flat-Unitree-G1 physics step, random actions (BASELINE.json configs[1])"),
  "go1_flat": ("go1_flat", "Mjlab-Velocity-Flat-Unitree-Go1 physics step, random actions"),
  "asimov_toe_flat": ("asimov_toe_flat", "Asimov-toe velocity task physics step on the plane (fixed ankle tendons + tendon position actuators; mesh collision geoms dropped), random actions"),
  "asimov_toe_rough": ("asimov_toe_rough", "Asimov-toe velocity task physics step on the reference's default box terrain, random actions"),
  "go1_rough_hf": ("go1_rough_hf", "Go1 rough task with the generator's height-field sub-terrains enabled (terrains/config.py:28-53, commented out in the reference), random actions (BASELINE.json configs[2], hfield variant)"),
  "asimov_toe_full_flat": ("asimov_toe_full_flat", "Asimov-toe, FULL_COLLISION (11 link meshes by convex hull + 20 capsules, self collisions) on the plane, random actions"),
  "asimov_flat": ("asimov_flat", "Mjlab-Velocity-Flat-Asimov physics step (12-dof biped, all-mesh collision geometry: the two foot meshes on the plane, plane-mesh manifolds), random actions"),
  "g1_flat_origins": ("g1_flat", "the G1-flat scene WITH the reference's per-env origin sites on the world body (terrain_importer.py:95,98-120; SURVEY.md §0 fact 10): num_envs static sites stored once, not per world (H3)"),
  "go1_rough": ("go1_rough", "Mjlab-Velocity-Rough-Unitree-Go1 physics step on the reference's default box terrain, random actions (BASELINE.json configs[2], box variant)"),
}


def alg_bytes(m, friction_geoms):
  """SURVEY.md §8(d) formula: 4 B x (state in + state/manager-visible out + per-world Model reads) per world per sub-step.
  in  = qpos + qvel + ctrl + qacc_warmstart + time + xfrc_applied + qfrc_applied
  out = qpos + qvel + qacc + qacc_warmstart + time + sensordata + xpos + xquat + subtree_com + cvel (robot bodies count
        as nbody) + site_xpos + site_xmat; Model = geom_friction of the geoms that are randomised per world."""
  nb = m.nbody
  r = m.nq + m.nv + m.nu + m.nv + 1 + 6 * nb + m.nv
  w = m.nq + 3 * m.nv + 1 + m.nsensordata + (3 + 4 + 3 + 6) * nb + 12 * m.nsite
  return 4 * (r + w + 3 * friction_geoms)
# DRAM traffic of one sub-step (the 7 stage kernels + the solver's second launch), dram__bytes_read.sum + dram__bytes_write.sum from
# the `ncu --set full` capture profiles/step_r2g_metrics.txt (G1-flat, 4096 worlds, upright regime): 144.6 MB read + 7.4 MB written.
# NOT measured live (ncu cannot run inside the timed region): the JSON names the file it comes from in `traffic_from_profile`.
NCU_TRAFFIC = {("g1_flat", 4096): (151800000, "profiles/step_r2g_metrics.txt (ncu --set full, sum over the 8 launches of one sub-step, upright regime)")}
DECIMATION = 4
RESET_PERIOD_STEPS = 250            # reset every 1000 sub-steps, forward() inside the timed region


def parse():
  ap = argparse.ArgumentParser()
  ap.add_argument("--gpus", type=int, default=1)
  ap.add_argument("--steps", type=int, default=300)
  ap.add_argument("--warmup", type=int, default=20)
  ap.add_argument("--envs", type=int, default=4096, help="worlds per GPU")
  ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
  ap.add_argument("--cpu-baseline", type=int, default=1)
  ap.add_argument("--ls-parallel", type=int, default=1)
  ap.add_argument("--workload", default="g1_flat", choices=sorted(WORKLOADS))
  ap.add_argument("--extras", type=int, default=1, help="also measure (outside the timed region) the state regimes, the other workloads and PPO")
  ap.add_argument("--overlap", type=int, default=-1, help="stage overlap on forked graph branches (-1: the library's per-model default, 0: serial stages, 1: all branches, 2: only the solver's second launch)")
  ap.add_argument("--cone", default="pyramidal", choices=["pyramidal", "elliptic"], help="friction cone (the reference's default is pyramidal)")
  return ap.parse_args()


def measured_peak():
  p = os.path.join(ROOT, "MEASURED_PEAKS.json")
  if os.path.exists(p):
    with open(p) as f:
      return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
  return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
  """nvidia-smi sampling DURING the timed region (B200_PROFILING.md clocks line)."""

  def __init__(self, index=0):
    self.rows, self.proc, self.index = [], None, index

  def start(self):
    try:
      self.proc = subprocess.Popen(
        ["nvidia-smi", f"--id={self.index}",
         "--query-gpu=clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap",
         "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
      self.thread = threading.Thread(target=self._read, daemon=True)
      self.thread.start()
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
      self.proc.wait(timeout=2)
    except Exception:
      self.proc.kill()
    sm = sorted(float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit())
    mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
    names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
    reasons = sorted({names[k] for r in self.rows if len(r) >= 7 for k in range(4) if r[3 + k].lower().startswith("active")})
    return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons, "samples": len(sm)}


def load_model(ls_parallel=1, workload="g1_flat", cone="pyramidal", origin_sites=0):
  from asimov_mjlab_b200 import scenes
  m = scenes.load_asset(WORKLOADS[workload][0])
  if origin_sites:
    m = scenes.with_env_origin_sites(m, origin_sites)
  m.opt.ls_parallel = bool(ls_parallel)
  m.opt.cone = 1 if cone == "elliptic" else 0
  return m


def action_targets(m, n_batches, envs, seed, device=None):
  """ctrl = default_joint_pos + G1_ACTION_SCALE * a, a ~ U(-1,1)."""
  import numpy as np
  from asimov_mjlab_b200 import scenes
  key = m.keys["robot/init_state"]
  jn = [n.split("/")[-1] for n in m.names["actuator"]]
  rng = np.random.default_rng(seed)
  a = rng.uniform(-1.0, 1.0, size=(n_batches, envs, m.nu)).astype(np.float32)
  if getattr(m, "ntendon", 0):
    # Asimov-toe (env_cfgs.py:136-160): joint_pos action on hips / knees / toes... and ankle_ab on the 4 tendon
    # actuators = AnklePrToTendonAction(default ankle angles + 1.0 * a) (its scale dict has no ankle entry -> 1.0)
    scale = scenes.asimov_action_scale(jn)
    ctrl = np.asarray(key["ctrl"], np.float32)[None, None, :] + scale.astype(np.float32)[None, None, :] * a
    jp = scenes.ASIMOV_KNEES_BENT["joint_pos"]
    ab = scenes.ankle_pr_to_tendon(jp["left_ankle_pitch_joint"] + a[..., 0], a[..., 1], jp["right_ankle_pitch_joint"] + a[..., 2], a[..., 3])
    for k, n in enumerate(("left_ankle_A", "left_ankle_B", "right_ankle_A", "right_ankle_B")):
      ctrl[..., jn.index(n)] = ab[..., k]
    return ctrl.astype(np.float32)
  if m.meta.get("task") == "Mjlab-Velocity-Flat-Asimov":
    return (np.asarray(key["ctrl"], np.float32)[None, None, :] + scenes.asimov12_action_scale(jn).astype(np.float32)[None, None, :] * a)
  scale = scenes.g1_action_scale(jn) if m.nu == 29 else scenes.go1_action_scale(jn)
  return (np.asarray(key["ctrl"], np.float32)[None, None, :] + scale.astype(np.float32)[None, None, :] * a)


# ------------------------------------------------------------------------------------------- reference arm (CPU)
class CpuSampler:
  """The CPU arm: the oracle (CPU restatement of the reference's step; bench.py may execute oracle/ only here) on the host
  cores, OpenMP over worlds.  One EPISODE = `episode_substeps` sub-steps from the reset state with a fresh action batch every
  DECIMATION sub-steps -- the same state distribution the GPU arm's timed region sees (its K steps follow a reset).  Only the
  rollout calls are timed: joint targets are written between them, outside the clock.  Both `cpu_baseline` (in the B200 arm's
  line) and `--impl reference` use this one procedure, so they agree."""

  def __init__(self, m, ls_parallel=1, worlds_per_core=16, seed=0):
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    from oracle import Oracle
    self.m, self.cores = m, os.cpu_count() or 1
    self.nw = worlds_per_core * self.cores
    self.orc = Oracle(m, "f32", 50, 300, ls_parallel=ls_parallel)
    self.key = m.keys["robot/init_state"]
    self.worlds = [self.orc.new_world(self.key["qpos"], None, self.key["ctrl"]) for _ in range(self.nw)]
    self.seed = seed

  def reset(self):
    import numpy as np
    rng = np.random.default_rng(self.seed)
    for w in self.worlds:
      w["qpos"][:] = self.key["qpos"]; w["qpos"][7:] += 0.02 * rng.standard_normal(self.m.nq - 7)
      w["qvel"][:] = 0; w["qacc_warmstart"][:] = 0; w["time"][:] = 0; w["ctrl"][:] = self.key["ctrl"]
      self.orc.forward(w)

  def episode(self, episode_substeps):
    """Returns the seconds spent inside the rollout calls of one episode."""
    nb = max(episode_substeps // DECIMATION, 1)
    ctrl = action_targets(self.m, nb, self.nw, self.seed)
    self.seed += 1
    self.reset()
    t = 0.0
    for b in range(nb):
      for w in range(self.nw):
        self.worlds[w]["ctrl"][:] = ctrl[b, w]
      t0 = time.perf_counter()
      self.orc.rollout(self.worlds, DECIMATION, self.cores)
      t += time.perf_counter() - t0
    return t, nb * DECIMATION * self.nw

  def run(self, episode_substeps, min_seconds, max_seconds=60.0, min_episodes=1):
    self.episode(min(episode_substeps, 8 * DECIMATION))          # warm: OpenMP team, page faults, caches
    t = n = 0.0
    ep = 0
    while (t < min_seconds or ep < min_episodes) and t < max_seconds:
      dt, dn = self.episode(episode_substeps)
      t += dt; n += dn; ep += 1
    return n / t, t, ep


def cpu_sample_desc(cs, episode_substeps, t, ep):
  return (f"{cs.nw} worlds x {ep} episode(s) of {episode_substeps} sub-steps from the reset state (the GPU arm's timed window), "
          f"oracle f32 with OpenMP on {cs.cores} threads, {t:.1f} s inside the rollout calls")


def run_reference(args):
  rank = int(os.environ.get("RANK", "0"))
  if rank != 0:
    return
  m = load_model(args.ls_parallel, args.workload, args.cone)
  cs = CpuSampler(m, args.ls_parallel)
  # a "step" of this arm = one episode of the GPU arm's timed window (K decimations after a reset) on cs.nw worlds; W warm-up
  # and K timed steps would take hours on the CPU, so the sample is bounded: at least 5 s, at most 60 s of rollout time
  window = min(args.steps, RESET_PERIOD_STEPS) * DECIMATION
  rate, t, ep = cs.run(window, min_seconds=5.0)
  out = {
    "impl": "reference", "metric": "env_steps_per_s", "value": rate, "unit": "physics sub-steps/s (1 env-step = 1 sim.step per world)",
    "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": t / max(ep, 1) / max(window // DECIMATION, 1) * 1e3,
    "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
    "config": {"workload": WORKLOADS[args.workload][1], "envs_per_step": cs.nw, "decimation": DECIMATION,
               "bounded_sample": f"{ep} episode(s) x {window} sub-steps x {cs.nw} worlds; one step = one decimation of {cs.nw} worlds"},
    "cpu_baseline": {"value": rate, "unit": "physics sub-steps/s", "cores": cs.cores, "kind": "port", "sample": cpu_sample_desc(cs, window, t, ep)},
    "e2e": {"value": rate, "unit": "physics sub-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
  }
  print(json.dumps(out), flush=True)


# ------------------------------------------------------------------------------------------------ B200 arm
class PhysBench:
  """One physics workload on one GPU: the Simulation, its reset state, a pool of action batches and the bench step."""

  def __init__(self, workload, E, dev, rank, ls_parallel=1, cone="pyramidal", overlap=-1):
    import numpy as np
    import torch
    from asimov_mjlab_b200 import engine
    from asimov_mjlab_b200.sim import MujocoCfg, Simulation, SimulationCfg
    self.torch, self.engine, self.E, self.dev = torch, engine, E, dev
    m = self.m = load_model(ls_parallel, workload, cone, origin_sites=E if workload.endswith("_origins") else 0)
    cfg = SimulationCfg(nconmax=50, njmax=300, ls_parallel=bool(ls_parallel),
                        mujoco=MujocoCfg(timestep=0.005, iterations=10, ls_iterations=20, cone=cone))
    sim = self.sim = Simulation(E, cfg, m, dev)
    M, D = self.M, self.D = sim.model, sim.data
    # foot-friction domain randomisation ON (per-world Model reads stay honest; velocity_env_cfg.py:230-240)
    sim.expand_model_fields(("geom_friction",))
    g = torch.Generator(device="cpu").manual_seed(rank)
    self.foot = foot = [i for i, n in enumerate(m.names["geom"]) if ("foot" in n or "_toe" in n or "ankle_roll_link" in n) and n.endswith("_collision")]
    fr = 0.3 + 0.9 * torch.rand(E, len(foot), generator=g)
    M.geom_friction[:, foot, 0] = fr.to(dev)
    if int(overlap) >= 0:
      M.opt.stage_overlap = int(overlap)
    n0 = engine.launch_count()
    sim.create_graph(decimation=DECIMATION)
    assert engine.launch_count() - n0 >= 7 * DECIMATION
    # kernels per forward() / per step, counted on eager calls (the graphs replay the same launches)
    n1 = engine.launch_count(); engine.forward(M, D); n2 = engine.launch_count(); engine.step(M, D, 1); n3 = engine.launch_count()
    self.launches = {"forward": n2 - n1, "step": n3 - n2, "decimation": (n3 - n2) * DECIMATION}
    key = m.keys["robot/init_state"]
    self.q0 = torch.tensor(key["qpos"], dtype=torch.float32, device=dev)
    self.c0 = torch.tensor(key["ctrl"], dtype=torch.float32, device=dev)
    self.nbatch = 64
    self.ctrl_pool = torch.as_tensor(action_targets(m, self.nbatch, E, seed=1000 + rank), device=dev)
    origins = self.origins = torch.zeros(E, 3, device=dev)
    if "terrain_origins" in m.meta:     # curriculum spawn points of the box terrain (terrain_importer.py:203-223)
      from asimov_mjlab_b200 import terrains
      eo, _, _ = terrains.env_origins_curriculum(E, np.asarray(m.meta["terrain_origins"]), seed=rank)
      origins[:] = torch.as_tensor(eo, dtype=torch.float32, device=dev)
      origins[:, 0:2] += (torch.rand(E, 2, generator=g) - 0.5).to(dev) * 2.0     # off the platform centre, onto the steps
      # spawn ON the terrain: highest surface under the robot's footprint (the generator's own hfield-pyramid origins
      # lie inside the field's slab, heightfield_terrains.py:243-246)
      xy = origins[:, 0:2].cpu().numpy()
      surf = np.max([terrains.surface_height(m, xy + np.array(o)) for o in ((0, 0), (0.3, 0.2), (0.3, -0.2), (-0.3, 0.2), (-0.3, -0.2))], axis=0)
      origins[:, 2] = torch.as_tensor(surf, dtype=torch.float32, device=dev) + 0.05
    else:
      side = int(np.ceil(np.sqrt(E)))
      idx = torch.arange(E, device=dev)
      origins[:, 0] = (idx // side).float() * 2.5
      origins[:, 1] = (idx % side).float() * 2.5
    self.noise = 0.02 * torch.randn(E, m.nq - 7, generator=g).to(dev)

  def reset_all(self):
    D = self.D
    D.qpos[:] = self.q0
    D.qpos[:, 0:3] += self.origins
    D.qpos[:, 7:] += self.noise
    D.qvel.zero_(); D.qacc_warmstart.zero_(); D.time.zero_()
    D.ctrl[:] = self.c0
    self.sim.forward()

  def one_step(self, i):
    if i % RESET_PERIOD_STEPS == 0:
      self.reset_all()
    self.D.ctrl.copy_(self.ctrl_pool[i % self.nbatch])
    self.sim.step_decimation()

  def timed(self, steps, warmup, first=0, barrier=None):
    """W warm-up steps, then exactly K timed steps between CUDA events; step indices start at `first` (a multiple of the reset
    period resets first).  Returns (milliseconds, resets inside the timed region)."""
    torch = self.torch
    for i in range(warmup):
      self.one_step(first + i)
    if barrier:
      barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n_resets = 0
    if barrier:
      barrier()
    e0.record()
    for i in range(steps):
      n_resets += int((first + i) % RESET_PERIOD_STEPS == 0)
      self.one_step(first + i)
    e1.record()
    if barrier:
      barrier()
    else:
      torch.cuda.synchronize()
    return e0.elapsed_time(e1), n_resets

  def rate(self, ms, steps, world=1):
    return world * self.E * DECIMATION * steps / (ms * 1e-3)


def extra_measurements(args, world, rank, dev, pb):
  """Driver-visible numbers for the other BASELINE.json configs, OUTSIDE the headline's timed region: the two state regimes of
  the headline workload, configs[2] (Go1 rough, box and height-field terrain), configs[3] (G1 tracking env step) and configs[4]
  (PPO rollout + update at --gpus N, the one place a collective is on the path)."""
  import torch
  import torch.distributed as dist
  ex = {}
  # ---- regimes of the headline workload (same sim): upright right after a reset vs. fallen after 150 random-action steps
  ms, _ = pb.timed(20, 3, first=0)
  up = pb.rate(ms, 20)
  for i in range(23, 150):
    pb.one_step(i)
  ms, _ = pb.timed(20, 0, first=150)
  ex["regimes"] = {"upright_first_80_substeps_after_reset": up, "fallen_substeps_600_to_680": pb.rate(ms, 20),
                   "unit": "physics sub-steps/s per GPU", "note": "robots driven by random joint targets stand for ~40 control steps (two loaded feet: ~50 constraint rows, ~5.4 Newton iterations), then lie on the ground (~15 rows, ~3.3 iterations)"}
  if rank == 0 and world == 1:
    wl = {}
    for name in ("g1_flat_origins", "go1_rough", "go1_rough_hf", "asimov_flat", "asimov_toe_flat", "asimov_toe_full_flat"):
      try:
        p2 = PhysBench(name, args.envs, dev, rank, args.ls_parallel)
        ms, _ = p2.timed(40, 5)
        wl[name] = {"value": p2.rate(ms, 40), "unit": "physics sub-steps/s", "envs": args.envs, "steps": 40, "workload": WORKLOADS[name][1]}
        del p2
      except Exception as e:      # an extra must never take the headline line down
        wl[name] = {"error": repr(e)[:200]}
    ex["workloads"] = wl
  # ---- env step of the tracking task (configs[3]) and PPO on the velocity task (configs[4]); all ranks take part
  try:
    from asimov_mjlab_b200.rl.ppo import PPO
    from asimov_mjlab_b200.rl.tracking_env import G1TrackingFlatEnv
    from asimov_mjlab_b200.rl.velocity_env import G1VelocityFlatEnv
    if world == 1:
      env = G1TrackingFlatEnv(args.envs, dev, seed=rank)
      act = torch.zeros(args.envs, env.num_actions, device=dev)
      for _ in range(5):
        env.step(act)
      e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
      e0.record()
      for k in range(30):
        act.uniform_(-1.0, 1.0); env.step(act)
      e1.record(); torch.cuda.synchronize()
      ex.setdefault("workloads", {})["g1_tracking_env_step"] = {
        "value": args.envs * 30 / (e0.elapsed_time(e1) * 1e-3), "unit": "env-steps/s (4 physics sub-steps + fused manager terms each, random actions)",
        "envs": args.envs, "steps": 30, "workload": "Mjlab-Tracking-Flat-Unitree-G1, synthetic motion (BASELINE.json configs[3])"}
      del env
    env = G1VelocityFlatEnv(args.envs, dev, seed=rank)
    algo = PPO(env, torch.device(dev), steps_per_env=24, seed=0)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    t_roll = t_upd = 0.0
    iters, warm = 3, 1
    for it in range(warm + iters):
      if it == warm and world > 1:
        dist.barrier()
      torch.cuda.synchronize()
      ev[0].record(); algo.rollout(); ev[1].record(); algo.update(); ev[2].record()
      torch.cuda.synchronize()
      if it >= warm:
        t_roll += ev[0].elapsed_time(ev[1]); t_upd += ev[1].elapsed_time(ev[2])
    t = torch.tensor([t_roll, t_upd], device=dev)
    if world > 1:
      dist.all_reduce(t, op=dist.ReduceOp.MAX)
    n = iters * 24 * args.envs * world
    ex["ppo"] = {"value": n / (float(t[0] + t[1]) * 1e-3), "unit": "env-steps/s over all GPUs (rollout + update; 4 physics sub-steps per env-step)",
                 "rollout_env_steps_per_s": n / (float(t[0]) * 1e-3), "rollout_ms_per_iter": float(t[0]) / iters, "update_ms_per_iter": float(t[1]) / iters,
                 "envs_per_gpu": args.envs, "n_gpus": world, "iters": iters, "steps_per_env": 24, "epochs": 5, "mini_batches": 4,
                 "collectives": "per mini-batch ONE NCCL all-reduce of the flattened gradients + KL (0.44 M floats, 20 per iteration); per iteration: advantage moments, observation-normaliser statistics, episode statistics",
                 "workload": "Mjlab-Velocity-Flat-Unitree-G1, rsl_rl PPO hyper-parameters (BASELINE.json configs[4])"}
  except Exception as e:
    ex["ppo"] = {"error": repr(e)[:300]}
  return ex


def run_b200(args):
  import torch
  import torch.distributed as dist
  from asimov_mjlab_b200 import engine

  world = int(os.environ.get("WORLD_SIZE", "1"))
  rank = int(os.environ.get("RANK", "0"))
  local = int(os.environ.get("LOCAL_RANK", "0"))
  if not torch.cuda.is_available():
    raise SystemExit("bench.py: no CUDA device; the B200 arm has no CPU fallback (use --impl reference for the CPU arm)")
  torch.cuda.set_device(local)
  dev = f"cuda:{local}"
  if world > 1:
    dist.init_process_group("nccl", device_id=torch.device(dev))
  E = args.envs
  pb = PhysBench(args.workload, E, dev, rank, args.ls_parallel, args.cone, args.overlap)
  m, M, D = pb.m, pb.M, pb.D
  per_graph = pb.launches

  def barrier():
    if world > 1:
      dist.barrier()
    torch.cuda.synchronize()

  # ---- headline: W warm-up steps, then exactly K timed steps (CUDA events, barrier + synchronize on both sides)
  sampler = ClockSampler(local)
  if rank == 0:
    sampler.start()
  ms, n_resets = pb.timed(args.steps, args.warmup, first=0, barrier=barrier)
  clocks = sampler.stop() if rank == 0 else None
  # ---- e2e: same work through the host-buffer C-ABI call, pinned host buffers, copies inside the timed region
  ctrl_host = torch.as_tensor(action_targets(m, 8, E, seed=2000 + rank)).pin_memory()
  qpos_h = torch.empty(E, m.nq).pin_memory(); qvel_h = torch.empty(E, m.nv).pin_memory(); sens_h = torch.empty(E, m.nsensordata).pin_memory()
  pb.reset_all()
  for i in range(max(3, args.warmup)):
    engine.step_host(M, D, ctrl_host[i % 8], DECIMATION, qpos_h, qvel_h, sens_h)
  pb.reset_all()
  barrier()
  e2e_steps = args.steps
  t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
  t0.record()
  for i in range(e2e_steps):
    if i % RESET_PERIOD_STEPS == 0 and i:
      pb.reset_all()
    engine.step_host(M, D, ctrl_host[i % 8], DECIMATION, qpos_h, qvel_h, sens_h)
  t1.record()
  barrier()
  ms_e2e = t0.elapsed_time(t1)
  # ---- per-kernel shares (CUDA events between launches, live, outside the timed region), in the state the timed region ends in
  pb.reset_all()
  for i in range(min(args.steps, 40)):
    pb.one_step(i + 1)
  torch.cuda.synchronize()
  stages = engine.profile_step(M, D, 20)
  extras = extra_measurements(args, world, rank, dev, pb) if args.extras else {}
  # ---- max over ranks
  rank_ms = [ms / args.steps]
  if world > 1:
    t = torch.tensor([ms, ms_e2e], device=dev)
    allt = [torch.zeros_like(t) for _ in range(world)]
    dist.all_gather(allt, t)
    rank_ms = [float(x[0]) / args.steps for x in allt]
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, ms_e2e = float(t[0]), float(t[1])
  if rank != 0:
    if world > 1:
      dist.destroy_process_group()
    return
  total_sub = world * E * DECIMATION * args.steps
  value = total_sub / (ms * 1e-3)
  e2e_value = world * E * DECIMATION * e2e_steps / (ms_e2e * 1e-3)
  peak, peak_src = measured_peak()
  per_gpu_substep_s = (ms * 1e-3) / (args.steps * DECIMATION)
  # the per-env origin sites are static and stored once for all worlds: they add no per-world bytes
  ab = ALG_BYTES_PER_SUBSTEP_WORLD if args.workload in ("g1_flat", "g1_flat_origins") else alg_bytes(m, len(pb.foot))
  achieved = E * ab / per_gpu_substep_s / 1e9
  tot = sum(stages.values())
  top = max(stages, key=stages.get)
  traffic = NCU_TRAFFIC.get((args.workload, E), (None, None))
  out = {
    "metric": "env_steps_per_s", "value": value, "unit": "physics sub-steps/s (1 env-step = 1 sim.step per world)",
    "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
    "rank_ms_per_step": rank_ms,
    "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
    "config": {"workload": WORKLOADS[args.workload][1],
               "envs_per_gpu": E, "decimation": DECIMATION, "substeps_per_step": DECIMATION, "timestep": 0.005,
               "nconmax": 50, "njmax": 300, "iterations": 10, "ls_iterations": 20, "ls_parallel": bool(args.ls_parallel), "cone": args.cone,
               "friction_dr": True, "resets_in_timed_region": n_resets, "parallelism": f"env-shard x{world}, no collective on the step path",
               "state_window": f"sub-steps 0..{min(args.steps, RESET_PERIOD_STEPS) * DECIMATION} after a reset (see `regimes`: upright robots cost more than fallen ones)",
               "l2": f"working set {D.bytes / 2**20:.0f} MiB per GPU (> 126 MB L2); no explicit flush"},
    "e2e": {"value": e2e_value, "unit": "physics sub-steps/s", "h2d_bytes_per_step": int(E * m.nu * 4),
            "d2h_bytes_per_step": int(E * (m.nq + m.nv + m.nsensordata) * 4), "ms_per_step": ms_e2e / e2e_steps,
            "api": "b2s_step_host (C ABI) with pinned host buffers"},
    "gpu_launches": int(args.steps * per_graph["decimation"] + n_resets * per_graph["forward"]),
    "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                 "traffic": traffic[0], "traffic_from_profile": traffic[1],
                 "algorithmic_bytes_per_launch": E * ab,
                 "peak_source": peak_src, "scope": f"whole sub-step (7 stage kernels); algorithmic bytes {ab} B per world per sub-step",
                 "dominant_kernel": top, "dominant_share": stages[top] / tot, "dominant_kernel_ms": stages[top],
                 # the same algorithmic bytes over the dominant kernel's own launch duration (CUDA events around that launch)
                 "achieved_over_dominant_kernel_only": E * ab / (stages[top] * 1e-3) / 1e9,
                 "stage_ms": stages},
    "clocks": clocks,
  }
  out.update(extras)
  if args.cpu_baseline and world == 1:
    cs = CpuSampler(m, args.ls_parallel)
    window = min(args.steps, RESET_PERIOD_STEPS) * DECIMATION
    rate, t, ep = cs.run(window, min_seconds=8.0)
    out["cpu_baseline"] = {"value": rate, "unit": "physics sub-steps/s", "cores": cs.cores, "kind": "port",
                           "sample": cpu_sample_desc(cs, window, t, ep)}
  print(json.dumps(out), flush=True)
  if world > 1:
    dist.destroy_process_group()


if __name__ == "__main__":
  a = parse()
  if a.impl == "reference":
    run_reference(a)
  else:
    run_b200(a)
