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
# DRAM traffic of one sub-step (7 stage kernels), dram__bytes_read.sum + dram__bytes_write.sum from the `ncu --set full`
# capture profiles/step_r1k_metrics.txt (G1-flat, 4096 worlds): 118 477 568 B read + 978 432 B written
NCU_TRAFFIC = {("g1_flat", 4096): (119456000, "profiles/step_r1k_metrics.txt (ncu --set full, sum over the 7 stage kernels of one sub-step)")}
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
  ap.add_argument("--overlap", type=int, default=1, help="stage overlap on forked graph branches (0: serial stages)")
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


def load_model(ls_parallel=1, workload="g1_flat", cone="pyramidal"):
  from asimov_mjlab_b200 import scenes
  m = scenes.load_asset(WORKLOADS[workload][0])
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
  scale = scenes.g1_action_scale(jn) if m.nu == 29 else scenes.go1_action_scale(jn)
  return (np.asarray(key["ctrl"], np.float32)[None, None, :] + scale.astype(np.float32)[None, None, :] * a)


# ------------------------------------------------------------------------------------------- reference arm (CPU)
def cpu_rollout(m, nworld, nsub_total, nthread, seed=0, ls_parallel=1):
  """Oracle (CPU restatement of the reference's step) on the host cores: nworld worlds x nsub_total sub-steps,
  actions resampled every DECIMATION sub-steps.  Returns sub-steps/s.  (bench.py may execute oracle/ only here.)"""
  sys.path.insert(0, os.path.join(ROOT, "oracle"))
  from oracle import Oracle
  orc = Oracle(m, "f32", 50, 300, ls_parallel=ls_parallel)
  key = m.keys["robot/init_state"]
  worlds = [orc.new_world(key["qpos"], None, key["ctrl"]) for _ in range(nworld)]
  nb = nsub_total // DECIMATION
  ctrl = action_targets(m, nb, nworld, seed)
  t0 = time.perf_counter()
  for b in range(nb):
    for w in range(nworld):
      worlds[w]["ctrl"][:] = ctrl[b, w]
    orc.rollout(worlds, DECIMATION, nthread)
  dt = time.perf_counter() - t0
  return nworld * nb * DECIMATION / dt, dt


def run_reference(args):
  rank = int(os.environ.get("RANK", "0"))
  if rank != 0:
    return
  m = load_model(args.ls_parallel, args.workload, args.cone)
  cores = os.cpu_count() or 1
  nworld = 16 * cores
  # warm-up then K timed steps, each a bounded sample: one decimation of `nworld` worlds
  cpu_rollout(m, nworld, DECIMATION * max(args.warmup, 1), cores)
  rate, dt = cpu_rollout(m, nworld, DECIMATION * args.steps, cores)
  sample = f"{nworld} worlds x {DECIMATION * args.steps} sub-steps, oracle f32 (CPU restatement of the step), OpenMP {cores} threads"
  out = {
    "impl": "reference", "metric": "env_steps_per_s", "value": rate, "unit": "physics sub-steps/s (1 env-step = 1 sim.step per world)",
    "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
    "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
    "config": {"workload": WORKLOADS[args.workload][1], "envs_per_step": nworld, "decimation": DECIMATION},
    "cpu_baseline": {"value": rate, "unit": "physics sub-steps/s", "cores": cores, "kind": "port", "sample": sample},
    "e2e": {"value": rate, "unit": "physics sub-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
  }
  print(json.dumps(out), flush=True)


# ------------------------------------------------------------------------------------------------ B200 arm
def run_b200(args):
  import numpy as np
  import torch
  import torch.distributed as dist
  from asimov_mjlab_b200 import engine
  from asimov_mjlab_b200.sim import MujocoCfg, Simulation, SimulationCfg

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
  m = load_model(args.ls_parallel, args.workload, args.cone)
  cfg = SimulationCfg(nconmax=50, njmax=300, ls_parallel=bool(args.ls_parallel),
                      mujoco=MujocoCfg(timestep=0.005, iterations=10, ls_iterations=20, cone=args.cone))
  sim = Simulation(E, cfg, m, dev)
  M, D = sim.model, sim.data
  # foot-friction domain randomisation ON (per-world Model reads stay honest; velocity_env_cfg.py:230-240)
  sim.expand_model_fields(("geom_friction",))
  g = torch.Generator(device="cpu").manual_seed(rank)
  foot = [i for i, n in enumerate(m.names["geom"]) if ("foot" in n or "_toe" in n) and n.endswith("_collision")]
  fr = 0.3 + 0.9 * torch.rand(E, len(foot), generator=g)
  M.geom_friction[:, foot, 0] = fr.to(dev)
  M.opt.stage_overlap = int(args.overlap)
  n0 = engine.launch_count()
  sim.create_graph(decimation=DECIMATION)
  per_graph = {"step": 7, "forward": 6, "decimation": 7 * DECIMATION}
  assert engine.launch_count() - n0 >= per_graph["decimation"]
  key = m.keys["robot/init_state"]
  q0 = torch.tensor(key["qpos"], dtype=torch.float32, device=dev)
  c0 = torch.tensor(key["ctrl"], dtype=torch.float32, device=dev)
  nbatch = 64
  ctrl_pool = torch.as_tensor(action_targets(m, nbatch, E, seed=1000 + rank), device=dev)
  origins = torch.zeros(E, 3, device=dev)
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
  noise = 0.02 * torch.randn(E, m.nq - 7, generator=g).to(dev)

  def reset_all():
    D.qpos[:] = q0
    D.qpos[:, 0:3] += origins
    D.qpos[:, 7:] += noise
    D.qvel.zero_(); D.qacc_warmstart.zero_(); D.time.zero_()
    D.ctrl[:] = c0
    sim.forward()

  def one_step(i):
    if i % RESET_PERIOD_STEPS == 0:
      reset_all()
    D.ctrl.copy_(ctrl_pool[i % nbatch])
    sim.step_decimation()

  def barrier():
    if world > 1:
      dist.barrier()
    torch.cuda.synchronize()

  for i in range(args.warmup):
    one_step(i)
  barrier()
  sampler = ClockSampler(local)
  if rank == 0:
    sampler.start()
  e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
  n_resets = 0
  barrier()
  e0.record()
  for i in range(args.steps):
    n_resets += int(i % RESET_PERIOD_STEPS == 0)
    one_step(i)
  e1.record()
  barrier()
  ms = e0.elapsed_time(e1)
  clocks = sampler.stop() if rank == 0 else None
  # ---- e2e: same work through the host-buffer C-ABI call, pinned host buffers, copies inside the timed region
  ctrl_host = torch.as_tensor(action_targets(m, 8, E, seed=2000 + rank)).pin_memory()
  qpos_h = torch.empty(E, m.nq).pin_memory(); qvel_h = torch.empty(E, m.nv).pin_memory(); sens_h = torch.empty(E, m.nsensordata).pin_memory()
  reset_all()
  for i in range(max(3, args.warmup // 4)):
    engine.step_host(M, D, ctrl_host[i % 8], DECIMATION, qpos_h, qvel_h, sens_h)
  barrier()
  e2e_steps = args.steps
  t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
  t0.record()
  for i in range(e2e_steps):
    if i % RESET_PERIOD_STEPS == 0 and i:
      reset_all()
    engine.step_host(M, D, ctrl_host[i % 8], DECIMATION, qpos_h, qvel_h, sens_h)
  t1.record()
  barrier()
  ms_e2e = t0.elapsed_time(t1)
  # ---- per-kernel shares (CUDA events between launches, live, outside the timed region)
  reset_all()
  for i in range(40):
    one_step(i + 1)
  torch.cuda.synchronize()
  stages = engine.profile_step(M, D, 20)
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
  ab = ALG_BYTES_PER_SUBSTEP_WORLD if args.workload == "g1_flat" else alg_bytes(m, len(foot))
  achieved = E * ab / per_gpu_substep_s / 1e9
  tot = sum(stages.values())
  top = max(stages, key=stages.get)
  out = {
    "metric": "env_steps_per_s", "value": value, "unit": "physics sub-steps/s (1 env-step = 1 sim.step per world)",
    "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
    "rank_ms_per_step": rank_ms,
    "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
    "config": {"workload": WORKLOADS[args.workload][1],
               "envs_per_gpu": E, "decimation": DECIMATION, "substeps_per_step": DECIMATION, "timestep": 0.005,
               "nconmax": 50, "njmax": 300, "iterations": 10, "ls_iterations": 20, "ls_parallel": bool(args.ls_parallel), "cone": args.cone,
               "friction_dr": True, "resets_in_timed_region": n_resets, "parallelism": f"env-shard x{world}, no collective on the step path",
               "l2": f"working set {D.bytes / 2**20:.0f} MiB per GPU (> 126 MB L2); no explicit flush"},
    "e2e": {"value": e2e_value, "unit": "physics sub-steps/s", "h2d_bytes_per_step": int(E * m.nu * 4),
            "d2h_bytes_per_step": int(E * (m.nq + m.nv + m.nsensordata) * 4), "ms_per_step": ms_e2e / e2e_steps,
            "api": "b2s_step_host (C ABI) with pinned host buffers"},
    "gpu_launches": int(args.steps * per_graph["decimation"] + n_resets * per_graph["forward"]),
    "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                 "traffic": NCU_TRAFFIC.get((args.workload, E), (None, None))[0], "traffic_source": NCU_TRAFFIC.get((args.workload, E), (None, None))[1],
                 "algorithmic_bytes_per_launch": E * ab,
                 "peak_source": peak_src, "scope": f"whole sub-step (7 stage kernels); algorithmic bytes {ab} B per world per sub-step",
                 "dominant_kernel": top, "dominant_share": stages[top] / tot, "dominant_kernel_ms": stages[top],
                 # the same algorithmic bytes over the dominant kernel's own launch duration (CUDA events around that launch)
                 "achieved_over_dominant_kernel_only": E * ab / (stages[top] * 1e-3) / 1e9,
                 "stage_ms": stages},
    "clocks": clocks,
  }
  if args.cpu_baseline and world == 1:
    cores = os.cpu_count() or 1
    # bounded sample sized for roughly 10-20 s of host work: a short probe sets the number of sub-steps
    nw = 16 * cores
    probe, _ = cpu_rollout(m, nw, 2 * DECIMATION, cores, ls_parallel=args.ls_parallel)
    nsub = int(min(max(12.0 * probe / nw, 2 * DECIMATION), 4000)) // DECIMATION * DECIMATION
    rate, dt = cpu_rollout(m, nw, nsub, cores, ls_parallel=args.ls_parallel)
    out["cpu_baseline"] = {"value": rate, "unit": "physics sub-steps/s", "cores": cores, "kind": "port",
                           "sample": f"{nw} worlds x {nsub} sub-steps, oracle f32 with OpenMP on {cores} threads ({dt:.1f} s)"}
  print(json.dumps(out), flush=True)
  if world > 1:
    dist.destroy_process_group()


if __name__ == "__main__":
  a = parse()
  if a.impl == "reference":
    run_reference(a)
  else:
    run_b200(a)
