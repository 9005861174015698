#!/usr/bin/env python
"""bench.py — body-substeps/s of the physics substep loop on B200 (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]

One "step" = one 60 Hz frame = 8 substeps over the whole scene.
  N = 1   workload pile100k  (BASELINE.json configs[2], the config north_star's target is quoted
          on): 100 000 mixed capsule / rounded-rect / circle bodies with friction + restitution.
          The pile is first stepped `--settle` frames (default 240, SURVEY.md §8d: timing window
          is the settled pile, contacts dominant); those frames are not timed.
  N > 1   workload worlds512 (configs[4]) AS STATED: 8192 independent 512-body worlds partitioned
          over the N GPUs (8192 / N per GPU, "strong" scaling), world ranges partitioned over
          ranks, no data-path collective.  The line carries its own N = 1 point (`n1_value`: rank 0
          alone steps all 8192 worlds while the others wait) and `efficiency`.
          --worlds-per-gpu W runs the weak variant (W worlds on every GPU) instead.
Prints ONE JSON line on rank 0.  value = units all ranks processed / max-over-ranks device time.

--impl reference: the reference's own CPU step cannot be built or run (tree not mounted, no Rust
toolchain), so this arm times the CPU oracle (oracle/libsfp_ref.so, "kind":
"port") on the SAME config — the full 100 003-body pile, settled on the CPU — with all host
threads in the solver loops.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

DT, SUBSTEPS = 1.0 / 60.0, 8
TOTAL_WORLDS, BODIES_PER_WORLD = 8192, 512


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """samples nvidia-smi clocks + throttle reasons during the timed region"""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.rows = []
        self._proc = None
        self._t = threading.Thread(target=self._run, daemon=True)

    def _run(self):
        # one long-lived nvidia-smi streaming a line every 50 ms (start-up of a fresh
        # nvidia-smi per sample costs more than the whole timed region)
        try:
            self._proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-i", str(self.index), "-lms", "50"], stdout=subprocess.PIPE, text=True)
            for line in self._proc.stdout:
                parts = [x.strip() for x in line.strip().split(",")]
                if len(parts) >= 6:
                    self.rows.append(parts)
        except Exception:
            pass

    def _once(self):
        try:
            out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                  "-i", str(self.index)], capture_output=True, text=True, timeout=10).stdout
            parts = [x.strip() for x in out.strip().split(",")]
            return parts if len(parts) >= 6 else None
        except Exception:
            return None

    def __enter__(self):
        # a sample right before the timed region (the GPU is busy with the warm-up), the stream
        # during it, one right after: a region shorter than the 50 ms period is still bracketed
        self.before = self._once()
        self._t.start()
        time.sleep(0.25)  # let the first samples arrive before the timed region starts
        self.rows.clear()
        return self

    def __exit__(self, *a):
        self.after = self._once()
        try:
            if self._proc:
                self._proc.terminate()
        except Exception:
            pass
        self._t.join(timeout=3)
        self.during = len(self.rows)
        for r in (getattr(self, "before", None), self.after):
            if r:
                self.rows.append(r)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        sm = sorted(float(r[0]) for r in self.rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for k, n in enumerate(names) if any(r[2 + k].lower() == "active" for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(self.rows[0][1]), "reasons": reasons,
                "samples": len(self.rows), "samples_during": getattr(self, "during", 0),
                "note": "streamed every 50 ms during the timed region, plus one sample right before "
                        "and one right after it"}


def ncu_traffic(workload, worlds=None):
    """DRAM bytes per launch of the dominant kernel, from the committed ncu capture (or None);
    the batched kernel's capture is scaled to the number of worlds this launch steps"""
    p = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    try:
        with open(p) as f:
            e = json.load(f).get(workload, {})
        b = e.get("bytes_per_launch")
        if b is not None and worlds and e.get("worlds_per_launch"):
            b = b * worlds / e["worlds_per_launch"]
        return b
    except Exception:
        return None


def b_alg(n, pairs, contacts, joints, substeps=SUBSTEPS, whole_step=True):
    """algorithmic bytes per body-substep (SURVEY.md §8d / BASELINE.md §3)"""
    p, c, j = pairs / n, contacts / n, joints / n
    b = 136.0 + 72.0 * p + 344.0 * c + 116.0 * j
    if whole_step:
        b += (84.0 + 24.0 * p) / substeps
    return b


def stage_rooflines(n, st, br_ms, stage_ns, peak, substeps=SUBSTEPS):
    """Algorithmic-bytes fraction of the measured HBM peak per stage of one frame (SURVEY.md §8a/d:
    every stage reads its inputs once and writes its outputs once).  br_ms: CUDA-event times of
    broadphase / colouring (ms per frame); stage_ns: sfp_last_step_stage_ns (ns per frame)."""
    p, c, j = st["n_pairs"] / n, st["n_contacts"] / n, st["n_joints"] / n
    per_frame = {  # bytes per body per FRAME
        "broadphase": 84.0 + 8.0 * p,               # A2: 48 in + 8 keys + 16 sort + 12 ranges + 8 p out
        "colouring": 16.0 * p,                      # A8: pairs in, colour-sorted pairs out
        "integrate": 88.0 * substeps,               # A3
        "narrowphase": (72.0 * p + 48.0 * c) * substeps,        # A4
        "position": (168.0 * c + 116.0 * j) * substeps,         # A5 + A6
        "derive": 48.0 * substeps,                  # A7 i
        "velocity": 128.0 * c * substeps,           # A7 ii
    }
    secs = {"broadphase": br_ms["broadphase"] * 1e-3, "colouring": br_ms["colour"] * 1e-3}
    small = stage_ns.get("small_colours", 0) * 1e-9  # both passes' small colours: split evenly
    for k in ("integrate", "narrowphase", "position", "derive", "velocity"):
        secs[k] = stage_ns[k] * 1e-9 + (0.5 * small if k in ("position", "velocity") else 0.0)
    out = {}
    for k, b in per_frame.items():
        if secs[k] > 0:
            gbs = b * n / secs[k] / 1e9
            out[k] = {"us_per_frame": secs[k] * 1e6, "algorithmic_gbytes_per_s": gbs, "frac": gbs / peak}
    return out


def shuffle_scene(bodies, joints, seed=20261107):
    """the same scene with the body array in a random order (joints re-indexed)"""
    perm = np.random.default_rng(seed).permutation(len(bodies))
    inv = np.empty_like(perm)
    inv[perm] = np.arange(len(bodies))
    j = joints.copy()
    if len(j):
        j["body_a"] = inv[joints["body_a"]]
        j["body_b"] = inv[joints["body_b"]]
    return bodies[perm], j


def make_scene(workload: str):
    from starframe_b200 import scenes

    if workload == "pile100k":
        return scenes.mixed_pile(100_000)
    if workload == "circles10k":
        return scenes.circles_container(10_000)
    if workload == "ropes2k":
        return scenes.ropes()
    if workload == "stack100":
        return scenes.stack100()
    if workload.startswith("polys"):  # a mixed pile in which half of the bodies are general polygons
        return scenes.polygon_pile(int(workload[5:] or 100_000))
    if workload.startswith("shelves"):
        return scenes.shelves(int(workload[7:] or 4_000_000))
    if workload.startswith("pile"):
        return scenes.mixed_pile(int(workload[4:]))
    raise SystemExit(f"unknown workload {workload}")


# ------------------------------------------------------------------------------------------
# single scene on one GPU
# ------------------------------------------------------------------------------------------
def run_single(args, rank, world_size):
    import torch

    import starframe_b200 as sf

    bodies, joints = make_scene(args.workload)
    n = len(bodies)
    api = sf.product_api()
    w = sf.World(api, device=torch.cuda.current_device())
    w.upload_bodies(bodies)
    w.set_joints(joints)
    # SURVEY.md §8d also asks for the free-fall window, frames [0,60): broadphase dominant
    ff_frames, ff_ms, ff_pairs = min(60, args.settle), 0.0, 0
    for k in range(args.settle):
        w.step(DT, SUBSTEPS)
        if k < ff_frames:
            w.sync()
            ff_ms += w.last_step_ms()
            if k == ff_frames - 1:
                ff_pairs = w.stats()["n_pairs"]
    w.sync()
    settled = w.download_bodies()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def one_frame():
        flush.fill_(1)  # L2 flush between frames (outside the timed window)
        torch.cuda.synchronize()
        w.step(DT, SUBSTEPS)
        w.sync()
        return w.last_step_breakdown_ms()

    for _ in range(args.warmup):
        one_frame()
    barrier(world_size)
    torch.cuda.synchronize()
    t_dev = t_sub = t_bp = t_col = 0.0
    stage_acc = {}
    with ClockSampler(torch.cuda.current_device()) as clk:
        t0 = time.perf_counter()
        for _ in range(args.steps):
            br = one_frame()
            t_dev += br["step"]
            t_sub += br["substeps"]
            t_bp += br["broadphase"]
            t_col += br["colour"]
            for k, v in w.last_step_stage_ns().items():
                stage_acc[k] = stage_acc.get(k, 0) + v / args.steps
        torch.cuda.synchronize()
        wall = time.perf_counter() - t0
    barrier(world_size)
    st = w.stats()

    # end to end through the public call with host buffers (H2D + step + D2H inside the window)
    # (the engine keeps its body states in two page-locked arrays from sfp_alloc_host_states)
    buf = [api.alloc_host_states(n), api.alloc_host_states(n)]
    buf[0][:] = w.download_bodies()
    e2e_steps = max(3, min(args.steps, 30))
    w.step_host(buf[0], DT, SUBSTEPS, out=buf[1])
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for k in range(e2e_steps):
        w.step_host(buf[(k + 1) & 1], DT, SUBSTEPS, out=buf[k & 1])
    e2e_s = time.perf_counter() - t0

    t_dev_max = allreduce_max(t_dev, world_size)
    units = float(n) * SUBSTEPS * args.steps * world_size
    value = units / (t_dev_max * 1e-3)
    peak, peak_src = peaks()
    bytes_sub = b_alg(n, st["n_pairs"], st["n_contacts"], st["n_joints"], whole_step=False) * n * SUBSTEPS
    ach = bytes_sub / (t_sub / args.steps * 1e-3) / 1e9
    bytes_step = b_alg(n, st["n_pairs"], st["n_contacts"], st["n_joints"]) * n * SUBSTEPS
    out = {
        "metric": "body-substeps/sec", "value": value, "unit": "body-substeps/s",
        "n_gpus": world_size, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": t_dev_max / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {
            "workload": args.workload if world_size == 1 else f"{world_size} replicas of {args.workload}",
            "bodies": n, "dt": "1/60", "substeps": SUBSTEPS, "settle_frames": args.settle,
            "pairs_per_body": st["n_pairs"] / n, "contacts_per_body": st["n_contacts"] / n,
            "joints_per_body": st["n_joints"] / n, "colours": st["n_colours"],
            "colour_rounds": st["colour_rounds"],
            "l2": "256 MiB device fill between frames, outside the per-frame CUDA-event window; "
                  "the 100k-body working set (~30 MB) is L2-resident within a frame by design",
            "timing": "sum of per-frame CUDA-event times on the library's stream, max over ranks",
        },
        "wall_ms_per_step_incl_flush": wall / args.steps * 1e3,
        "breakdown_ms": {"broadphase": t_bp / args.steps, "colouring": t_col / args.steps,
                         "substep_kernel": t_sub / args.steps},
        "roofline": {
            "bound": "hbm", "kernel": "k_substeps_coop (all 8 substeps of a frame, one launch)",
            "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
            "peak_source": peak_src, "traffic": ncu_traffic(args.workload),
            "peak_nominal": 8000.0, "frac_nominal": ach / 8000.0,
            "algorithmic_bytes_per_launch": bytes_sub,
            "note": "algorithmic bytes = (136 + 72 p + 344 c + 116 j) * N * S, SURVEY.md §8d; "
                    "the kernel is grid-sync latency bound, not bandwidth bound (DESIGN.md)",
        },
        "roofline_stages": stage_rooflines(n, st, {"broadphase": t_bp / args.steps, "colour": t_col / args.steps},
                                           stage_acc, peak),
        "roofline_whole_step": {"achieved": bytes_step / (t_dev / args.steps * 1e-3) / 1e9,
                                "peak": peak, "unit": "GB/s",
                                "frac": bytes_step / (t_dev / args.steps * 1e-3) / 1e9 / peak},
        "e2e": {"value": float(n) * SUBSTEPS * e2e_steps / e2e_s, "unit": "body-substeps/s",
                "h2d_bytes_per_step": 32 * n, "d2h_bytes_per_step": 32 * n,
                "call": "sfp_step_host with host SfpBodyState arrays from sfp_alloc_host_states (pinned)",
                "steps": e2e_steps},
        "gpu_launches": int(st["launches"]) * args.steps,
        "launches_per_step": int(st["launches"]),
        "clocks": clk.summary(),
        "stats": {k: st[k] for k in ("n_pairs", "n_contacts", "n_manifolds", "max_penetration",
                                     "kinetic_energy")},
        "free_fall_window": {"frames": f"[0,{ff_frames})",
                             "value": float(n) * SUBSTEPS * ff_frames / (ff_ms * 1e-3) if ff_ms else None,
                             "unit": "body-substeps/s",
                             "ms_per_step": ff_ms / ff_frames if ff_frames else None,
                             "pairs_per_body_at_end": ff_pairs / n},
    }
    if rank == 0 and not args.no_cpu_baseline and world_size == 1:
        out["cpu_baseline"] = cpu_baseline_single(bodies, joints, settled, args.cpu_frames)
    if rank == 0 and world_size == 1 and not args.no_other_configs and args.workload == "pile100k":
        # the other single-GPU configs of BASELINE.json, measured the same way (short runs);
        # they are parity-test cases first, these lines are context for the headline
        del w
        out["other_configs"] = {name: quick_config(name, args) for name in ("circles10k", "ropes2k", "polys100000")}
        out["other_configs"]["worlds512_8192_1gpu"] = quick_worlds(args)
        # the same pile uploaded in a random body order (the library re-sorts internally)
        out["shuffled_order"] = quick_config(args.workload, args, shuffle=True)
        out["shuffled_order"]["vs_spawn_order"] = out["shuffled_order"]["ms_per_step"] / (t_dev / args.steps)
        if not args.no_scale_point:
            # the scale at which HBM bandwidth, not phase latency, bounds
            # the step (state + pair data several times the L2)
            out["scale_point"] = scale_point(args)
    return out


def quick_config(name, args, frames=30, shuffle=False):
    import torch

    import starframe_b200 as sf

    scene = make_scene(name)
    bodies, joints = scene[0], scene[1]
    polys = scene[2] if len(scene) > 2 else None
    if shuffle:
        bodies, joints = shuffle_scene(bodies, joints)
    w = sf.World(sf.product_api(), device=torch.cuda.current_device())
    if polys is not None:
        w.set_polygons(polys)
    w.upload_bodies(bodies)
    w.set_joints(joints)
    for _ in range(args.settle):
        w.step(DT, SUBSTEPS)
    w.sync()
    t = 0.0
    for _ in range(frames):
        w.step(DT, SUBSTEPS)
        w.sync()
        t += w.last_step_ms()
    st = w.stats()
    n = len(bodies)
    return {"bodies": n, "value": n * SUBSTEPS * frames / (t * 1e-3), "unit": "body-substeps/s",
            "ms_per_step": t / frames, "pairs_per_body": st["n_pairs"] / n,
            "contacts_per_body": st["n_contacts"] / n, "joints_per_body": st["n_joints"] / n,
            "colours": st["n_colours"]}


def scale_point(args, name="shelves4000000", settle=200, frames=10):
    import torch

    import starframe_b200 as sf

    bodies, joints = make_scene(name)
    n = len(bodies)
    peak, _ = peaks()

    def run(b, j):
        w = sf.World(sf.product_api(), device=torch.cuda.current_device())
        w.upload_bodies(b)
        for _ in range(settle):
            w.step(DT, SUBSTEPS)
        w.sync()
        acc = {"step": 0.0, "substeps": 0.0, "broadphase": 0.0, "colour": 0.0}
        stages = {}
        for _ in range(frames):
            w.step(DT, SUBSTEPS)
            w.sync()
            for k, v in w.last_step_breakdown_ms().items():
                acc[k] += v / frames
            for k, v in w.last_step_stage_ns().items():
                stages[k] = stages.get(k, 0) + v / frames
        return acc, stages, w.stats()

    acc, stages, st = run(bodies, joints)
    b_sub = b_alg(n, st["n_pairs"], st["n_contacts"], 0, whole_step=False) * n * SUBSTEPS
    b_all = b_alg(n, st["n_pairs"], st["n_contacts"], 0) * n * SUBSTEPS
    out = {"workload": name, "bodies": n, "value": n * SUBSTEPS / (acc["step"] * 1e-3),
           "unit": "body-substeps/s", "ms_per_step": acc["step"],
           "state_mbytes": n * 88 / 1e6, "pairs_per_body": st["n_pairs"] / n,
           "contacts_per_body": st["n_contacts"] / n, "colours": st["n_colours"],
           "breakdown_ms": {"broadphase": acc["broadphase"], "colouring": acc["colour"],
                            "substep_kernel": acc["substeps"]},
           "roofline_substep_kernel_frac": b_sub / (acc["substeps"] * 1e-3) / 1e9 / peak,
           "roofline_whole_step_frac": b_all / (acc["step"] * 1e-3) / 1e9 / peak,
           "roofline_stages": stage_rooflines(n, st, acc, stages, peak)}
    acc2, _, _ = run(*shuffle_scene(bodies, joints))
    out["shuffled_order"] = {"ms_per_step": acc2["step"], "vs_spawn_order": acc2["step"] / acc["step"]}
    return out


def quick_worlds(args, frames=10, nw=None):
    """config 5's N = 1 point: all 8192 worlds on one GPU"""
    import torch

    import starframe_b200 as sf
    from starframe_b200 import scenes

    nw = nw or (args.worlds_per_gpu or TOTAL_WORLDS)
    b = sf.Batch(sf.product_api(), nw, BODIES_PER_WORLD, device=torch.cuda.current_device())
    b.upload(scenes.worlds(nw, BODIES_PER_WORLD))
    for _ in range(args.settle):
        b.step(DT, SUBSTEPS)
    b.sync()
    t = 0.0
    for _ in range(frames):
        b.step(DT, SUBSTEPS)
        b.sync()
        t += b.last_step_ms()
    n = nw * BODIES_PER_WORLD
    return {"bodies": n, "worlds": nw, "value": n * SUBSTEPS * frames / (t * 1e-3),
            "unit": "body-substeps/s", "ms_per_step": t / frames,
            "note": "BASELINE.json configs[4] on ONE GPU: the N = 1 point of the strong-scaling curve"}


def cpu_baseline_single(bodies, joints, settled_state, frames):
    """the oracle in reference order (sequential Gauss-Seidel, 1 core) on the same settled state"""
    import starframe_b200 as sf
    from oracle import loader
    from starframe_b200 import _ffi

    ref = sf.World(loader.api(), solver_order=_ffi.ORDER_REFERENCE)
    ref.upload_bodies(bodies)
    ref.set_joints(joints)
    state = settled_state
    t0 = time.perf_counter()
    done = 0
    for _ in range(frames):
        state = ref.step_host(state, DT, SUBSTEPS)
        done += 1
        if time.perf_counter() - t0 > 30.0:
            break
    s = time.perf_counter() - t0
    return {"value": len(bodies) * SUBSTEPS * done / s, "unit": "body-substeps/s", "cores": 1,
            "kind": "port",
            "sample": f"{done} frames x {SUBSTEPS} substeps of the same scene from the same settled "
                      f"state ({s:.1f} s), sequential Gauss-Seidel in reference order; host has "
                      f"{os.cpu_count()} logical cores"}


# ------------------------------------------------------------------------------------------
# batched independent worlds, world range partitioned over ranks
# ------------------------------------------------------------------------------------------
def run_worlds(args, rank, world_size):
    import torch

    import starframe_b200 as sf
    from starframe_b200 import scenes

    # config 5 as BASELINE.json states it: 8192 worlds partitioned over the GPUs (strong scaling);
    # --worlds-per-gpu W: W worlds on every GPU instead (weak scaling)
    strong = args.worlds_per_gpu is None
    total = TOTAL_WORLDS if strong else args.worlds_per_gpu * world_size
    wr = world_range(rank, world_size, total)
    nw = len(wr)
    bodies = scenes.worlds(nw, BODIES_PER_WORLD, first_world=wr.start)
    api = sf.product_api()
    b = sf.Batch(api, nw, BODIES_PER_WORLD, device=torch.cuda.current_device())
    b.upload(bodies)
    for _ in range(args.settle):
        b.step(DT, SUBSTEPS)
    b.sync()
    # the curve's own N = 1 point, measured in the same run: rank 0 alone steps the single-GPU
    # workload (all worlds when strong, one GPU's share when weak) while the others wait
    n1 = None
    if world_size > 1:
        if rank == 0:
            n1 = quick_worlds(args, frames=10, nw=total if strong else nw)
        barrier(world_size)
    for _ in range(args.warmup):
        b.step(DT, SUBSTEPS)
        b.sync()
    barrier(world_size)
    torch.cuda.synchronize()
    t_dev = 0.0
    with ClockSampler(torch.cuda.current_device()) as clk:
        for _ in range(args.steps):
            b.step(DT, SUBSTEPS)
            b.sync()
            t_dev += b.last_step_ms()
        torch.cuda.synchronize()
    barrier(world_size)
    st = b.stats()
    n = nw * BODIES_PER_WORLD
    pairs, contacts = float(st["n_pairs"].sum()), float(st["n_contacts"].sum())
    # statistics gathered over ranks (the only collective in the path)
    tot = allreduce_sum([pairs, contacts, float(n)], world_size)
    t_dev_max = allreduce_max(t_dev, world_size)
    units = tot[2] * SUBSTEPS * args.steps
    value = units / (t_dev_max * 1e-3)
    peak, peak_src = peaks()
    bytes_step = b_alg(tot[2], tot[0], tot[1], 0.0) * tot[2] * SUBSTEPS / world_size
    ach = bytes_step / (t_dev_max / args.steps * 1e-3) / 1e9
    # end to end through the public call with host buffers: every frame uploads all worlds'
    # states from a page-locked array, steps, and reads them back into the other array
    buf = [api.alloc_host_states(n), api.alloc_host_states(n)]
    buf[0][:] = b.download().reshape(-1)
    e2e_steps = max(3, min(args.steps, 10))
    b.step_host(buf[0], DT, SUBSTEPS, out=buf[1])
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for k in range(e2e_steps):
        b.step_host(buf[(k + 1) & 1], DT, SUBSTEPS, out=buf[k & 1])
    e2e_s = time.perf_counter() - t0
    e2e_val = allreduce_sum([n * SUBSTEPS * e2e_steps / e2e_s], world_size)[0]
    # the same workload through the library's own multi-GPU entry, from ONE process (what a
    # single-process engine would call): rank 0 drives all N GPUs, the other ranks wait on the CPU
    single = None
    if world_size > 1:
        del b, buf
        cpu_barrier()
        if rank == 0:
            try:
                single = single_process_multi(args, world_size, total)
            except Exception as e:  # a box that hides the other GPUs from rank 0
                single = {"error": str(e)}
        cpu_barrier()
    return {
        "metric": "body-substeps/sec", "value": value, "unit": "body-substeps/s",
        "n_gpus": world_size, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": t_dev_max / args.steps, "higher_is_better": True,
        "scaling": "strong" if strong else "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "n1_value": (n1["value"] if n1 else (value if world_size == 1 else None)),
        "n1_ms_per_step": (n1["ms_per_step"] if n1 else None),
        "per_gpu_value": value / world_size,
        "efficiency": (value / (world_size * n1["value"]) if n1 else (1.0 if world_size == 1 else None)),
        "efficiency_note": "value / (n_gpus * n1_value); n1_value = rank 0 alone stepping the single-GPU "
                           "workload of this curve in this run (all 8192 worlds when strong)",
        "config": {"workload": f"worlds512 ({'strong' if strong else 'weak'}): {total} independent "
                               f"{BODIES_PER_WORLD}-body worlds over {world_size} GPU(s), {nw} on rank 0, "
                               f"partitioned by world index",
                   "bodies": int(tot[2]), "dt": "1/60", "substeps": SUBSTEPS,
                   "settle_frames": args.settle, "pairs_per_body": tot[0] / tot[2],
                   "contacts_per_body": tot[1] / tot[2],
                   "l2": f"per-GPU state {n * 88 / 1e6:.0f} MB + contacts exceeds L2; no flush",
                   "collective": "statistics only (all_reduce of 3 floats after the timed region)"},
        "roofline": {"bound": "hbm", "kernel": "k_batch_frame (one launch per frame; a CTA runs a world's frame or, for the last worlds, single stages from a queue)",
                     "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                     "peak_source": peak_src, "traffic": ncu_traffic("worlds512", nw),
                     "peak_nominal": 8000.0, "frac_nominal": ach / 8000.0,
                     "algorithmic_bytes_per_launch": bytes_step,
                     "note": "traffic < algorithmic bytes: the per-substep re-reads the algorithmic "
                             "count charges (bodies, pair lists) are served from shared memory / "
                             "L2; what reaches DRAM is mostly the contact records written each substep"},
        "e2e": {"value": e2e_val, "unit": "body-substeps/s", "h2d_bytes_per_step": 32 * n * world_size,
                "d2h_bytes_per_step": 32 * n * world_size, "steps": e2e_steps,
                "call": "sfp_batch_step_host with host SfpBodyState arrays from sfp_alloc_host_states "
                        "(pinned), one process per GPU; bytes are whole-job"},
        "gpu_launches": args.steps, "clocks": clk.summary(),
        "single_process_multi_gpu": single,
    }


def single_process_multi(args, n_dev, total, frames=10):
    """sfp_batch_create_multi over devices 0..n_dev-1 from this process: wall time per frame of
    step + sync (the fan-out to the per-device worker threads included)"""
    import torch

    import starframe_b200 as sf
    from starframe_b200 import scenes

    if torch.cuda.device_count() < n_dev:
        raise RuntimeError(f"rank 0 sees {torch.cuda.device_count()} of {n_dev} GPUs")
    m = sf.BatchMulti(sf.product_api(), total, BODIES_PER_WORLD, list(range(n_dev)))
    m.upload(scenes.worlds(total, BODIES_PER_WORLD))
    for _ in range(args.settle):
        m.step(DT, SUBSTEPS)
    m.sync()
    t0 = time.perf_counter()
    dev_ms = 0.0
    for _ in range(frames):
        m.step(DT, SUBSTEPS)
        m.sync()
        dev_ms += m.last_step_ms()
    wall = time.perf_counter() - t0
    n = total * BODIES_PER_WORLD
    return {"value": n * SUBSTEPS * frames / wall, "unit": "body-substeps/s",
            "ms_per_step_wall": wall / frames * 1e3, "ms_per_step_slowest_device": dev_ms / frames,
            "call": f"sfp_batch_create_multi({total} worlds, devices 0..{n_dev - 1}) + sfp_batch_multi_step "
                    f"/ _sync from rank 0's process; partition {m.partition()}"}


# ------------------------------------------------------------------------------------------
# reference arm: CPU oracle on the host cores
# ------------------------------------------------------------------------------------------
def run_reference(args, rank, world_size):
    """CPU arm.  Each measurement is a BOUNDED SAMPLE of the arm's workload, settled on the CPU for
    the same number of frames as the GPU arm so that the timed frames are contact-dominated too."""
    if rank != 0:
        return None
    import concurrent.futures as cf

    import starframe_b200 as sf
    from oracle import loader
    from starframe_b200 import _ffi, scenes

    api = loader.api()
    cores = os.cpu_count() or 1
    if args.workload != "worlds512":
        # All host threads: the oracle's colour-order step runs each colour's constraints (and the
        # per-body / per-pair loops) on SFP_REF_THREADS threads, with the same bits as one thread
        # (scripts/oracle_threads.py); its broadphase and colouring stay sequential.  The FULL
        # scene of the GPU arm, settled on the CPU (a couple of minutes for the 100 k pile).
        os.environ["SFP_REF_THREADS"] = str(cores)
        bodies, joints = make_scene(args.workload)
        what = f"the full {args.workload} scene ({len(bodies)} bodies, the GPU arm's config)"
        n = len(bodies)
        ref = sf.World(api, solver_order=_ffi.ORDER_COLOUR)
        ref.upload_bodies(bodies)
        ref.set_joints(joints)
        for _ in range(args.settle + args.warmup):
            ref.step(DT, SUBSTEPS)
        t1 = time.perf_counter()
        done = 0
        for _ in range(args.steps):
            ref.step(DT, SUBSTEPS)
            done += 1
            if time.perf_counter() - t1 > 60.0:
                break
        s = time.perf_counter() - t1
        st = ref.stats()
        value = n * SUBSTEPS * done / s
        used = cores
        sample = (f"{what}, settled {args.settle} frames on the CPU, then {done} timed frames; "
                  f"{cores} threads in the solver loops (pairs/body {st['n_pairs'] / n:.2f}, "
                  f"contacts/body {st['n_contacts'] / n:.2f}); host has {cores} logical cores")
        workload, steps = args.workload, done
    else:
        total = TOTAL_WORLDS if args.worlds_per_gpu is None else args.worlds_per_gpu * world_size
        nw = min(total, 4 * cores)
        worlds = scenes.worlds(nw, BODIES_PER_WORLD)
        ws = []
        for k in range(nw):
            r = sf.World(api, solver_order=_ffi.ORDER_REFERENCE)
            r.upload_bodies(worlds[k])
            ws.append(r)

        def run(r, frames):
            for _ in range(frames):
                r.step(DT, SUBSTEPS)

        frames = max(1, min(args.steps, 20))
        with cf.ThreadPoolExecutor(cores) as ex:
            list(ex.map(lambda r: run(r, args.settle + args.warmup), ws))
            t1 = time.perf_counter()
            list(ex.map(lambda r: run(r, frames), ws))
            s = time.perf_counter() - t1
        n = nw * BODIES_PER_WORLD
        value = n * SUBSTEPS * frames / s
        used = cores
        sample = (f"{nw} of the {total} worlds (every world is the same amount of work), settled {args.settle} "
                  f"frames on the CPU, then {frames} timed frames; one world per task on {cores} "
                  f"threads")
        workload, steps = "worlds512", frames
    return {
        "impl": "reference", "metric": "body-substeps/sec", "value": value,
        "unit": "body-substeps/s", "n_gpus": world_size, "steps": steps, "warmup": args.warmup,
        "ms_per_step": s / steps * 1e3, "higher_is_better": True,
        "scaling": "strong" if (workload == "worlds512" and args.worlds_per_gpu is None) else "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload, "bodies": n, "dt": "1/60", "substeps": SUBSTEPS,
                   "settle_frames": args.settle},
        "cpu_baseline": {"value": value, "unit": "body-substeps/s", "cores": used, "kind": "port",
                         "sample": sample},
        "e2e": {"value": value, "unit": "body-substeps/s", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
        "note": "reference CPU step unavailable (tree not mounted, BASELINE.md §0): this is the "
                "CPU oracle port (single scene: colour-order Gauss-Seidel, colours in parallel; "
                "worlds: one sequential world per task)",
    }


# ------------------------------------------------------------------------------------------
def barrier(world_size):
    if world_size > 1:
        import torch.distributed as dist

        dist.barrier()


_CPU_GROUP = None


def cpu_barrier():
    """a barrier that leaves the GPUs alone (an NCCL barrier parks a spinning kernel on every
    waiting rank's GPU): a gloo group next to the NCCL one"""
    global _CPU_GROUP
    import torch.distributed as dist

    if not dist.is_initialized() or dist.get_world_size() == 1:
        return
    if dist.get_backend() == "gloo":
        dist.barrier()
        return
    if _CPU_GROUP is None:
        _CPU_GROUP = dist.new_group(backend="gloo")
    dist.barrier(group=_CPU_GROUP)


def _reduce_device():
    import torch.distributed as dist

    return "cuda" if dist.get_backend() == "nccl" else "cpu"


def allreduce_max(x, world_size):
    if world_size == 1:
        return x
    import torch
    import torch.distributed as dist

    t = torch.tensor([x], dtype=torch.float64, device=_reduce_device())
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def allreduce_sum(xs, world_size):
    if world_size == 1:
        return list(xs)
    import torch
    import torch.distributed as dist

    t = torch.tensor(list(xs), dtype=torch.float64, device=_reduce_device())
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return [float(v) for v in t.tolist()]


def world_range(rank, world_size, total):
    """world indices owned by `rank`: contiguous block partition of [0, total) (no overlap, no
    gap; the first total % world_size ranks own one world more)"""
    base, extra = divmod(total, world_size)
    start = rank * base + min(rank, extra)
    return range(start, start + base + (1 if rank < extra else 0))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=120)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=None)
    ap.add_argument("--settle", type=int, default=240)
    ap.add_argument("--worlds-per-gpu", type=int, default=None,
                    help="weak scaling: this many worlds on every GPU (default: 8192 / N, strong)")
    ap.add_argument("--cpu-frames", type=int, default=12)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-other-configs", action="store_true")
    ap.add_argument("--no-scale-point", action="store_true")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world_size = int(os.environ.get("WORLD_SIZE", "1"))
    if world_size != args.gpus and world_size > 1:
        args.gpus = world_size
    if args.workload is None:
        args.workload = "pile100k" if world_size == 1 else "worlds512"
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    if args.impl == "reference":
        out = run_reference(args, rank, world_size)
        if out is not None:
            print(json.dumps(out), flush=True)
        return

    import torch

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world_size > 1:
        import torch.distributed as dist

        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    if args.workload == "worlds512":
        out = run_worlds(args, rank, world_size)
    else:
        out = run_single(args, rank, world_size)
    if rank == 0:
        print(json.dumps(out), flush=True)
    if world_size > 1:
        import torch.distributed as dist

        dist.destroy_process_group()


if __name__ == "__main__":
    main()
