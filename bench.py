#!/usr/bin/env python
"""Benchmark of the Dr.TVAM hot path on B200: rays/s and seconds per L-BFGS
iteration on the synthetic Benchy-sized workload (BASELINE.json config 2:
index-matched vial, 400x400 DMD, 400 angles, 400^3 dose grid, 1 spp).

    python bench.py --gpus N --steps K --warmup W [--impl reference] [--workload C2]

One "step" is the projection work of one L-BFGS iteration (optimize.py:294,308
and lbfgs.py:248): forward A p, thresholded loss + its gradient, adjoint A^T g,
forward A d (signed search direction) and one Armijo loss evaluation.  Prints
ONE JSON line (see the contract in the task description).  Beside the headline
(the configuration BASELINE.json quotes the metric on) the line carries
  default_sampling  the identical step with regular_sampling=false, the reference's default
                    (integrators/common.py:22);
  lbfgs_iteration   a whole LinearLBFGS iteration (history, two-loop recursion, Armijo search,
                    pattern update, clamp) timed around the same projections;
  multi_gpu_check   (N > 1) the N-rank dose and pattern gradient against the 1-rank ones.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="C2")
    ap.add_argument("--n", type=int, default=0, help="override DMD/grid edge (default: full size)")
    ap.add_argument("--jitter", action="store_true", help="headline on regular_sampling=false (per-ray jitter)")
    ap.add_argument("--general", action="store_true", help="force the general per-lane tracer")
    ap.add_argument("--surface-aware", action="store_true",
                    help="surface-aware discretization (two-channel film + target-mesh crossings; BASELINE config 4)")
    ap.add_argument("--emulate-shard", type=int, default=0, metavar="W",
                    help="profiling aid: run rank 0's shard of a W-rank job on one GPU (no collective)")
    ap.add_argument("--partition", default="auto", choices=["auto", "angle", "zslab"],
                    help="multi-GPU partitioning: by projection angle (one all-reduce of the dose per forward) or, "
                         "for planar scenes, by z-slab / DMD row (no dose collective)")
    ap.add_argument("--grid", default="", help="A,S: angle groups x z-slabs of the zslab partition (default: parallel.slab_grid)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-default-sampling", action="store_true", help="skip the regular_sampling=false record")
    ap.add_argument("--no-lbfgs", action="store_true", help="skip the whole-L-BFGS-iteration record")
    ap.add_argument("--no-verify", action="store_true", help="skip the N-rank == 1-rank check")
    ap.add_argument("--cpu-seconds", type=float, default=15.0)
    ap.add_argument("--profile-region", default="step", choices=["step", "lbfgs"],
                    help="which loop cudaProfilerStart/Stop brackets (ncu --profile-from-start off)")
    return ap.parse_args()


WORKLOAD_TEXT = {
    "C1": "C1: index-matched vial r=5, collimated 128 angles x 128x128 DMD, 128^3 dda grid, 1 spp",
    "C2": "C2: Benchy-sized synthetic target, index-matched vial r=7.5, collimated 400 angles x 400x400 DMD "
          "(25 um), 400^3 dda grid, 1 spp",
    "C3": "C3: as C2 with a cylindrical glass vial (refraction)",
    "C4": "C4: as C2 with a square glass vial (refraction)",
    "C5": "C5: as C3 with a Henyey-Greenstein scattering resin, 16 spp",
}

# shared-memory / L2 atomic peaks measured with csrc/microbench.cu on this pool's B200
# (profiles/r01_microbench_atomics.json); visits/s are set against them
ATOMIC_PEAKS = {
    "smem_lds_fadd_sts": (2.34482e12, "shared-memory ld + add + st per lane, conflict free"),
    "smem_atoms_f32_cas": (1.51375e12, "shared-memory fp32 CAS add"),
    "red_v4_l2_resident": (1.57674e12, "red.global.add.v4.f32, 64 MiB (L2 resident), per float"),
    "red_f32_scattered_l2": (1.9876e11, "red.global.add.f32, one 32-byte sector per lane, 64 MiB (L2 resident)"),
}


class ClockSampler:
    """SM clock and throttle reasons DURING the timed region: NVML polled every few milliseconds from a
    thread (nvidia-smi's own loop yields 2-3 samples in a 0.3 s region), nvidia-smi as the fallback."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.samples, self.reason_bits, self.max_mhz = [], 0, None
        self.stop_flag = threading.Event()
        self.proc = None
        self.rows = []
        self.mode = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nvml = pynvml
            # CUDA_VISIBLE_DEVICES may renumber: resolve through the PCI bus id of the torch device
            try:
                import torch
                bus = torch.cuda.get_device_properties(index).pci_bus_id
                dom = torch.cuda.get_device_properties(index).pci_domain_id
                dev_id = torch.cuda.get_device_properties(index).pci_device_id
                self.handle = pynvml.nvmlDeviceGetHandleByPciBusId(f"{dom:08X}:{bus:02X}:{dev_id:02X}.0")
            except Exception:
                self.handle = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
            self.mode = "nvml"
            self.thread = threading.Thread(target=self._poll, daemon=True)
            self.thread.start()
            return
        except Exception:
            self.mode = None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "50", "-i", str(index)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.mode = "nvidia-smi"
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _poll(self):
        n = self.nvml
        while not self.stop_flag.is_set():
            try:
                self.samples.append(float(n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM)))
                self.reason_bits |= int(n.nvmlDeviceGetCurrentClocksEventReasons(self.handle))
            except Exception:
                try:
                    self.reason_bits |= int(n.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle))
                except Exception:
                    pass
            time.sleep(0.004)

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if self.mode == "nvml":
            self.stop_flag.set()
            self.thread.join(timeout=2)
            sm = sorted(self.samples)
            bits = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4}
            reasons = sorted(k for k, b in bits.items() if self.reason_bits & b)
            return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": self.max_mhz, "reasons": reasons,
                    "samples": len(sm), "source": "nvml, 4 ms period, inside the timed region"}
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            parts = [p.strip() for p in r.split(",")]
            if len(parts) < 6:
                continue
            try:
                sm.append(float(parts[0]))
                mx = float(parts[1])
            except ValueError:
                continue
            for n, v in zip(names, parts[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm), "source": "nvidia-smi -lms 50"}


# mma.sync.aligned.m16n8k8 tf32 (SASS HMMA), 8 warps per SM x 16 accumulator tiles: dense TFLOP/s measured by
# tools/microbench.py on this pool's B200 (profiles/r02g_microbench.json)
MMA_SYNC_TF32_PEAK = 276.0
# mma.sync.aligned.m16n8k16 f16 -> f32 (SASS HMMA.16816.F32): the same issue rate, twice the flops per instruction
# (profiles/r03d_microbench.json)
MMA_SYNC_F16_K16_PEAK = 555.0


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return json.load(f).get("hbm_gbs", 6650.0), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def measured_traffic(workload, plan, world, edge):
    """DRAM bytes per forward pass from the ncu --set full capture committed under profiles/
    (profiles/ncu_traffic.json, written by tools/ncu_summary.py traffic); None when no capture of this
    workload / kernel family exists."""
    path = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if world != 1 or edge or not os.path.exists(path):
        return None, None
    try:
        with open(path) as f:
            rec = json.load(f).get(f"{workload}:plan{plan}")
    except (OSError, ValueError):
        return None, None
    if not rec:
        return None, None
    return rec["forward_dram_bytes"], rec["source"]


def oracle_sample(cfg, desc, seconds, threads=0):
    """Times the CPU oracle (fp32 build, OpenMP over rays) on a bounded sample of
    the workload: whole angles of the same scene, 3 passes each (A p, A^T g, A d)."""
    import numpy as np
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle as orc
    orc.build()
    o = orc.Oracle(desc.as_dict(), "f32", threads=threads)
    per_angle = desc.res_x * desc.res_y
    rng = np.random.default_rng(0)
    g = rng.standard_normal(desc.film_shape()).astype(np.float32)
    n_angles, rays, t_used = 1, 0, 0.0
    k0 = 0
    while True:
        ks = np.arange(k0, min(k0 + n_angles, desc.n_patterns))
        if ks.size == 0:
            break
        pix = (ks[:, None] * per_angle + np.arange(per_angle)[None, :]).reshape(-1).astype(np.uint32)
        p = rng.random(pix.size).astype(np.float32) + 0.01
        d = (rng.random(pix.size).astype(np.float32) - 0.5)
        t0 = time.perf_counter()
        o.forward(p, pix)
        o.adjoint(g, pix)
        o.forward(d, pix)
        t_used += time.perf_counter() - t0
        rays += 3 * pix.size
        k0 += ks.size
        if t_used > seconds:
            break
        n_angles = min(n_angles * 2, 16)
    return rays / t_used, rays, t_used, k0


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference" and rank != 0:
        return 0

    import numpy as np
    import torch
    from drtvam_b200.synthetic import make_config

    workdir = os.path.join(tempfile.gettempdir(), f"drtvam_bench_{os.getuid()}_{rank}")

    def config_for(regular):
        cfg = make_config(args.workload, workdir, n=args.n or None, regular_sampling=regular,
                          surface_aware=args.surface_aware)
        if args.workload == "C5":
            cfg["regular_sampling"] = False
        return cfg

    cfg = config_for(not args.jitter)
    spp = int(cfg["spp"])
    host_cores = os.cpu_count() or 1

    def workload_text(c):
        w = WORKLOAD_TEXT[args.workload] + (f" [edge overridden to {args.n}]" if args.n else "")
        w += ", regular_sampling=%s" % str(bool(c["regular_sampling"])).lower()
        if args.surface_aware:
            w += ", surface-aware discretization (2-channel film, target mesh traced)"
        return w

    workload = workload_text(cfg)

    # ------------------------------------------------------------ reference arm (CPU)
    if args.impl == "reference":
        from drtvam_b200.integrators import VolumeIntegrator
        from drtvam_b200.optimize import load_scene
        from drtvam_b200.scene import load_dict
        scene = load_dict(load_scene(cfg), device="cpu")
        integ = VolumeIntegrator({"max_depth": cfg["max_depth"], "rr_depth": cfg["rr_depth"],
                                  "regular_sampling": cfg["regular_sampling"], "print_time": 1.0})
        desc = scene.describe(integ, scene.sensor("sensor"))
        vals = []
        per_step = max(2.0, min(20.0, 120.0 / max(1, args.steps + args.warmup)))
        sample = ""
        for i in range(args.warmup + args.steps):
            rps, rays, t, n_ang = oracle_sample(cfg, desc, per_step, threads=host_cores)
            if i >= args.warmup:
                vals.append((rps, rays, t))
            sample = f"{n_ang} of {desc.n_patterns} angles x {desc.res_x}x{desc.res_y} rays, 3 passes (A p, A^T g, A d)"
        total_rays = sum(v[1] for v in vals)
        total_t = sum(v[2] for v in vals)
        value = total_rays / total_t
        full_rays = 3 * desc.n_patterns * desc.res_x * desc.res_y * spp
        line = {
            "impl": "reference", "metric": "rays_per_s", "value": value, "unit": "rays/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": full_rays / value * 1e3,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": workload, "rays_per_step": full_rays, "gpus_used": 0,
                       "note": "the reference's Dr.Jit LLVM path is unavailable on this box (mitsuba/drjit not "
                               "installable); this is the CPU oracle PORT, fp32, OpenMP over rays with one "
                               "`omp atomic` per deposit (an untuned baseline); every step times a bounded sample "
                               "of whole angles and ms_per_step is EXTRAPOLATED from it to the full step"},
            "cpu_baseline": {"value": value, "unit": "rays/s", "cores": host_cores, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": "rays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0,
        }
        print(json.dumps(line), flush=True)
        return 0

    # ------------------------------------------------------------ B200 arm
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the B200 arm has no CPU fallback (use --impl reference)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    import torch.distributed as dist
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    import drtvam_b200 as dv
    from drtvam_b200 import _lib
    from drtvam_b200.integrators import VolumeIntegrator
    from drtvam_b200.lbfgs import LinearLBFGS
    from drtvam_b200.loss import LazyAxpy, ThresholdedLoss
    from drtvam_b200.optimize import load_scene
    from drtvam_b200 import parallel
    from drtvam_b200.utils import discretize

    lib = _lib.load()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def rel_l2(a, b):
        return float(torch.linalg.norm((a.double() - b.double())) / torch.linalg.norm(b.double()).clamp_min(1e-300))

    def setup(c):
        """Scene, integrator, sensor, loss target and (N > 1) the shard of this rank."""
        scene = dv.load_dict(load_scene(c), device=dev)
        scene.force_general = bool(args.general)
        integ = VolumeIntegrator({"max_depth": c["max_depth"], "rr_depth": c["rr_depth"],
                                  "regular_sampling": c["regular_sampling"], "print_time": 1.0})
        sensor = scene.sensor("sensor")
        desc_full = scene.describe(integ, sensor)
        n_total = scene.projector.active_size()
        if sensor.film().surface_aware:
            target = sensor.compute_volume(scene)     # sensor.py:47-110, once per run (optimize.py:131-134)
        else:
            target = discretize(scene, sensor=sensor)
        return scene, integ, sensor, desc_full, n_total, target

    def verify_multi_gpu(scene, integ, sensor, n_total, c_spp, target):
        """N-rank result == 1-rank result, on hardware, before the timed loop: every rank draws the SAME full
        pattern vector; rank 0 projects it unsharded (no collective), forms the dose gradient of the
        thresholded loss there and back-projects it; the gradient is broadcast, then all ranks project
        their shards; the reduced dose and the concatenated pattern gradient must match."""
        gen = torch.Generator(device=dev).manual_seed(4242)
        p_full = torch.rand(n_total, device=dev, generator=gen) * 2e-3 + 1e-4
        film_shape = tuple(sensor.film().shape)
        g_full = torch.zeros(film_shape, device=dev)
        ref = {}
        if rank == 0:
            scene.projector.active_data = p_full
            with torch.no_grad():
                dose = integ._project(scene, sensor, 77, c_spp, p_full).clone()
            x = dose.clone().requires_grad_(True)
            ThresholdedLoss({})(x, target, p_full).backward()
            g_full = x.grad.detach().clone()
            with torch.no_grad():
                ref["dose"] = dose
                ref["grad"] = integ.render_backward(scene, None, g_full, sensor, 77, c_spp).clone()
        dist.broadcast(g_full, 0)
        return p_full, g_full, ref

    def measure(c, steps, warmup, headline):
        """Times `steps` projection steps of config `c`; returns a dict of numbers."""
        c_spp = int(c["spp"])
        scene, integ, sensor, desc_full, n_total, target = setup(c)
        check = None
        verify = world > 1 and headline and not args.no_verify
        if verify:
            p_full, g_full, ref = verify_multi_gpu(scene, integ, sensor, n_total, c_spp, target)
        partition = "single"
        grid = tuple(int(v) for v in args.grid.split(",")) if args.grid else None
        if world > 1:
            partition = scene.shard(rank, world, partition=args.partition, integrator=integ, sensor=sensor, grid=grid)[2]
        elif args.emulate_shard > 1:
            partition = scene.shard(0, args.emulate_shard, partition=args.partition, integrator=integ,
                                    sensor=sensor, grid=grid)[2]
            scene.world_size = 1
        n_local = scene.projector.active_size()
        pixels = scene.projector.active_pixels
        if verify:
            with torch.no_grad():
                idx = scene.shard_index                      # positions of this rank's entries in the full vector
                p_loc = p_full[idx].contiguous()
                scene.projector.active_data = p_loc
                dose_n = integ._project(scene, sensor, 77, c_spp, p_loc)
                dose_n = scene.gather_dose(dose_n)           # identity when every rank holds the full dose
                grad_loc = integ.render_backward(scene, None, scene.local_dose_grad(g_full), sensor, 77, c_spp)
                grad_n = torch.zeros(n_total, device=dev)
                grad_n[idx] = grad_loc
                dist.all_reduce(grad_n)
                if rank == 0:
                    check = {"partition": partition, "grid": list(scene.grid), "ranks": world,
                             "dose_rel_l2_vs_1_rank": rel_l2(dose_n, ref["dose"]),
                             "grad_rel_l2_vs_1_rank": rel_l2(grad_n, ref["grad"]),
                             # ranks that sum partial doses of different angles re-order the accumulation of the
                             # tensor-core products (their fp32 accumulators truncate): 2e-5 measured; slabs a
                             # rank owns alone must match to rounding
                             "bound": 1e-6 if scene.dose_group_size == 1 and partition == "zslab" else 5e-5,
                             "inputs": "same seeded full pattern vector on every rank; dose gradient = that of the "
                                       "thresholded loss at the 1-rank dose, broadcast from rank 0"}
                    check["ok"] = bool(check["dose_rel_l2_vs_1_rank"] <= check["bound"] and
                                       check["grad_rel_l2_vs_1_rank"] <= check["bound"])
            del p_full, g_full, ref, dose_n, grad_n
            torch.cuda.empty_cache()
        loss_fn = ThresholdedLoss({})
        if world > 1:
            loss_fn.dose_allsum = scene.dose_allsum()         # z-slab sharding: the dose term is a sharded sum
        target_loc = scene.local_target(target)

        gen = torch.Generator(device=dev).manual_seed(1234 + rank)
        # dense, strictly non-zero inputs so that no ray is skipped
        p_dev = (torch.rand(n_local, device=dev, generator=gen) * 2e-3 + 1e-4)
        d_dev = (torch.rand(n_local, device=dev, generator=gen) - 0.5) * 1e-3
        d_dev = torch.where(d_dev == 0, torch.full_like(d_dev, 1e-4), d_dev)
        pass_events = {}

        def step(seed, p, d, record=False, d_ready=None, on_grad=None):
            """One L-BFGS iteration's projections on device-resident inputs.  `d_ready` (an event) is
            waited for before A d reads the search direction; `on_grad` sees the pattern gradient as soon
            as the adjoint is queued (the end-to-end loop uses both to overlap its copies)."""
            p = p.detach().requires_grad_(True)
            scene.projector.active_data = p
            integ.events = pass_events if record else None
            prm = dv.traverse(scene)
            vol = dv.render(scene, prm, integrator=integ, sensor=sensor, spp=c_spp, seed=seed)
            loss = loss_fn(vol, target_loc, p)
            loss.backward()                                     # adjoint A^T g
            if on_grad is not None:
                on_grad(p.grad)
            with torch.no_grad():
                if d_ready is not None:
                    torch.cuda.current_stream(dev).wait_event(d_ready)
                scene.projector.active_data = d
                dvol = dv.render(scene, None, integrator=integ, sensor=sensor, spp=c_spp, seed=seed)   # A d
                f_new = loss_fn(LazyAxpy(vol.detach(), dvol, 1.0), target_loc, d)                      # Armijo trial
            return loss.detach(), f_new, p.grad

        for i in range(warmup):
            step(i, p_dev, d_dev)
        barrier()

        clocks = ClockSampler(local_rank) if (rank == 0 and headline) else None
        launches0 = lib.drtvam_launch_count()
        t0 = torch.cuda.Event(enable_timing=True)
        t1 = torch.cuda.Event(enable_timing=True)
        barrier()
        if headline and args.profile_region == "step":
            torch.cuda.cudart().cudaProfilerStart()   # ncu --profile-from-start off: only the timed steps are listed
        t0.record()
        for i in range(steps):
            out = step(100 + i, p_dev, d_dev, record=True)
        t1.record()
        barrier()
        if headline and args.profile_region == "step":
            torch.cuda.cudart().cudaProfilerStop()
        ms = t0.elapsed_time(t1)
        launches = lib.drtvam_launch_count() - launches0
        clock_info = clocks.stop() if clocks else None
        res = {"ms": ms, "launches": launches, "clocks": clock_info, "loss": float(out[0]), "check": check,
               "partition": partition, "grid": tuple(scene.grid), "n_local": n_local, "n_total": n_total, "spp": c_spp}

        def mean_ms(name):
            ev = pass_events.get(name, [])
            return sum(a.elapsed_time(b) for a, b in ev) / max(1, len(ev))
        res["fwd_ms"], res["adj_ms"], res["ar_ms"] = mean_ms("forward"), mean_ms("adjoint"), mean_ms("allreduce")
        h = scene.handle(integ, sensor)
        res["plan"] = int(lib.drtvam_scene_plan(h, n_local, c_spp))
        stat = _lib.C.c_uint64(0)
        res["stat"] = {}
        for name, which in (("pairs", 0), ("layers", 1), ("records", 2), ("coef_bytes", 3)):
            _lib.check(lib.drtvam_scene_stat(h, which, _lib.C.byref(stat)))
            res["stat"][name] = int(stat.value)

        if headline:
            # end to end through the integrator's own entry points (VolumeIntegrator.render / render_backward,
            # what mi.render and dr.backward call): host (pinned) inputs copied in and results copied out
            # every step.  The patterns are uploaded in blocks of angles under the projection of the previous
            # block (render(host_patterns=...)); the search direction is uploaded and the pattern gradient
            # downloaded on a copy stream, under A^T g and under A d respectively (render_backward(host_out=...) does
            # the same download inside the integrator; through it the step measured 1.7 ms slower, r03x).  Every copy of step i is
            # issued and completed inside step i (no prefetch across steps).
            from drtvam_b200.scene import sample_tea_32
            p_host = p_dev.cpu().pin_memory()
            d_host = d_dev.cpu().pin_memory()
            g_host = torch.empty(n_local, dtype=torch.float32).pin_memory()
            integ.events = None
            barrier()
            main_stream = torch.cuda.current_stream(dev)
            copy_stream = torch.cuda.Stream(dev)
            e0 = torch.cuda.Event(enable_timing=True)
            e1 = torch.cuda.Event(enable_timing=True)
            e0.record()
            loss_host = 0.0
            for i in range(steps):
                seed = 100 + i
                vol = integ.render(scene, sensor, seed=seed, spp=c_spp, host_patterns=p_host)      # upload + A p
                p_in = scene.projector.active_data
                copy_stream.wait_stream(scene.copy_stream())         # the direction follows the patterns on the bus
                with torch.cuda.stream(copy_stream):
                    d_in = d_host.to(dev, non_blocking=True)
                    d_ready = torch.cuda.Event()
                    d_ready.record(copy_stream)
                d_in.record_stream(main_stream)
                x = vol.detach().requires_grad_(True)
                loss = loss_fn(x, target_loc, p_in)
                loss.backward()
                grad = integ.render_backward(scene, None, x.grad, sensor, sample_tea_32(seed, 1)[0], c_spp)   # A^T g
                copy_stream.wait_stream(main_stream)                 # after the adjoint
                grad.record_stream(copy_stream)
                with torch.cuda.stream(copy_stream):
                    g_host.copy_(grad, non_blocking=True)
                with torch.no_grad():
                    main_stream.wait_event(d_ready)
                    scene.projector.active_data = d_in
                    dvol = dv.render(scene, None, integrator=integ, sensor=sensor, spp=c_spp, seed=seed)      # A d
                    f_new = loss_fn(LazyAxpy(vol.detach(), dvol, 1.0), target_loc, d_in)                        # Armijo trial
                main_stream.wait_stream(copy_stream)                 # the gradient is on the host before the step ends
                loss_host = float(loss)                              # D2H of the scalar: syncs the step
            e1.record()
            barrier()
            res["ms_e2e"] = e0.elapsed_time(e1)
            res["loss"] = loss_host
            n_vis = _lib.C.c_uint64(0)
            _lib.check(lib.drtvam_count_visits(h, pixels.data_ptr(), n_local, c_spp, 100, _lib.C.byref(n_vis)))
            res["visits"] = int(n_vis.value)
            res["film_elems"] = int(np.prod(desc_full.film_shape()))
            res["desc"] = desc_full

            if not args.no_lbfgs:
                # a whole LinearLBFGS iteration as optimize() runs it (optimize.py:286-320): A p, loss,
                # backward, history + two-loop recursion, A d, Armijo search, pattern update, clamp
                state = {"i": 0}

                def render_fn(vars_):
                    scene.projector.active_data = vars_["projector.active_data"].detach()
                    return dv.render(scene, None, integrator=integ, sensor=sensor, spp=c_spp, seed=state["i"])

                def loss2(y, patterns):
                    return loss_fn(y, target_loc, patterns)
                loss2.accepts_lazy = True
                loss2.armijo_sweep = lambda x, dx, direction, *a: loss_fn.armijo_sweep(x, dx, target_loc, direction, *a)
                opt = LinearLBFGS(loss_fn=loss2, render_fn=render_fn, dot=parallel.make_dot(None, world) if world > 1 else None,
                                  allsum=parallel.make_sum(None, world) if world > 1 else None)
                opt["projector.active_data"] = p_dev.clone()
                it_ev = []
                n_it = warmup + steps
                for i in range(n_it):
                    state["i"] = i
                    if i == n_it - 2 and args.profile_region == "lbfgs":
                        torch.cuda.cudart().cudaProfilerStart()
                    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    a.record()
                    p = opt["projector.active_data"]
                    scene.projector.active_data = p
                    prm = dv.traverse(scene)
                    vol = dv.render(scene, prm, integrator=integ, sensor=sensor, spp=c_spp, seed=i)
                    loss = loss_fn(vol, target_loc, p)
                    loss.backward()
                    opt.step(vol.detach(), loss.detach())
                    opt["projector.active_data"] = torch.clamp_min(opt["projector.active_data"].detach(), 0)
                    b.record()
                    it_ev.append((a, b, opt._alpha))      # a device tensor: read after the loop, no sync inside it
                barrier()
                if args.profile_region == "lbfgs":
                    torch.cuda.cudart().cudaProfilerStop()
                its = sorted(a.elapsed_time(b) for a, b, _ in it_ev[warmup:])
                res["lbfgs_ms_median"] = its[len(its) // 2]
                trials = [float(t[1]) if isinstance(t, torch.Tensor) else 1.0 for _, _, t in it_ev[warmup:]]
                res["lbfgs_trials_mean"] = sum(trials) / max(1, len(trials))
        scene.close()
        del scene, target, target_loc, p_dev, d_dev
        torch.cuda.empty_cache()
        return res

    R = measure(cfg, args.steps, args.warmup, headline=True)

    D = None
    if not args.no_default_sampling and cfg["regular_sampling"] and args.workload != "C5":
        D = measure(config_for(False), max(2, min(args.steps, 5)), 3, headline=False)

    times = torch.tensor([R["ms"], R["ms_e2e"], D["ms"] if D else 0.0, R.get("lbfgs_ms_median", 0.0)], device=dev,
                         dtype=torch.float64)
    if world > 1:
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
    ms, ms_e2e, ms_default, ms_lbfgs = times.tolist()

    n_active_total, n_local = R["n_total"], R["n_local"]
    rays_per_step = 3 * n_active_total * spp
    value = rays_per_step * args.steps / (ms * 1e-3)
    e2e_value = rays_per_step * args.steps / (ms_e2e * 1e-3)

    # ---- roofline of the dominant kernel family: the forward projection A p
    fwd_ms, adj_ms, ar_ms, visits, plan_now = R["fwd_ms"], R["adj_ms"], R["ar_ms"], R["visits"], R["plan"]
    peak, how = measured_peaks()
    # algorithmic HBM bytes of one forward pass of this rank: pattern values + pixel indices read once, the
    # (local part of the) dose volume written once
    film_local = R["film_elems"] if R["partition"] != "zslab" else R["film_elems"] // max(R["grid"][1], 1)
    hbm_alg_bytes = 4.0 * n_local * 2 + 4.0 * film_local
    achieved = hbm_alg_bytes / (fwd_ms * 1e-3) * 1e-9
    visits_per_s = visits / (fwd_ms * 1e-3)
    atomic_key = {2: "smem_lds_fadd_sts", 3: "smem_atoms_f32_cas", 1: "smem_atoms_f32_cas"}.get(plan_now,
                                                                                                "red_f32_scattered_l2")
    traffic, traffic_src = measured_traffic(args.workload, plan_now, world, args.n)
    pairs, layers = R["stat"]["pairs"], R["stat"]["layers"]
    on_tensor = plan_now == 2 and pairs > 0 and os.environ.get("DRTVAM_PLANAR_FORWARD", "") != "ffma"
    roofline = {
        "kernel": {2: "forward projection A p: zpack + planarg_mma_kernel (patch gather as mma.sync m16n8k8 tf32, "
                      "3xTF32 split)" if on_tensor else "forward projection A p: zpack + planarg_kernel (patch gather, FFMA)",
                   3: "forward projection A p: jtable_kernel + planarj_kernel (jittered rays, shared-memory tiles)",
                   0: "forward projection A p: general per-lane tracer"}.get(plan_now, "forward projection A p"),
        "traffic": traffic, "traffic_source": traffic_src,
        "frac_hbm": achieved / peak, "hbm_algorithmic_GBps": achieved, "hbm_peak_GBps": peak, "hbm_peak_source": how,
        "hbm_algorithmic_bytes": hbm_alg_bytes,
        "frac_atomic": visits_per_s / ATOMIC_PEAKS[atomic_key][0],
        "atomic_peak": {"name": atomic_key, "ops_per_s": ATOMIC_PEAKS[atomic_key][0], "what": ATOMIC_PEAKS[atomic_key][1],
                        "source": "profiles/r01_microbench_atomics.json (csrc/microbench.cu)"},
        "visits_per_launch": visits, "visits_per_s": visits_per_s, "kernel_ms": fwd_ms,
        "adjoint_ms": adj_ms, "adjoint_visits_per_s": visits / (adj_ms * 1e-3) if adj_ms else None,
        "allreduce_ms": ar_ms, "plan": plan_now,
        "evidence": "profiles/ (ncu --set full details and launch lists of this round, named r02*)",
    }
    if on_tensor:
        # the dominant kernel is a block-sparse x dense product on the tensor pipe (ncu: the tensor pipe is the
        # busiest unit; profiles/r02i_*, r03b_*).  Dense-equivalent flops of one pass = products x pairs x 16 voxels x
        # layers x 2: the split-fp16 kernel issues all four products of its {hi, lo} operands as two m16n8k16
        # instructions per tile, the 3xTF32 kernel (DRTVAM_PLANAR_FORWARD=tf32) three m16n8k8 products.
        h16 = os.environ.get("DRTVAM_PLANAR_FORWARD", "") == ""
        products, mma_peak = (4.0, MMA_SYNC_F16_K16_PEAK) if h16 else (3.0, MMA_SYNC_TF32_PEAK)
        flops = products * pairs * 16.0 * layers * 2.0
        tf = flops / (fwd_ms * 1e-3) * 1e-12
        roofline.update({
            "kernel": "forward projection A p: absmax + zpack_h16 + planarg_h16_kernel (patch gather as mma.sync m16n8k16 "
                      "f16 on {hi, lo} fp16 splits of both operands)" if h16 else roofline["kernel"],
            "bound": "tensor", "achieved": tf, "peak": mma_peak, "unit": "TFLOP/s",
            "frac": tf / mma_peak,
            "peak_source": "measured on this pool's B200 for the instruction the kernel issues (mma.sync m16n8k16 f16 -> "
                           "SASS HMMA.16816.F32, or m16n8k8 tf32; csrc/microbench.cu, profiles/r03d_microbench.json); "
                           "MEASURED_PEAKS.json only holds the cuBLAS bf16 tcgen05 figure, which this gathered 16x8xK "
                           "product cannot be fed into (DESIGN.md section 5)",
            "flops_per_launch": flops, "coefficient_pairs": pairs, "layers": layers,
            "useful_flops_per_launch": 2.0 * visits,
            "note": "achieved = dense-equivalent flops of one forward pass / CUDA-event time of the pass; 80 % "
                    "of them multiply zeros of the block-sparse coefficient blocks and the operand split multiplies "
                    "them by 4 (useful = 2 flops per voxel visit). "
                    "frac_hbm = algorithmic HBM bytes (pattern values + pixel indices read, dose written) / time "
                    "against the measured copy peak; frac_atomic = voxel visits/s against the measured rate of the "
                    "shared-memory read-add-write the scatter form would need (accumulators live in registers here, so "
                    "it can exceed 1)."})
    else:
        roofline.update({
            "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
            "peak_source": how,
            "note": "achieved = algorithmic HBM bytes of one forward pass (pattern values + pixel indices read, dose "
                    "written) / CUDA-event time of the pass, against the measured copy peak.  The pass is NOT near "
                    "that roof: its work is ~478 voxel visits per ray on data that stays on chip, and ncu shows it "
                    "bound by instruction issue and on-chip atomics (plan 3: shared-memory red.add, plan 0: L2 "
                    "red.global.add); frac_atomic = voxel visits/s against the measured peak rate of that atomic "
                    "(SURVEY 8(d)) is the fraction that describes it."})

    line = None
    if rank == 0:
        cpu_baseline = None
        if not args.no_cpu_baseline and world == 1:
            rps, rays, t, n_ang = oracle_sample(cfg, R["desc"], args.cpu_seconds, threads=host_cores)
            d_ = R["desc"]
            cpu_baseline = {"value": rps, "unit": "rays/s", "cores": host_cores, "kind": "port",
                            "sample": f"{n_ang} of {d_.n_patterns} angles x {d_.res_x}x{d_.res_y} "
                                      f"rays, 3 passes, {t:.1f} s of oracle (fp32, OpenMP, omp atomic per deposit) time; "
                                      "whole-step figures derived from it are extrapolations"}
        par = {"single": "1 GPU", "angle": f"angle-sharded x{world}, 1 all_reduce(dose) per forward",
               "zslab": f"{R['grid'][0]} angle groups x {R['grid'][1]} z-slabs (DMD rows): a rank projects its angles "
                        f"into its slab of layers; the dose of a slab is all-reduced among the {R['grid'][0]} ranks that "
                        "share it (no collective when that is one rank), scalar all-reduces for the loss and the "
                        "L-BFGS Gram matrix"}[R["partition"]]
        line = {
            "metric": "rays_per_s", "value": value, "unit": "rays/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "sec_per_projection_step": ms / args.steps * 1e-3,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": workload, "rays_per_step": rays_per_step,
                       "step": "A p + thresholded loss + A^T g + A d + 1 Armijo trial (one L-BFGS iteration's "
                               "projections), dense non-zero patterns",
                       "l2": "inputs (patterns 256 MB, dose 256 MB at C2) exceed the 126 MB L2",
                       "parallelism": par},
            "clocks": R["clocks"],
            "e2e": {"value": e2e_value, "unit": "rays/s", "ms_per_step": ms_e2e / args.steps,
                    "h2d_bytes_per_step": 8 * n_local, "d2h_bytes_per_step": 4 * n_local + 4},
            "gpu_launches": int(R["launches"]),
            "roofline": roofline,
            "cpu_baseline": cpu_baseline,
            "loss": R["loss"],
        }
        if D is not None:
            line["default_sampling"] = {
                "workload": workload_text(config_for(False)), "ms_per_step": ms_default / D_steps(args),
                "fwd_ms": D["fwd_ms"], "adj_ms": D["adj_ms"], "allreduce_ms": D["ar_ms"], "plan": D["plan"],
                "steps": D_steps(args), "rays_per_s": rays_per_step * D_steps(args) / (ms_default * 1e-3),
                "note": "the identical step with regular_sampling=false, the reference's default "
                        "(integrators/common.py:22): every ray has its own pixel jitter"}
        if "lbfgs_ms_median" in R:
            line["lbfgs_iteration"] = {
                "ms_median": ms_lbfgs, "armijo_trials_mean": R["lbfgs_trials_mean"],
                "what": "one whole LinearLBFGS iteration as optimize() runs it (optimize.py:286-320, lbfgs.py:198-275): "
                        "A p, loss, backward, history update, two-loop recursion, A d, Armijo search, pattern update, "
                        "clamp; median over the timed iterations"}
        if R["check"] is not None:
            line["multi_gpu_check"] = R["check"]
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def D_steps(args):
    return max(2, min(args.steps, 5))


if __name__ == "__main__":
    sys.exit(main())
