#!/usr/bin/env python
"""bench.py — one JSON line per run (driver contract in the task description).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload p3m|pm|pp|pk] [--impl ours|reference]

A "step" is one pass of the hot path over one batch of synthetic input:
  p3m  (default) BASELINE config 5: one P3M kick-drift step, 512^3 particles / 1024^3 mesh
  pm             BASELINE config 4: one PM step, 256^3 particles / 512^3 mesh
  pp             BASELINE config 3: one PP kick-drift step, 128^3 particles
  pk             BASELINE config 2: snapshot P(k), 128^3 particles / 256^3 mesh
`value` is measured with inputs resident in HBM; `e2e` goes through the numpy-level call with
pinned HOST buffers (H2D of pos+mom and D2H of the updated pos+mom inside the timed region).
`--impl reference` times the reference-side CPU path (the oracle port: float64 numpy/scipy +
C loops, all host threads) on a bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
_P3M_V = os.environ.get("NB_P3M_KERNEL", "7")
P3M_KERNEL_LABEL = {"7": "p3m_short (K5, v6: CTA tile shared by 4 z-adjacent cells, 8 i x 4 j-streams, test all / "
                         "evaluate in-range)",
                    "6": "p3m_short (K5, v5: v4's pair loop on a sliding z-window tile)"}.get(
    _P3M_V, f"p3m_short (K5, v{_P3M_V}: tile per warp, 8 i x 4 j-streams, test all / evaluate in-range)")
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

WORKLOADS = {
    # name: (method, n particles per side, mesh, boxsize, metric, unit)
    "p3m": ("p3m", 512, 1024, 1000.0, "p3m_particle_steps_per_s", "particle-steps/s"),
    "pm": ("pm", 256, 512, 1000.0, "pm_particle_steps_per_s", "particle-steps/s"),
    "pp": ("pp", 128, 0, 500.0, "pp_ginteractions_per_s", "Ginteractions/s"),
    "pk": ("pk", 128, 256, 1000.0, "pk_particles_per_s", "particles/s"),
}
CPU_SAMPLE = {"p3m": (128, 256), "pm": (128, 256), "pp": (64, 0), "pk": (128, 256)}   # bounded CPU samples
# --impl reference runs the CPU arm alone, so it can afford a sample 8x larger (1/8 of config 5 / the whole of
# config 4's particle count at half its mesh): ~20 s per step on 16 host threads
CPU_SAMPLE_REFERENCE = {"p3m": (256, 512), "pm": (256, 512), "pp": (64, 0), "pk": (128, 256)}
FP32_LANES_PER_SM = 128


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)", float(d.get("sm_max_mhz", 1965.0))
    return 6650.0, "fallback (B200_PROFILING.md)", 1965.0


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.index), "-lms", "200"], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, pw = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1])); pw.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                "power_w_max": max(pw), "samples": len(sm)}


def surrogate_pk_torch(k):
    q = k / 0.02
    return 2.0e4 * q / (1.0 + q ** 2.2) ** 1.4


def zeldovich_on_device(eng, n, L, seed, rms_disp_spacing=1.0):
    """Synthetic Zel'dovich set in lattice order: the reference's recipe
    (CiC/3D startbetingelser1.py:62-173) with a surrogate P(k); noise from torch's device RNG,
    δ_k/ψ synthesis by the library's own K7 kernel, ψ sampled at the lattice nodes."""
    import torch
    from nbody_b200 import tables
    dev = eng.device
    g = torch.Generator(device=dev).manual_seed(seed)
    c1 = torch.randn((n, n, n), generator=g, device=dev, dtype=torch.float64) / math.sqrt(2)
    c2 = torch.randn((n, n, n), generator=g, device=dev, dtype=torch.float64) / math.sqrt(2)
    kv = torch.from_numpy(tables.k_vals(n, L)).to(dev)
    k2 = kv * kv
    km = torch.sqrt((k2[:, None, None] + k2[None, :, None]) + k2[None, None, :])
    km[0, 0, 0] = 1e-10
    pk = surrogate_pk_torch(km).contiguous()
    del km
    _d, psi = eng.grf_synthesize(pk, c1, c2, L, n, guard=True, dtype=torch.float32, want_delta=False, want_psi=True)
    del pk, c1, c2
    disp = psi.reshape(3, -1).t().to(torch.float64)
    del psi
    disp *= rms_disp_spacing * (L / n) / disp.std()
    lat = torch.arange(n, device=dev, dtype=torch.float64) * (L / n)
    pos = torch.stack(torch.meshgrid(lat, lat, lat, indexing="ij"), -1).reshape(-1, 3).contiguous()
    eng.displace_wrap(pos, disp.contiguous(), L)
    del disp
    torch.cuda.empty_cache()
    return pos


def zeldovich_on_host(n, L, seed, rms_disp_spacing=1.0):
    """Same recipe in numpy for the CPU baseline samples."""
    from oracle import restate
    rs = np.random.RandomState(seed)
    c1 = rs.normal(0, 1 / np.sqrt(2), size=(n, n, n))
    c2 = rs.normal(0, 1 / np.sqrt(2), size=(n, n, n))
    R = restate.hermitian_R(c1, c2, guard=True)
    km = restate.k_mag_grid(n, L)
    km[0, 0, 0] = 1e-10
    q = km / 0.02
    _d, psi = restate.grf_delta_psi(2.0e4 * q / (1.0 + q ** 2.2) ** 1.4, R, L, n)
    disp = np.column_stack([p.ravel() for p in psi])
    disp *= rms_disp_spacing * (L / n) / disp.std()
    pos = restate.lattice(L, n) + disp
    pos %= L
    return pos


# ------------------------------------------------------------------------------- CPU arm
def cpu_config(workload, n, N, L):
    from nbody_b200.params import ForceConfig
    method = WORKLOADS[workload][0]
    return ForceConfig(method=method if method != "pk" else "pm", npart=n ** 3, boxsize=L, gridsize=N,
                       interlace=(True, True), deconvolve=(True, True))


def cpu_step_once(workload, cfg, pos, mom, workers):
    """One step of the reference-side CPU path (oracle port) on host arrays."""
    from oracle import cpu_step
    if workload == "pk":
        return cpu_step.snapshot_pk(pos, cfg.boxsize, cfg.gridsize, workers=workers)
    if workload == "pp":
        # direct sum is O(Np^2): time an i-subset against all j and extrapolate (stated in `sample`)
        from oracle import cport
        return cport.pp_accel(pos, cfg.G, cfg.mass, cfg.eps, 0.0, 0, min(pos.shape[0], 2048))
    return cpu_step.step(cfg, pos, mom, 1e-3, 1e-3, workers=workers)


def host_threads():
    """Threads this process may use: its CPU affinity mask, not an inherited OMP_NUM_THREADS (under
    torch.distributed.run every rank is started with OMP_NUM_THREADS=1)."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return os.cpu_count() or 1


def cpu_measure(workload, steps=1, warmup=0, sample=None):
    from oracle import cport
    n, N = sample or CPU_SAMPLE[workload]
    L = WORKLOADS[workload][3]
    cfg = cpu_config(workload, n, N, L)
    pos = zeldovich_on_host(n, L, seed=11)
    mom = np.zeros_like(pos)
    cores = os.cpu_count() or 1
    threads = cport.set_num_threads(host_threads())     # OpenMP loops; scipy.fft gets workers=threads
    for _ in range(warmup):
        cpu_step_once(workload, cfg, pos, mom, threads)
    t0 = time.perf_counter()
    for _ in range(steps):
        cpu_step_once(workload, cfg, pos, mom, threads)
    dt = (time.perf_counter() - t0) / steps
    np3 = float(n ** 3)
    if workload == "pp":
        isub = min(n ** 3, 2048)
        value = isub * np3 / dt / 1e9
        sample_txt = f"{n}^3 particles, {isub} i-particles x all j per step (rate is size-independent)"
    else:
        value = np3 / dt
        sample_txt = (f"{n}^3 particles / {N}^3 mesh per step (1/{round((WORKLOADS[workload][1] / n) ** 3, 1)} of the "
                      "workload; particle-steps/s of the sample, i.e. extrapolated linearly in the particle count)")
    sample = sample_txt
    return {"value": value, "unit": WORKLOADS[workload][5], "cores": threads, "host_cores": cores, "kind": "port",
            "sample": sample, "sample_particles": n ** 3, "sample_mesh": N, "seconds_per_sample_step": dt,
            "extrapolated": workload != "pp" and n != WORKLOADS[workload][1],
            "omp_num_threads_env": os.environ.get("OMP_NUM_THREADS"),
            "threads_note": "deposit loop serial like the reference's numba loop; gather, pair loops (OpenMP) and "
                            "scipy.fft use all threads"}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    wl = args.workload
    method, n, N, L, metric, unit = WORKLOADS[wl]
    steps = max(1, min(args.steps, 2))
    warm = min(args.warmup, 1) if wl in ("pp", "pk") else 0      # a P3M sample step is ~20 s: no warm-up step
    cb = cpu_measure(wl, steps=steps, warmup=warm, sample=CPU_SAMPLE_REFERENCE[wl])
    cfgd = workload_config(wl, n, N, L)
    # what was actually timed: the sample's particle count and mesh, with the full workload beside it
    cfgd.update(particles=cb["sample_particles"], mesh=cb["sample_mesh"], extrapolated=cb["extrapolated"],
                full_workload={"particles": n ** 3, "mesh": N},
                note="CPU arm = the repo's float64 C/OpenMP + scipy.fft restatement (the reference has no force "
                     "solver of its own: CONCEPT is external), timed on a bounded sample of the workload")
    line = {"impl": "reference", "metric": metric, "value": cb["value"], "unit": unit, "n_gpus": args.gpus,
            "steps": steps, "warmup": warm, "ms_per_step": cb["seconds_per_sample_step"] * 1e3,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": cfgd,
            "cpu_baseline": cb,
            "e2e": {"value": cb["value"], "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def workload_config(wl, n, N, L):
    method = WORKLOADS[wl][0]
    c = {"workload": {"p3m": f"P3M step, {n}^3 particles on a {N}^3 mesh (BASELINE config 5)",
                      "pm": f"PM step, {n}^3 particles on a {N}^3 mesh (BASELINE config 4)",
                      "pp": f"PP direct-sum kick-drift step, {n}^3 particles (BASELINE config 3)",
                      "pk": f"snapshot P(k): {n}^3 particles on a {N}^3 mesh (BASELINE config 2)"}[wl],
         "particles": n ** 3, "mesh": N, "boxsize": L, "input": "Zel'dovich lattice set, rms displacement 1 spacing",
         "l2": "inputs larger than L2 (no flush needed)" if n >= 256 else "L2 flushed between timed iterations"}
    if method in ("pm", "p3m"):
        c.update(interlace=[True, True], deconvolve=[True, True], differentiation="Fourier")
    if method == "p3m":
        c.update(r_s_cells=1.25, r_cut_over_rs=4.5)
    if method in ("pp", "p3m"):
        c.update(softening="Plummer, eps = 0.025 L / cbrt(Np)")
    return c


# ------------------------------------------------------------------------------- GPU arm
def run_ours(args):
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        args.gpus = world
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from nbody_b200.engine import Engine
    from nbody_b200.force import ForceStepper
    from nbody_b200.params import ForceConfig

    wl = args.workload
    method, n, N, L, metric, unit = WORKLOADS[wl]
    if args.n:
        n = args.n
    if args.mesh:
        N = args.mesh
    eng = Engine(local)
    hbm_peak, peak_src, sm_max_mhz = peaks()
    fp32_nominal_tf = 2.0 * FP32_LANES_PER_SM * eng_num_sms(torch, local) * sm_max_mhz * 1e6 / 1e12
    # measured denominators: pure-FMA microkernels of the library, scalar and packed
    fp32_meas = {"ffma_tflops": eng.fp32_peak(packed=False), "ffma2_tflops": eng.fp32_peak(packed=True),
                 "nominal_tflops": fp32_nominal_tf,
                 "how": "nb_fp32_peak: 64 independent FMAs per thread and iteration, 2 x 1024 threads per SM, best of 5 "
                        "launches, CUDA events; nominal = 2 x 128 lanes x SMs x clocks.max.sm"}
    fp32_peak_tf = max(fp32_meas["ffma_tflops"], fp32_meas["ffma2_tflops"])
    np_total = n ** 3
    dt_kick, dt_drift = 1e-3, 1e-3
    cfg = ForceConfig(method=method if method != "pk" else "pm", npart=np_total, boxsize=L, gridsize=N,
                      interlace=(True, True), deconvolve=(True, True))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- inputs (resident) -------------------------------------------------------------
    if world > 1:
        from nbody_b200.dist import DistForceStepper, shard_particles
        pos_all = zeldovich_on_device(eng, n, L, seed=20250611)
        pos = shard_particles(pos_all, method, L, rank, world, N)
        del pos_all
        torch.cuda.empty_cache()
        if method == "pk":
            from nbody_b200.dist import DistPowerSpectrum, TorchComm
            stepper = DistPowerSpectrum(eng, TorchComm(), N, L)
        else:
            stepper = DistForceStepper(cfg, eng, rank, world)
    else:
        pos = zeldovich_on_device(eng, n, L, seed=20250611)
        stepper = ForceStepper(cfg, eng) if method != "pk" else None
    state = {"pos": pos, "mom": torch.zeros_like(pos)}
    mom = state["mom"]
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=eng.device) if n < 256 else None
    phases = {}

    def ev():
        e = torch.cuda.Event(enable_timing=True)
        e.record()
        return e

    def one_step(record=None):
        """One pass of the hot path on resident data; `record` collects (phase, start, end)."""
        def mark(name, fn):
            if record is None:
                return fn()
            a = ev(); r = fn(); b = ev()
            record.append((name, a, b))
            return r
        if world > 1:
            if args.subphases and getattr(stepper, "mesh", None) is not None:
                stepper.mesh.submark = mark
            if method == "pk":
                return stepper(state["pos"], np_total, mark=mark)
            # migration may hand back new tensors: the particle set of this rank is `state`
            state["pos"], state["mom"] = stepper.step(state["pos"], state["mom"], dt_kick, dt_drift, mark=mark,
                                                      n_total=np_total)
            return None
        c = cfg
        if method == "pk":
            g = mark("deposit", lambda: eng.cic_deposit(pos, L / N, N))
            mark("overdensity", lambda: eng.grid_affine(g, a=(L / N) ** 3, b=np_total / L ** 3, c=1.0))
            return mark("fft+shell_bin", lambda: eng.power_spectrum_bins(g, L, N, 300, norm=L ** 3 / N ** 6))
        if method == "pp":
            half = 0.5 * L
            p4 = mark("pack", lambda: eng.pp_pack(pos, origin=(half, half, half), mass=1.0))
            acc = stepper._acc_buf(pos.shape[0])
            mark("pp_accel", lambda: eng.pp_accel(p4, p4, c.eps, scale=c.G * c.mass, out=acc))
            return mark("kick_drift", lambda: eng.kick_drift(pos, mom, acc, c.mass * dt_kick,
                                                             dt_drift / (c.mass * c.a ** 2), L))
        f, fs = mark("pm_mesh(deposit x2, R2C x2, kspace, C2R x6)", lambda: stepper.pm_mesh(pos))
        if method == "pm":
            return mark("gather_kick_drift", lambda: eng.gather_kick_drift(
                f, fs, N, c.cell, pos, mom, None, 1.0, c.mass * dt_kick, dt_drift / (c.mass * c.a ** 2), L))
        acc = stepper._acc_buf(pos.shape[0])
        mark("gather", lambda: eng.gather_kick_drift(f, fs, N, c.cell, pos, None, acc, 1.0, 0.0, 0.0, L))
        srt = mark("cell_sort", lambda: eng.cell_sort(pos, L, stepper.ncs, mass=1.0))
        mark("p3m_short", lambda: eng.p3m_short(srt[0], srt[1], srt[2], L, stepper.ncs, c.r_s, c.r_cut, c.eps,
                                                scale=c.G * c.mass, out=acc, accumulate=True, **stepper._near(pos)))
        return mark("kick_drift", lambda: eng.kick_drift(pos, mom, acc, c.mass * dt_kick,
                                                         dt_drift / (c.mass * c.a ** 2), L))

    def product_step():
        """The step a user runs: ForceStepper.step / DistForceStepper.step; P(k): deposit ->
        overdensity -> FFT + bins."""
        if world > 1 or method == "pk":
            return one_step()
        return stepper.step(pos, mom, dt_kick, dt_drift)

    # ---- warm-up + timed region (device clock, max over ranks) -----------------------------
    for _ in range(max(args.warmup, 3 if args.warmup >= 3 else args.warmup)):
        product_step()
    sampler = ClockSampler(local)
    barrier()
    if rank == 0:
        sampler.start()
    launches0 = eng.launch_count()
    eng.profile_read()
    eng.profile_enable(True)
    recs = []
    t_start = ev()
    step_ms = []
    for _ in range(args.steps):
        if flush is not None:
            flush.zero_()
        a = ev()
        if world > 1 or method == "pk":
            one_step(recs)
        else:
            product_step()
        b = ev()
        step_ms.append((a, b))
    t_end = ev()
    barrier()
    launches = eng.launch_count() - launches0
    invariants = None
    if method != "pk":
        # state after W + K steps from the same initial set at every N: conserved / partition-independent
        # quantities, so the SCALE lines carry a correctness signal
        p_, m_ = state["pos"], state["mom"]
        inv = torch.cat([torch.tensor([float(p_.shape[0])], device=eng.device, dtype=torch.float64),
                         m_.sum(0), m_.abs().sum(0), p_.sum(0), (m_ * m_).sum().reshape(1)])
        if world > 1:
            dist.all_reduce(inv)
        inv = inv.cpu().tolist()
        invariants = {"n_particles_total": int(round(inv[0])), "n_particles_expected": int(np_total),
                      "mom_sum": inv[1:4], "mom_abs_sum": inv[4:7], "pos_sum": inv[7:10], "mom_sq_sum": inv[10],
                      "after_steps": int(args.warmup + args.steps),
                      "note": "identical initial set and step count at every N: n_particles_total is conserved, "
                              "|mom_sum| << mom_abs_sum (Newton's third law), pos_sum / mom_sq_sum agree across N to "
                              "the float32 force tolerance"}
    if world == 1 and method != "pk":
        # phase pass (not part of `value`): the same kernels one after the other on one stream,
        # CUDA events around each, for phase_ms and the per-kernel rooflines
        eng.profile_read()
        for _ in range(min(args.steps, 3)):
            if flush is not None:
                flush.zero_()
            one_step(recs)
        torch.cuda.synchronize()
    eng.profile_enable(False)
    clocks = sampler.stop() if rank == 0 else None
    ms = [a.elapsed_time(b) for a, b in step_ms]
    total_ms = sum(ms)
    if world > 1:
        t = torch.tensor([total_ms], device=eng.device, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms = float(t.item())
    ms_per_step = total_ms / args.steps
    for name, a, b in recs:
        phases.setdefault(name, []).append(a.elapsed_time(b))
    phase_ms = {k: sum(v) / len(v) for k, v in phases.items()}
    phase_by_rank = None
    if world > 1:      # every rank's phase times (diagnosis of load imbalance / exposed waits)
        allp = [None] * world
        dist.all_gather_object(allp, {k: round(v, 3) for k, v in phase_ms.items()})
        phase_by_rank = {k: [r.get(k) for r in allp] for k in phase_ms}
    work = float(np_total) ** 2 / 1e9 if method == "pp" else float(np_total)
    value = work / (ms_per_step * 1e-3)

    line = {"metric": metric, "value": value, "unit": unit, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": workload_config(wl, n, N, L),
            "phase_ms": {k: round(v, 4) for k, v in phase_ms.items()}, "gpu_launches": int(launches),
            "gpu_launches_note": "kernels of libnbody_b200.so only (cuFFT/cub launches are library code)"}
    if world == 1 and method != "pk":
        line["phase_ms_note"] = ("phases timed in a separate sequential pass (one stream); the timed step is "
                                 "ForceStepper.step" + (" (short range into its own buffer, mesh chain, one fused "
                                                        "gather + kick + drift tail)" if method == "p3m" else ""))
        ov = getattr(stepper, "overlap", False)
        if method == "p3m" and (ov is True or (ov == "auto" and n ** 3 >= stepper.overlap_min_particles)):
            line["phase_ms_note"] += (f"; in the timed step the pair kernel (persistent, first {stepper.p3m_sms} SMs) and "
                                      "the mesh chain (other SMs) run side by side after the cell sort, so ms_per_step "
                                      "is less than the sum of the phases")
            line["overlap"] = {"pair_kernel_sms": stepper.p3m_sms, "sequential_sum_ms": round(sum(phase_ms.values()), 3)}
    if world > 1 and method == "p3m" and "p3m_short(join)" in phase_ms:
        line["phase_ms_note"] = (f"phases are CUDA-event intervals on the step's own stream inside the timed steps; the pair "
                                 f"kernel runs on a side stream, confined to the first {stepper.p3m_sms} SMs, from the end of "
                                 "cell_sort to p3m_short(join) (the wait that is left when the mesh chain is done), so the "
                                 "mesh phases are times measured beside it on the other SMs")
        line["overlap"] = {"pair_kernel_sms": stepper.p3m_sms}
    if phase_by_rank is not None:
        line["phase_ms_by_rank"] = phase_by_rank
    if invariants is not None:
        line["invariants"] = invariants
    line["config"]["parallelism"] = ("single GPU" if world == 1 else
                                     f"{world} ranks, i-sharded + all-gather of j" if method == "pp" else
                                     f"{world} x-slabs: slab FFT all-to-all, ghost planes, halo particles")

    # ---- rooflines ------------------------------------------------------------------------
    line.update(rooflines(eng, method, cfg, state["pos"], phase_ms, n, N, L, hbm_peak, peak_src, fp32_peak_tf, world,
                          stepper, torch))
    line["fp32_peak"] = fp32_meas
    # ---- end to end through host buffers ----------------------------------------------------
    if world == 1:
        line["e2e"] = e2e_measure(torch, eng, method, cfg, stepper, pos, mom, one_step, args, work, unit, L, N,
                                  np_total)
    else:
        line["e2e"] = e2e_measure_dist(torch, dist, eng, method, state, one_step, args, work, unit, barrier,
                                       stepper=stepper, dts=(dt_kick, dt_drift), np_total=np_total)
    if wl == "p3m" and not args.no_secondary and not args.n:
        # BASELINE.json's metric names PM particle-steps/s and PP Ginteractions/s too: short runs of configs 4
        # and 3 after the headline workload, so that they are driver-observed
        del pos, mom, stepper
        state.clear()
        eng.release_scratch()
        torch.cuda.empty_cache()
        line["secondary"] = {w: secondary_measure(torch, dist, eng, w, rank, world, fp32_peak_tf, hbm_peak, barrier)
                             for w in ("pm", "pp")}
        pos = mom = None
    if rank == 0:
        line["clocks"] = clocks
        if world == 1 and not args.no_cpu:
            del pos, mom
            torch.cuda.empty_cache()
            line["cpu_baseline"] = cpu_measure(wl, steps=1)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def secondary_measure(torch, dist, eng, wl, rank, world, fp32_peak_tf, hbm_peak, barrier, steps=3, warmup=1):
    """Device-timed product steps of another BASELINE workload (inputs resident, CUDA events, max over
    ranks): config 4 (PM, 256^3 / 512^3) and config 3 (PP, 128^3)."""
    from nbody_b200.force import ForceStepper
    from nbody_b200.params import ForceConfig
    method, n, N, L, metric, unit = WORKLOADS[wl]
    try:
        cfg = ForceConfig(method=method, npart=n ** 3, boxsize=L, gridsize=N, interlace=(True, True),
                          deconvolve=(True, True))
        pos = zeldovich_on_device(eng, n, L, seed=20250611)
        if world > 1:
            from nbody_b200.dist import DistForceStepper, shard_particles
            shard = shard_particles(pos, method, L, rank, world, N)
            del pos
            st = {"pos": shard, "mom": torch.zeros_like(shard)}
            stepper = DistForceStepper(cfg, eng, rank, world)

            def step():
                st["pos"], st["mom"] = stepper.step(st["pos"], st["mom"], 1e-3, 1e-3, n_total=n ** 3)
        else:
            st = {"pos": pos, "mom": torch.zeros_like(pos)}
            stepper = ForceStepper(cfg, eng)

            def step():
                stepper.step(st["pos"], st["mom"], 1e-3, 1e-3)
        for _ in range(warmup):
            step()
        barrier()
        a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(steps):
            step()
        b.record()
        barrier()
        t = torch.tensor([a.elapsed_time(b) / steps], device=eng.device, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
        work = float(n ** 3) ** 2 / 1e9 if method == "pp" else float(n ** 3)
        out = {"metric": metric, "value": work / (ms * 1e-3), "unit": unit, "ms_per_step": ms, "steps": steps,
               "warmup": warmup, "n_gpus": world, "config": workload_config(wl, n, N, L),
               "timing": "whole product step (ForceStepper.step / DistForceStepper.step), inputs resident"}
        if method == "pp":
            # the step is the pair kernel plus a 2 M-particle pack and kick/drift (< 0.1 %): step-level fraction
            ach = float(n ** 3) ** 2 * PP_FLOP / (ms * 1e-3) / 1e12
            out["roofline"] = {"bound": "fp32", "achieved": ach, "peak": fp32_peak_tf * world, "unit": "TFLOP/s",
                               "frac": ach / (fp32_peak_tf * world), "flop_per_interaction": PP_FLOP,
                               "peak_source": "measured FMA microkernel x n_gpus"}
        del st, stepper
        eng.release_scratch()
        torch.cuda.empty_cache()
        return out
    except Exception as e:      # a secondary block must never take the headline line down
        return {"metric": metric, "error": f"{type(e).__name__}: {e}"}


def eng_num_sms(torch, dev):
    return torch.cuda.get_device_properties(dev).multi_processor_count


# flops per pair (stated in DESIGN.md): PP convention 20 (3 sub, 3 fma, rsqrt, 3 mul, 3 fma);
# the short-range split adds u²,u (2 mul + rsqrt), t (fma), degree-10 Horner (10 fma),
# u³Q (mul + fma), ·f (mul): 29 flop  ->  49 per in-range pair.
PP_FLOP, P3M_FLOP = 20.0, 49.0


def rooflines(eng, method, cfg, pos, phase_ms, n, N, L, hbm_peak, peak_src, fp32_peak_tf, world, stepper, torch):
    """Per-kernel rooflines of THIS rank (rank 0's are printed): algorithmic bytes / flops of the
    rank's share of the work over the kernel's CUDA-event duration."""
    out = {}
    prof = eng.profile_read()
    avg = {k: v[0] / v[1] for k, v in prof.items()}
    if world > 1 and "kspace" in phase_ms:
        avg["kspace"] = phase_ms["kspace"]      # pure kernel phases of dist.DistForceStepper
    if world > 1 and "shell_bin" in phase_ms:
        avg["pk_bin"] = phase_ms["shell_bin"]
    np_local = float(pos.shape[0])
    ng = float(N) ** 3
    modes = float(N) * N * (N // 2 + 1)
    hbm = []
    if "deposit" in avg:
        b = 24.0 * np_local + 8.0 * ng / world
        hbm.append({"kernel": "cic_deposit (K1, memset + scatter)", "bound": "hbm", "ms": avg["deposit"],
                    "achieved": b / avg["deposit"] / 1e6, "alg_bytes": b})
    if "kspace" in avg:
        b = (16.0 + 48.0) * modes / world
        hbm.append({"kernel": "kspace_fused (K2)", "bound": "hbm", "ms": avg["kspace"],
                    "achieved": b / avg["kspace"] / 1e6, "alg_bytes": b})
    if "pk_bin" in avg:
        b = 8.0 * modes / world
        hbm.append({"kernel": "pk_shell_bin (K6, lane-per-row)", "bound": "hbm", "ms": avg["pk_bin"],
                    "achieved": b / avg["pk_bin"] / 1e6, "alg_bytes": b})
    for name in ("gather_kick_drift", "gather"):
        if name in phase_ms:
            b = (96.0 if name == "gather_kick_drift" else 48.0) * np_local + 24.0 * ng / world
            hbm.append({"kernel": f"{name} (K3)", "bound": "hbm", "ms": phase_ms[name],
                        "achieved": b / phase_ms[name] / 1e6, "alg_bytes": b})
    if method == "pk" and "deposit" in phase_ms:
        b = 24.0 * np_local + 8.0 * ng
        hbm.append({"kernel": "cic_deposit (K1, memset + scatter)", "bound": "hbm", "ms": phase_ms["deposit"],
                    "achieved": b / phase_ms["deposit"] / 1e6, "alg_bytes": b})
    for h in hbm:
        h.update(peak=hbm_peak, unit="GB/s", frac=h["achieved"] / hbm_peak, peak_source=peak_src, traffic=None)
    fp = None
    if method == "pp" and "pp_accel" in phase_ms:
        inter = np_local * float(n) ** 3        # this rank's i-particles x all j
        ach = inter * PP_FLOP / (phase_ms["pp_accel"] * 1e-3) / 1e12
        fp = {"kernel": "pp_accel (K4, FFMA2 j-tile)", "bound": "fp32", "ms": phase_ms["pp_accel"], "achieved": ach,
              "peak": fp32_peak_tf, "unit": "TFLOP/s", "frac": ach / fp32_peak_tf, "traffic": None,
              "flop_per_interaction": PP_FLOP, "interactions_per_launch": inter,
              "peak_source": "measured FMA microkernel (max of FFMA / FFMA2, see fp32_peak)"}
    if method == "p3m" and ("p3m_short" in phase_ms or world > 1):
        if world == 1:
            srt = eng.cell_sort(pos, L, stepper.ncs, mass=1.0)
            cand, inr = eng.p3m_count_pairs(srt[0], srt[2], L, stepper.ncs, cfg.r_cut)
        else:   # own + halo set of the last step; halo copies are sources only
            halo = stepper.halo_exchange(pos)
            srt = eng.cell_sort(torch.cat([pos, halo]), L, stepper.ncs, mass=1.0)
            cand, inr = eng.p3m_count_pairs(srt[0], srt[2], L, stepper.ncs, cfg.r_cut, perm=srt[1],
                                            i_limit=pos.shape[0])
        k5_note = None
        if "p3m_short" not in phase_ms:
            # overlapped sharded step: the pair kernel ran on a side stream beside the mesh chain; for the roofline
            # it is timed alone here, on all SMs, on the same sorted own + halo set
            lay = stepper.mesh.lay
            cs_ = L / stepper.ncs
            plane = stepper.ncs ** 2
            lo = max(0, int(math.floor(lay.x0 * cfg.cell / cs_)) - 1) * plane
            hi = min(stepper.ncs, int(math.floor((lay.x0 + lay.nxl) * cfg.cell / cs_)) + 2) * plane
            acc_ = stepper._acc_buf(pos.shape[0])
            tms = []
            for it in range(3):
                a_ = torch.cuda.Event(enable_timing=True); b_ = torch.cuda.Event(enable_timing=True)
                a_.record()
                eng.p3m_short(srt[0], srt[1], srt[2], L, stepper.ncs, cfg.r_s, cfg.r_cut, cfg.eps, scale=cfg.G * cfg.mass,
                              out=acc_, accumulate=False, cell_lo=lo, cell_hi=hi, i_limit=pos.shape[0])
                b_.record(); b_.synchronize()
                tms.append(a_.elapsed_time(b_))
            phase_ms = dict(phase_ms, p3m_short=sum(tms[1:]) / 2)
            k5_note = "timed alone after the timed region (in the step it runs beside the mesh chain on an SM subset)"
        t = phase_ms["p3m_short"] * 1e-3
        ach = inr * P3M_FLOP / t / 1e12
        fp = {"kernel": P3M_KERNEL_LABEL, "bound": "fp32", "ms": phase_ms["p3m_short"],
              "achieved": ach, "peak": fp32_peak_tf, "unit": "TFLOP/s", "frac": ach / fp32_peak_tf, "traffic": None,
              "flop_per_inrange_pair": P3M_FLOP, "inrange_pairs": inr, "candidate_pairs": cand,
              "candidate_pairs_per_s": cand / t,
              "peak_source": "measured FMA microkernel (max of FFMA / FFMA2, see fp32_peak)"}
        if k5_note:
            fp["ms_note"] = k5_note
        if world == 1 and cfg.eps > 0 and 20.0 * cfg.eps < cfg.r_cut:
            # the same launch in the exact mode (pairs inside 20 eps and the cut-off shell in float64; off by default)
            acc_ = stepper._acc_buf(pos.shape[0])
            cnt_ = torch.zeros(1, dtype=torch.int32, device=eng.device)
            tms = []
            for it in range(3):
                a_ = torch.cuda.Event(enable_timing=True); b_ = torch.cuda.Event(enable_timing=True)
                a_.record()
                eng.p3m_short(srt[0], srt[1], srt[2], L, stepper.ncs, cfg.r_s, cfg.r_cut, cfg.eps, scale=cfg.G * cfg.mass,
                              out=acc_, accumulate=False, pos=pos, r_near=20.0 * cfg.eps, near_count=cnt_)
                b_.record(); b_.synchronize()
                tms.append(a_.elapsed_time(b_))
            fp["exact_mode"] = {"ms": sum(tms[1:]) / 2, "near_pairs_listed": int(cnt_.item()),
                                "note": "nb_p3m_short_exact (NB_P3M_NEAR=1): 1e-5 in the max norm, not the timed default"}
    allk = hbm + ([fp] if fp else [])
    # DRAM traffic per launch from the committed ncu --set full captures (profiles/ncu_traffic.json);
    # it cannot be measured live, so it is only filled in for the exact size that was captured
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as fh:
            traffic_tab = json.load(fh)["entries"]
    except (OSError, ValueError, KeyError):
        traffic_tab = []
    for r in allk:
        for e in traffic_tab:
            if world == 1 and e["match"] in r["kernel"] and e["particles"] == n ** 3 and e["mesh"] == N:
                r["traffic"] = float(e["bytes"])
                r["traffic_source"] = e["source"]
    if allk:
        dom = max(allk, key=lambda r: r["ms"])
        out["roofline"] = dom
        out["rooflines_all"] = allk
    if prof:
        out["library_phase_ms"] = {k: round(v[0] / v[1], 4) for k, v in prof.items()}
    return out


def e2e_measure(torch, eng, method, cfg, stepper, pos, mom, one_step, args, work, unit, L, N, np_total):
    """Same step through HOST buffers: pinned pos/mom -> device, step, updated pos/mom -> pinned
    host, every step, copies inside the timed region (CUDA events on the launching stream)."""
    steps = max(1, min(args.steps, 3))
    if method == "pk":
        hpos = torch.empty(pos.shape, dtype=torch.float64, pin_memory=True)
        hpos.copy_(pos)
        hout = torch.empty(2 * 299, dtype=torch.float64, pin_memory=True)
        h2d, d2h = hpos.numel() * 8, 299 * 16
    else:
        hpos = torch.empty(pos.shape, dtype=torch.float64, pin_memory=True)
        hmom = torch.empty(pos.shape, dtype=torch.float64, pin_memory=True)
        hpos.copy_(pos); hmom.copy_(mom)
        opos = torch.empty(pos.shape, dtype=torch.float64, pin_memory=True)
        omom = torch.empty(pos.shape, dtype=torch.float64, pin_memory=True)
        h2d, d2h = 2 * hpos.numel() * 8, 2 * hpos.numel() * 8
    torch.cuda.synchronize()

    def host_step():
        if method == "pk":
            pos.copy_(hpos, non_blocking=True)
            _e, counts, sums = one_step()
            hout[:299].copy_(counts.to(torch.float64), non_blocking=True)
            hout[299:].copy_(sums, non_blocking=True)
        else:   # the public host-buffer step (mom upload overlapped with the force computation)
            stepper.step_host(hpos, hmom, opos, omom, 1e-3, 1e-3)

    host_step()
    torch.cuda.synchronize()
    a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps):
        host_step()
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / steps
    out = {"value": work / (ms * 1e-3), "unit": unit, "ms_per_step": ms, "steps": steps,
           "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
           "api": ("nbody_b200.force.ForceStepper.step_host on pinned host pos/mom (H2D pos, mom upload overlapped "
                   "with the force computation, D2H pos+mom)" if method != "pk" else
                   "deposit + overdensity + power_spectrum_bins on pinned host positions, (count,sum) bins back")}
    if method in ("pm", "p3m"):
        # the same work for a stream of independent particle sets: two calls in flight (double-buffered device arrays),
        # so the upload of call k+1 overlaps the computation and download of call k; every call moves the same bytes
        calls = 8
        for _ in range(2):
            stepper.step_host_pipelined(hpos, hmom, opos, omom, 1e-3, 1e-3)
        stepper.finish_host()
        torch.cuda.synchronize()
        a.record()
        for _ in range(calls):
            stepper.step_host_pipelined(hpos, hmom, opos, omom, 1e-3, 1e-3)
        stepper.finish_host()
        b.record()
        torch.cuda.synchronize()
        msp = a.elapsed_time(b) / calls
        out["pipelined"] = {"value": work / (msp * 1e-3), "unit": unit, "ms_per_step": msp, "calls": calls,
                            "calls_in_flight": 2,
                            "api": "ForceStepper.step_host_pipelined + finish_host: independent particle sets back to back, "
                                   "same H2D / D2H bytes per call as e2e; NOT the headline e2e (whose calls do not overlap)"}
    return out


def e2e_measure_dist(torch, dist, eng, method, state, one_step, args, work, unit, barrier, stepper=None, dts=(0.0, 0.0),
                     np_total=0):
    """N > 1: every rank stages ITS shard through pinned host memory each step (H2D of pos+mom,
    the sharded step, D2H of the updated shard); device time, max over ranks.  Force steps go through
    DistForceStepper.step_host (momentum upload on a side stream under the force computation)."""
    steps = max(1, min(args.steps, 3))
    pos0 = state["pos"]
    binding = None
    if os.environ.get("NB_BIND_HOST", "1") == "1":
        from nbody_b200.dist import bind_host_near_device
        binding = bind_host_near_device(eng.device.index if eng.device.index is not None else 0)
    hpos = torch.empty(pos0.shape, dtype=torch.float64, pin_memory=True)
    hpos.copy_(pos0)
    if method == "pk":
        hout = torch.empty(2 * 299, dtype=torch.float64, pin_memory=True)
        h2d, d2h = hpos.numel() * 8, 299 * 16
    else:
        hmom = torch.empty(pos0.shape, dtype=torch.float64, pin_memory=True)
        hmom.copy_(state["mom"])
        cap = int(pos0.shape[0] * 1.05) + 1024            # room for particles that migrate in
        opos = torch.empty((cap, 3), dtype=torch.float64, pin_memory=True)
        omom = torch.empty((cap, 3), dtype=torch.float64, pin_memory=True)
        h2d, d2h = 2 * hpos.numel() * 8, 2 * hpos.numel() * 8
    dpos, dmom = torch.empty_like(pos0), torch.empty_like(pos0)
    torch.cuda.synchronize()

    def host_step():
        if method != "pk":
            return stepper.step_host(hpos, hmom, opos, omom, dts[0], dts[1], n_total=np_total)
        dpos.copy_(hpos, non_blocking=True)
        state["pos"] = dpos
        if method != "pk":
            dmom.copy_(hmom, non_blocking=True)
            state["mom"] = dmom
        r = one_step()
        if method == "pk":
            _e, counts, sums = r
            hout[:299].copy_(counts.to(torch.float64), non_blocking=True)
            hout[299:].copy_(sums, non_blocking=True)
        else:
            k = min(state["pos"].shape[0], opos.shape[0])
            opos[:k].copy_(state["pos"][:k], non_blocking=True)
            omom[:k].copy_(state["mom"][:k], non_blocking=True)

    host_step()
    barrier()
    a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps):
        host_step()
    b.record()
    barrier()
    t = torch.tensor([a.elapsed_time(b) / steps], device=eng.device, dtype=torch.float64)
    bt = torch.tensor([float(h2d), float(d2h)], device=eng.device, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dist.all_reduce(bt)
    ms = float(t.item())
    return {"value": work / (ms * 1e-3), "unit": unit, "ms_per_step": ms, "steps": steps,
            "h2d_bytes_per_step": int(bt[0].item()), "d2h_bytes_per_step": int(bt[1].item()),
            "host_binding_rank0": binding if binding is not None else "none (topology unreadable or already bound)",
            "api": ("nbody_b200.dist.DistPowerSpectrum, each rank's shard staged through pinned host memory" if method == "pk"
                    else "nbody_b200.dist.DistForceStepper.step_host on each rank's pinned host shard (H2D pos, mom upload "
                         "under the force computation, D2H of the migrated shard)")}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="p3m", choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--n", type=int, default=0, help="override particles per side (development only)")
    ap.add_argument("--mesh", type=int, default=0, help="override mesh size (development only)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg (development only)")
    ap.add_argument("--no-secondary", action="store_true", help="skip the PM / PP secondary blocks of the default line")
    ap.add_argument("--subphases", action="store_true", help="finer CUDA-event marks inside the slab FFT (development only)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_ours(args)


if __name__ == "__main__":
    main()
