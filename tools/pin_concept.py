"""Full-run pin of the force path against the reference's own output files.

The reference holds no source for its force solver (CONCEPT), but it holds what full CONCEPT runs
produced: the binned P(k), σ₈ and mode counts of 64³-particle, L = 500 Mpc, a = 0.02 → 1 runs with
the PP, PM and P³M methods (PPPMP3M/{PP,PM,P3M}/powerspec/powerspec_a=1.00) and of a 64³,
L = 1000 Mpc PM run at a = 0.17 and 1 (SnapVsPK/powerspec/powerspec_a={0.17,1.00};
SnapVsPK/Uncertain.param).  This script repeats those runs with this package — ICs at a_begin from
the files' own linear spectrum, the comoving leapfrog of timeloop.evolve around ForceStepper.step,
the CONCEPT-style estimator of powerspec.py — and compares column by column:

  * `modes`, `k`           exact (grid geometry),
  * `P_linear`             the input spectrum, reconstructed so that its binned mean is the file's,
  * `P` / `P (corrected)`  per bin, our realisation-corrected mean over `--seeds` realisations
                           against the file's corrected column (the realisation correction removes
                           the linear part of the cosmic variance of both sides),
  * P³M / PM ratio per bin (same seed on each side, so cosmic variance cancels to first order):
                           the signature of the short-range force,
  * σ₈                     of the data column against the header values.

    python tools/pin_concept.py --backend gpu --seeds 4 --out profiles/r2_pin_concept.json
    python tools/pin_concept.py --backend cpu --seeds 1 --methods pm      # float64 oracle loop

`--backend cpu` drives the same loop with the float64 C/numpy oracle (oracle/cpu_step.py) and the
numpy estimator (oracle/concept_ps.py): test infrastructure, used to cross-check the device result.
"""
from __future__ import annotations

import argparse
import json
import math
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from nbody_b200 import powerspec as PS  # noqa: E402
from nbody_b200 import timeloop  # noqa: E402
from nbody_b200.cosmo import Cosmology  # noqa: E402
from nbody_b200.params import ForceConfig, G_NEWTON  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden", "concept_powerspec")

#: the reference's runs: file tag -> (method, boxsize, a of the file, A_s)
RUNS = {
    "P3M_a=1.00": ("p3m", 500.0, 1.0), "PM_a=1.00": ("pm", 500.0, 1.0), "PP_a=1.00": ("pp", 500.0, 1.0),
    "SnapVsPK_a=0.17": ("pm", 1000.0, 1.0 / 6.0), "SnapVsPK_a=1.00": ("pm", 1000.0, 1.0),
}


class TabulatedLinear:
    """Linear P(k, a=1) whose mode-averaged bin means equal the `P_linear` column of a reference file:
    the closed-form spectrum of cosmo.py (Eisenstein-Hu with baryons, standing in for CLASS) times a
    correction interpolated linearly in ln k through the 16 bins, constant outside them.  Fixed-point
    passes make the binned mean match the column to < 1e-6."""

    def __init__(self, cosmo, ref: PS.PowerSpec, boxsize, a_file=1.0, A_s=2.105e-9, n_s=0.9649):
        self.cosmo, self.A_s, self.n_s = cosmo, A_s, n_s
        self.lnk = np.log(ref.k)
        self.corr = np.ones(ref.k.size)
        geom = PS.bin_geometry(ref.gridsize or 64)
        target = ref.linear / (cosmo.growth(a_file) / cosmo.growth(1.0)) ** 2
        for _ in range(24):
            have, _p = PS.linear_column(geom, boxsize, self)
            self.corr *= target / have
        self.residual = float(np.abs(PS.linear_column(geom, boxsize, self)[0] / target - 1).max())

    def __call__(self, k):
        k = np.asarray(k, dtype=np.float64)
        out = np.zeros_like(k)
        ok = k > 0
        base = self.cosmo.linear_power(k[ok], A_s=self.A_s, n_s=self.n_s)
        out[ok] = base * np.interp(np.log(k[ok]), self.lnk, self.corr)
        return out


def force_cfg(method, n, L, cosmo, mesh=None):
    rho_m = 3.0 * cosmo.H0 ** 2 / (8.0 * math.pi * G_NEWTON) * cosmo.Omega_m
    cfg = ForceConfig(method=method, npart=n ** 3, boxsize=L, gridsize=int(mesh or n), interlace=(True, True),
                      deconvolve=(True, True), G=G_NEWTON, mass=rho_m * L ** 3 / n ** 3)
    cfg.H0, cfg.Omega_b, cfg.Omega_cdm, cfg.a_begin = cosmo.H0, cosmo.Omega_b, cosmo.Omega_cdm, 0.02
    return cfg


# ------------------------------------------------------------------------------------ back ends
def run_gpu(cfg, cosmo, n, L, lin, seed, a_outs, dlna, lpt, offset=0.0, est_interlace=True, sub=None):
    """ICs, leapfrog and estimator on the device (the product path)."""
    import torch
    from nbody_b200 import api, force as F
    from nbody_b200.engine import default_engine
    eng = default_engine()
    c1, c2 = api._draws(n, None, None, seed)
    pos, mom = timeloop.zeldovich_ics(cosmo, L, n, cfg.a_begin, cfg.mass, c1=c1, c2=c2, engine=eng, lpt=lpt,
                                      linear_power=lin, lattice_offset=offset)
    noise2 = np.abs(api.createR(n, c1=c1, c2=c2, guard=True)) ** 2
    stepper = F.ForceStepper(cfg, eng)
    if cfg.method == "pp":
        stepper = timeloop.PeriodicPP(stepper)
    edges, idx = timeloop.edges_through(cfg.a_begin, a_outs, dlna)
    out = {}

    def emit(a, p, m):
        g = (cosmo.growth(a) / cosmo.growth(1.0)) ** 2
        ps = PS.concept_powerspec(p, L, n, linear=lambda k: g * lin(k), noise_abs2=noise2, interlace=est_interlace,
                                  tophat=8.0 / cosmo.h, a=a, t=cosmo.cosmic_time(a), engine=eng)
        out[round(a, 6)] = (ps, float(torch.sqrt((m ** 2).mean()).item()))

    stats = []
    if sub and cfg.method != "pm":
        timeloop.evolve_subcycled(stepper, cosmo, pos, mom, edges, output_at=idx, on_output=emit, eta=sub[0],
                                  max_sub=sub[1], stats=stats)
    else:
        timeloop.evolve(stepper, cosmo, pos, mom, edges, output_at=idx, on_output=emit)
    return out, (len(edges) - 1, int(sum(stats)) if stats else len(edges) - 1)


def run_cpu(cfg, cosmo, n, L, lin, seed, a_outs, dlna, lpt, offset=0.0, est_interlace=True, sub=None):
    """Same run with the float64 oracle: numpy ICs, C/numpy force loop, numpy estimator."""
    from oracle import concept_ps, cpu_step, restate
    if lpt != 1:
        raise SystemExit("the oracle loop has 1LPT ICs only")
    rs = np.random.RandomState(seed)
    c1 = rs.normal(0, 1 / np.sqrt(2), size=(n, n, n))
    c2 = rs.normal(0, 1 / np.sqrt(2), size=(n, n, n))
    R = restate.hermitian_R(c1, c2, guard=True)
    pk = timeloop.linear_pk_grid(cosmo, L, n, 0, 0, linear_power=lin)
    _, psi = restate.grf_delta_psi(pk, R, L, n)
    psi = np.stack([p.ravel() for p in psi], axis=1)
    d_rel = cosmo.growth(cfg.a_begin) / cosmo.growth(1.0)
    q = restate.lattice(L, n) + offset * L / n
    pos = (q + d_rel * psi) % L
    mom = cfg.mass * cfg.a_begin ** 2 * float(cosmo.H(cfg.a_begin)) * cosmo.growth_rate(cfg.a_begin) * d_rel * psi

    from oracle import cport, force as oforce

    class Stepper:
        """numpy stand-in with ForceStepper's interface (in-place updates like the device tensors)."""

        def __init__(self):
            self.cfg = cfg

        def step(self, p, m, dk, dd):
            return cpu_step.step(cfg, p, m, dk, dd)

        def accel_long(self, p):
            return None if cfg.method == "pp" else cpu_step.pm_accel(cfg, p)

        def accel_short(self, p):
            if cfg.method == "pm":
                return None
            if cfg.method == "pp":
                return oforce.pp_accel_ewald(p, L, cfg.G, cfg.mass, cfg.eps)
            return cport.p3m_short_accel(p, L, cfg.r_s, cfg.r_cut, cfg.G, cfg.mass, cfg.eps)

        def kick(self, p, m, acc, k):
            m += acc * k

        def kick_drift(self, p, m, acc, k, d):
            if acc is not None:
                m += acc * k
            p += m * d
            p %= L

        def max_norm2(self, acc):
            return float((acc * acc).sum(1).max())

    geom = PS.bin_geometry(n)
    ranges = [(int(f), int(l)) for f, l in zip(geom.first_k2, list(geom.first_k2[1:] - 1) + [geom.k2_end - 1])]
    edges, idx = timeloop.edges_through(cfg.a_begin, a_outs, dlna)
    out = {}

    def emit(a, p, m):
        g = (cosmo.growth(a) / cosmo.growth(1.0)) ** 2
        modes, P = concept_ps.powerspec(p, L, n, ranges, interlace=est_interlace)
        ratio = concept_ps.noise_ratio(R, L, n, ranges, lin)
        linc, _ = PS.linear_column(geom, L, lambda k: g * lin(k))
        ps = PS.PowerSpec(2 * math.pi / L * geom.k_grid, modes, P, corrected=P * ratio, linear=linc, gridsize=n,
                          boxsize=L, a=a, tophat=8.0 / cosmo.h)
        ps.sigma = tuple(PS.sigma_tophat(ps.k, c, ps.tophat) for c in ps.columns())
        out[round(a, 6)] = (ps, float(np.sqrt((m ** 2).mean())))

    stats = []
    if sub and cfg.method != "pm":
        timeloop.evolve_subcycled(Stepper(), cosmo, pos, mom, edges, output_at=idx, on_output=emit, eta=sub[0],
                                  max_sub=sub[1], stats=stats)
    else:
        timeloop.evolve(Stepper(), cosmo, pos, mom, edges, output_at=idx, on_output=emit)
    return out, (len(edges) - 1, int(sum(stats)) if stats else len(edges) - 1)


def giant32(a, cosmo, run):
    """The reference's two 32³, L = 1000 Mpc, a = 1 snapshots (Different_Sizes/newrun2/3giant32{pm,p3m},
    Giant.param with _size = 32: A_s = 2.1e-9) hold positions AND momenta: compare the rms momentum of our
    runs of the same set-up (mean over seeds) and, for P³M, the estimator's P(k) of the reference's
    particle positions with that of ours (32³: 10 bins, cosmic variance of single realisations on both sides)."""
    d = np.load(os.path.join(ROOT, "tests", "golden", "snap_giant32_mom.npz"))
    n, L = 32, 1000.0
    ref100 = PS.read_powerspec(os.path.join(GOLD, "SnapVsPK_a=1.00.txt"))
    lin0 = TabulatedLinear(cosmo, ref100, L, a_file=1.0)                 # CLASS shape of the L = 1000 file ...
    lin = lambda k: lin0(k) * (2.1e-9 / 2.105e-9)                        # noqa: E731  ... at Giant.param's A_s
    out = {}
    for method in ("pm", "p3m"):
        cfg = force_cfg(method, n, L, cosmo)
        if a.deconv_order is not None:
            cfg.deconvolve = (a.deconv_order >= 1, a.deconv_order >= 2)
        rms, P = [], []
        for s in range(a.seeds):
            res, _ = run(cfg, cosmo, n, L, lin, a.seed0 + s, [1.0], a.dlna, a.lpt, a.lattice_offset,
                         bool(a.estimator_interlace), (a.eta, a.max_sub) if a.max_sub > 1 else None)
            ps, r = res[1.0]
            rms.append(r)
            P.append(ps.power)
        ref_rms = float(np.sqrt((d[f"mom_{method}"] ** 2).mean()))
        rec = {"mom_rms_ref": ref_rms, "mom_rms_mean": float(np.mean(rms)), "mom_rms_seeds": [float(x) for x in rms],
               "mom_rms_over_ref": float(np.mean(rms) / ref_rms), "mass": cfg.mass}
        if method == "p3m":
            if a.backend == "gpu":
                pr = PS.concept_powerspec(d["pos_p3m"], L, n, interlace=bool(a.estimator_interlace)).power
            else:
                from oracle import concept_ps
                geom = PS.bin_geometry(n)
                rg = [(int(f), int(l)) for f, l in zip(geom.first_k2, list(geom.first_k2[1:] - 1) + [geom.k2_end - 1])]
                pr = concept_ps.powerspec(d["pos_p3m"], L, n, rg, interlace=bool(a.estimator_interlace))[1]
            rec["P_of_reference_snapshot"] = pr.tolist()
            rec["P_mean"] = np.mean(P, 0).tolist()
            rec["P_over_reference_snapshot"] = (np.mean(P, 0) / pr).tolist()
            rec["modes"] = PS.bin_geometry(n).modes.tolist()
        out[method] = rec
        print(f"giant32 {method}: mom rms {np.mean(rms):.1f} (ref {ref_rms:.1f}, ratio {np.mean(rms) / ref_rms:.3f})"
              + (f", P / P(reference snapshot): {np.array2string(np.array(rec['P_over_reference_snapshot']), precision=2)}"
                 if method == "p3m" else ""), flush=True)
    return out


# ------------------------------------------------------------------------------------ driver
def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--backend", choices=("gpu", "cpu"), default="gpu")
    ap.add_argument("--methods", default="pm,p3m", help="comma list of pm,p3m,pp (L=500 runs) and snapvspk (L=1000 PM)")
    ap.add_argument("--seeds", type=int, default=4)
    ap.add_argument("--seed0", type=int, default=1000)
    ap.add_argument("--n", type=int, default=64)
    ap.add_argument("--dlna", type=float, default=0.025)
    ap.add_argument("--lpt", type=int, default=1)
    ap.add_argument("--softening", type=float, default=None, help="Plummer ε in units of L/cbrt(N)")
    ap.add_argument("--eta", type=float, default=0.025, help="accuracy parameter of the sub-step criterion")
    ap.add_argument("--max-sub", type=int, default=32, help="largest number of short-range sub-steps per base step; 1: off")
    ap.add_argument("--lattice-offset", type=float, default=0.0, help="IC lattice offset in particle spacings")
    ap.add_argument("--estimator-interlace", type=int, default=1, help="0: estimator without interlacing")
    ap.add_argument("--deconv-order", type=int, default=None, help="powers of the CIC window divided out of the mesh force")
    ap.add_argument("--out", default=None)
    a = ap.parse_args(argv)
    cosmo = Cosmology()
    run = run_gpu if a.backend == "gpu" else run_cpu
    report = {"backend": a.backend, "n": a.n, "dlna_max": a.dlna, "seeds": a.seeds, "lpt": a.lpt,
              "lattice_offset": a.lattice_offset, "eta": a.eta, "max_sub": a.max_sub, "estimator_interlace": bool(a.estimator_interlace), "runs": {}}
    spectra = {}
    for meth in a.methods.split(","):
        if meth == "giant32":
            report["giant32"] = giant32(a, cosmo, run)
            continue
        if meth == "snapvspk":
            method, L, tags, outs = "pm", 1000.0, ["SnapVsPK_a=0.17", "SnapVsPK_a=1.00"], [1.0 / 6.0, 1.0]
        else:
            method, L, tags, outs = meth, 500.0, [f"{meth.upper()}_a=1.00"], [1.0]
        refs = [PS.read_powerspec(os.path.join(GOLD, t + ".txt")) for t in tags]
        lin = TabulatedLinear(cosmo, refs[-1], L, a_file=1.0)
        cfg = force_cfg(method, a.n, L, cosmo)
        if a.deconv_order is not None:
            cfg.deconvolve = (a.deconv_order >= 1, a.deconv_order >= 2)
        if a.softening is not None:
            cfg.softening_frac = a.softening
        acc = {round(o, 6): [] for o in outs}
        t0 = time.time()
        for s in range(a.seeds):
            res, (nsteps, nshort) = run(cfg, cosmo, a.n, L, lin, a.seed0 + s, outs, a.dlna, a.lpt, a.lattice_offset,
                                        bool(a.estimator_interlace), (a.eta, a.max_sub) if a.max_sub > 1 else None)
            for k, v in res.items():
                acc[k].append(v)
        wall = time.time() - t0
        for tag, ref, o in zip(tags, refs, outs):
            got = acc[round(o, 6)]
            P = np.array([g[0].power for g in got])
            Pc = np.array([g[0].corrected for g in got])
            sig = np.array([g[0].sigma for g in got])
            rec = {
                "method": method, "boxsize": L, "a": o, "steps": nsteps, "short_range_evaluations": nshort,
                "wall_s": round(wall, 2),
                "modes_exact": bool(np.array_equal(got[0][0].modes, ref.modes)),
                "k_8_significant_figures": bool(all("%.7e" % x == "%.7e" % y for x, y in zip(got[0][0].k, ref.k))),
                "linear_column_max_rel": float(np.abs(got[0][0].linear / ref.linear - 1).max()),
                "k": ref.k.tolist(), "modes": ref.modes.tolist(),
                "P_ref": ref.power.tolist(), "P_ref_corrected": ref.corrected.tolist(),
                "P_mean": P.mean(0).tolist(), "P_corrected_mean": Pc.mean(0).tolist(),
                "P_corrected_over_ref_corrected": (Pc.mean(0) / ref.corrected).tolist(),
                "P_corrected_scatter": (Pc.std(0) / Pc.mean(0)).tolist() if len(got) > 1 else None,
                "sigma8_ref": list(ref.sigma), "sigma8_mean": sig.mean(0).tolist(),
                "mom_rms": float(np.mean([g[1] for g in got])),
            }
            report["runs"][tag] = rec
            spectra[tag] = Pc.mean(0)
            r = np.array(rec["P_corrected_over_ref_corrected"])
            print(f"{tag}: steps {nsteps} ({nshort} short-range), {wall:.1f} s, modes exact {rec['modes_exact']}, k exact "
                  f"{rec['k_8_significant_figures']}, sigma8 {sig.mean(0)[0]:.4f} (ref {ref.sigma[0]:.3f}), "
                  f"corrected P / ref: {np.array2string(r, precision=3)}", flush=True)
    if "P3M_a=1.00" in spectra and "PM_a=1.00" in spectra:
        r3, rm = (PS.read_powerspec(os.path.join(GOLD, t + ".txt")) for t in ("P3M_a=1.00", "PM_a=1.00"))
        ours = spectra["P3M_a=1.00"] / spectra["PM_a=1.00"]
        theirs = r3.corrected / rm.corrected
        report["p3m_over_pm"] = {"ours": ours.tolist(), "reference": theirs.tolist(), "ratio": (ours / theirs).tolist()}
        print("P3M/PM ours     :", np.array2string(ours, precision=3))
        print("P3M/PM reference:", np.array2string(theirs, precision=3))
    if a.out:
        os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
        with open(a.out, "w") as f:
            json.dump(report, f, indent=1)
    return report


if __name__ == "__main__":
    main()
