"""TEST INFRASTRUCTURE ONLY — generates tests/golden/* by running the REFERENCE's own
functions (loaded by AST extraction, oracle/ref_loader.py) in the build container.

    python oracle/gen_golden.py          # needs /root/reference; writes tests/golden/

What is stored
  snap_{tag}.npz        pos (Np,3) f8 + boxsize of four 32^3 CONCEPT snapshots shipped with
                        the reference, and `delta_golden`: the reference's stored
                        delta_{test|train}_id-{i}.npy for that snapshot
  ref_pk_snapshot.npz   compute_power_spectrum_from_snapshot (Power_compare.py:135-200) for
                        the 4 flag combinations, the Plotter.py:148-202 and the
                        1vs2tlptplotter.py:81-116 variants, on snapshot 'test1' and on the
                        L=1000 3giant32pm snapshot
  ref_pk_field.npz      compute_power_spectrum (digitize loop; Power_compare.py:422-465,
                        CiC/3D startbetingelser1.py:274-315, Plotter.py:294-333) on a 16^3 field
  ref_grf.npz           createR (Cosmo data create (new).py:24-47), find_delta (:114-137),
                        CLASS_simualtor (Power_compare.py:279-331) and particle_assigment
                        (:336-418) with CLASS replaced by a surrogate P(k) (CLASS is absent)
  ref_window.npz        cic_window_ft both variants at N=8,16
  concept_powerspec/    the five CONCEPT `powerspec_a=*` text files the reference ships
                        (PPPMP3M/{PP,PM,P3M}/powerspec/powerspec_a=1.00, SnapVsPK/powerspec/
                        powerspec_a={0.17,1.00}), copied byte for byte: the known answers of the
                        CONCEPT-style estimator (modes / k columns) and of the full-run P(k) pin
  snap_giant32_mom.npz  pos + mom + attributes of Different_Sizes/newrun2/3giant32p3m (32^3, L=1000,
                        a=1) and mom of the 3giant32pm snapshot (momentum statistics of a full run)
The CLASS substitution: the reference functions call `power_spectrum(...)` /
`get_matter_power_spectrum(...)`; a stand-in with the same signature returning
`surrogate_pk(k)` is injected into their namespace.  Everything else is reference code.
"""
from __future__ import annotations

import glob
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)

from oracle import ref_loader  # noqa: E402
from nbody_b200 import snapshot  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def surrogate_pk(k):
    """Smooth BBKS-like P(k) stand-in for CLASS (same array goes to oracle and GPU)."""
    k = np.asarray(k, dtype=np.float64)
    q = k / 0.02
    return 2.0e4 * q / (1.0 + q ** 2.2) ** 1.4


class FakeH5File:
    """h5py.File stand-in so reference functions that open the snapshot themselves run
    unmodified: f['components/matter/pos'][:], f.attrs['boxsize']."""

    def __init__(self, path, mode="r"):
        self.path = path
        self.attrs = snapshot.read_attrs(path)

    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False

    def __getitem__(self, key):
        return snapshot._H5(self.path).dataset(key)


class FakeH5py:
    File = FakeH5File


def golden_pairs():
    base = os.path.join(ref_loader.REF_ROOT, "TheFeedingPipeline", "Datafolder4")
    test = snapshot.find_hdf5_files(os.path.join(base, "Test_Data_concept"))
    train = snapshot.find_hdf5_files(os.path.join(base, "TrainingVal_Concept"))
    pick = {"test1": (test, 1, "Test data_CONCEPT/delta_test_id-1.npy"),
            "test57": (test, 57, "Test data_CONCEPT/delta_test_id-57.npy"),
            "train1": (train, 1, "Training & val data_CONCEPT/delta_train_id-1.npy"),
            "train200": (train, 200, "Training & val data_CONCEPT/delta_train_id-200.npy")}
    for tag, (files, idx, rel) in pick.items():
        yield tag, files[idx - 1], os.path.join(base, rel)


def main():
    assert ref_loader.available(), "reference tree not found"
    os.makedirs(OUT, exist_ok=True)

    # ---- 1. snapshots + shipped golden deltas -------------------------------------
    frap = ref_loader.load("frap", ["cic_density", "compute_delta", "cic_window_ft",
                                    "deconvolve_cic"])
    for tag, snap, gold in golden_pairs():
        pos = snapshot.read_positions(snap)
        box = snapshot.read_boxsize(snap)
        g = np.load(gold)
        d = frap["deconvolve_cic"](frap["compute_delta"](frap["cic_density"](pos, box, 32)), box, 32)
        if tag.startswith("train"):
            d -= d.mean()
        err = np.abs(d - g).max()
        print(f"{tag}: reference fn vs shipped golden max-abs {err:.2e}")
        assert err < 1e-12
        if tag == "test1":   # the raw CONCEPT file itself: fixture of the a14 reader test (data, not source)
            import shutil
            shutil.copyfile(snap, os.path.join(OUT, "snapshot_test1.hdf5"))
        np.savez(os.path.join(OUT, f"snap_{tag}.npz"), pos=pos, boxsize=box, delta_golden=g,
                 source=os.path.relpath(snap, ref_loader.REF_ROOT))
    giant = os.path.join(ref_loader.REF_ROOT, "Different_Sizes/newrun2/3giant32pm/snapshots/snapshot_a=1.00.hdf5")
    gpos = snapshot.read_positions(giant)
    np.savez(os.path.join(OUT, "snap_giant32pm.npz"), pos=gpos, boxsize=snapshot.read_boxsize(giant),
             source=os.path.relpath(giant, ref_loader.REF_ROOT))

    # ---- 2. P(k) of a snapshot, all variants ---------------------------------------
    first = next(golden_pairs())[1]
    out = {}
    pc = ref_loader.load("power_compare", ["cic_window_ft", "deconvolve_cic", "cic_density",
                                           "compute_power_spectrum_from_snapshot"],
                         extra_ns={"h5py": FakeH5py, "N": 32})
    for dec in (False, True):
        for fac in (False, True):
            k, p = pc["compute_power_spectrum_from_snapshot"](first, 500.0, 32, deconvolve=dec, use_fac=fac)
            out[f"pc_dec{int(dec)}_fac{int(fac)}_k"] = k
            out[f"pc_dec{int(dec)}_fac{int(fac)}_p"] = p
    # finer mesh than particles (the BASELINE config-2 shape, scaled down): 32^3 -> 64^3
    k, p = pc["compute_power_spectrum_from_snapshot"](first, 500.0, 64, deconvolve=False, use_fac=False)
    out["pc_mesh64_k"], out["pc_mesh64_p"] = k, p
    pl = ref_loader.load("plotter", ["cic_assign", "compute_power_spectrum_from_snapshot"],
                         extra_ns={"h5py": FakeH5py})
    k, p = pl["compute_power_spectrum_from_snapshot"](giant, 1000.0, 32)
    out["plotter_giant_k"], out["plotter_giant_p"] = k, p
    k, p = pl["compute_power_spectrum_from_snapshot"](first, 500.0, 32)
    out["plotter_test1_k"], out["plotter_test1_p"] = k, p
    lp = ref_loader.load("lpt", ["cic_assign", "compute_power_spectrum_from_snapshot"],
                         extra_ns={"h5py": FakeH5py})
    k, p = lp["compute_power_spectrum_from_snapshot"](giant, 1000.0, 32)
    out["lpt_giant_k"], out["lpt_giant_p"] = k, p
    np.savez(os.path.join(OUT, "ref_pk_snapshot.npz"), **out)

    # ---- 3. GRF: createR / find_delta / CLASS_simualtor / particle_assigment ---------
    N, L = 16, 500.0
    out = {}
    kv = 2 * np.pi * np.fft.fftfreq(N, d=L / N)
    kx, ky, kz = np.meshgrid(kv, kv, kv, indexing="ij")
    k_mag = np.sqrt(kx ** 2 + ky ** 2 + kz ** 2)
    k_mag[0, 0, 0] = 1e-10

    def fake_power_spectrum(A_s=0, n_s=0, omega_b=0, omega_cdm=0, z=0, k=None, plot=False, returnk=False):
        pk = surrogate_pk(k.ravel())
        return (k.ravel(), pk) if returnk else pk

    grf = ref_loader.load("grf", ["createR", "find_delta"],
                          extra_ns={"power_spectrum": fake_power_spectrum, "k_mag": k_mag})
    np.random.seed(1234)
    out["createR_seed1234_N8"] = grf["createR"](8)
    np.random.seed(4321)
    out["find_delta_seed4321_N16"] = grf["find_delta"](0, L, N)
    out["find_delta_Pk_grid"] = surrogate_pk(k_mag)

    def fake_get_mps(z, k=None, returnk=False):
        kk = np.logspace(-5, 1, 2000)
        return kk, surrogate_pk(kk)

    pcg = ref_loader.load("power_compare", ["CLASS_simualtor", "particle_assigment",
                                            "compute_power_spectrum"],
                          use_numba=False, extra_ns={"get_matter_power_spectrum": fake_get_mps})
    np.random.seed(99)
    delta_x, psi, _k = pcg["CLASS_simualtor"](L, N, 0)
    out["sim_seed99_delta"] = delta_x
    out["sim_seed99_psi"] = np.stack(psi)
    # the P_k_grid the simulator used (interp1d of the tabulated surrogate on clipped k)
    import scipy.interpolate as scint
    kk, pp = fake_get_mps(0)
    out["sim_Pk_grid"] = scint.interp1d(kk, pp, bounds_error=False, fill_value="extrapolate")(
        np.clip(k_mag, kk.min(), kk.max()))
    dcic = pcg["particle_assigment"](psi, L, N)
    out["sim_seed99_delta_cic"] = dcic
    np.savez(os.path.join(OUT, "ref_grf.npz"), **out)

    # ---- 4. compute_power_spectrum (digitize loop), 3 variants -----------------------
    outp = {"delta_x": delta_x, "delta_cic": dcic}
    for wf in (False, True):
        k, p = pcg["compute_power_spectrum"](delta_x, fromcic=False, kx=kx, ky=ky, kz=kz, L=L, N=N, withfac=wf)
        outp[f"pc_fft_fac{int(wf)}_k"], outp[f"pc_fft_fac{int(wf)}_p"] = k, p
        k, p = pcg["compute_power_spectrum"](dcic, fromcic=True, kx=kx, ky=ky, kz=kz, L=L, N=N, withfac=wf)
        outp[f"pc_cic_fac{int(wf)}_k"], outp[f"pc_cic_fac{int(wf)}_p"] = k, p
    plf = ref_loader.load("plotter", ["compute_power_spectrum"])
    k, p = plf["compute_power_spectrum"](dcic, fromcic=True, kx=kx, ky=ky, kz=kz, L=L, N=N)
    outp["plotter_cic_k"], outp["plotter_cic_p"] = k, p
    np.savez(os.path.join(OUT, "ref_pk_field.npz"), **outp)

    # ---- 5. windows ------------------------------------------------------------------
    outw = {}
    for n in (8, 16):
        outw[f"frap_W_{n}"] = frap["cic_window_ft"](n, 500.0)
        outw[f"pc_W_{n}"] = pc["cic_window_ft"](n, 500.0)
    np.savez(os.path.join(OUT, "ref_window.npz"), **outw)
    print("golden vectors written to", OUT)


def copy_concept_powerspec():
    """The reference's five powerspec files -> tests/golden/concept_powerspec/ (byte copies) and the
    momenta of the two 32^3 a=1 snapshots under Different_Sizes/newrun2."""
    import shutil
    dst = os.path.join(OUT, "concept_powerspec")
    os.makedirs(dst, exist_ok=True)
    R = ref_loader.REF_ROOT
    for tag, rel in (("P3M_a=1.00", "PPPMP3M/P3M/powerspec/powerspec_a=1.00"),
                     ("PM_a=1.00", "PPPMP3M/PM/powerspec/powerspec_a=1.00"),
                     ("PP_a=1.00", "PPPMP3M/PP/powerspec/powerspec_a=1.00"),
                     ("SnapVsPK_a=0.17", "SnapVsPK/powerspec/powerspec_a=0.17"),
                     ("SnapVsPK_a=1.00", "SnapVsPK/powerspec/powerspec_a=1.00")):
        shutil.copyfile(os.path.join(R, rel), os.path.join(dst, tag + ".txt"))
    out = {}
    for tag in ("pm", "p3m"):
        f = os.path.join(R, "Different_Sizes", "newrun2", f"3giant32{tag}", "snapshots", "snapshot_a=1.00.hdf5")
        out[f"mom_{tag}"] = snapshot.read_momenta(f)
        if tag == "p3m":
            out["pos_p3m"] = snapshot.read_positions(f)
        at = snapshot.read_attrs(f)
        for k in ("boxsize", "a", "H0"):
            out[f"{k}_{tag}"] = float(at[k])
    np.savez_compressed(os.path.join(OUT, "snap_giant32_mom.npz"), **out)


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "concept":
        copy_concept_powerspec()
    else:
        main()
        copy_concept_powerspec()
