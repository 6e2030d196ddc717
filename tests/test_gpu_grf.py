"""GPU parity: Hermitian noise, GRF synthesis (find_delta, δ and ψ), gather and
particle_assigment against the reference's golden vectors (tests/golden/ref_grf.npz,
generated from the reference's own createR / find_delta / CLASS_simualtor /
particle_assigment with a surrogate P(k)) and against the float64 oracle.
Tolerances: float64 device path 1e-11 norm-wise; float32 path 1e-5 norm-wise."""
import numpy as np
import pytest

from conftest import golden, surrogate_pk
from oracle import restate

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def api():
    from nbody_b200 import api as _api
    return _api


def nrm(a, ref):
    return np.abs(np.asarray(a) - ref).max() / np.abs(ref).max()


def test_createR_bit_exact(api):
    g = golden("ref_grf.npz")
    R = api.createR(8, seed=1234)
    assert np.array_equal(R, g["createR_seed1234_N8"])                 # unguarded createR
    np.random.seed(7)
    c1 = np.random.normal(0, 1 / np.sqrt(2), size=(12, 12, 12))
    c2 = np.random.normal(0, 1 / np.sqrt(2), size=(12, 12, 12))
    for guard in (False, True):
        assert np.array_equal(api.createR(12, c1=c1, c2=c2, guard=guard), restate.hermitian_R(c1, c2, guard))
    for n in (5, 9):                                                     # odd N
        c1 = np.random.normal(size=(n, n, n)); c2 = np.random.normal(size=(n, n, n))
        assert np.array_equal(api.createR(n, c1=c1, c2=c2, guard=True), restate.hermitian_R(c1, c2, True))


@pytest.mark.parametrize("prec,tol", [("f64", 1e-11), ("f32", 1e-5)])
def test_find_delta_golden(api, prec, tol):
    g = golden("ref_grf.npz")
    d = api.find_delta(0, 500.0, 16, Pk_grid=g["find_delta_Pk_grid"], seed=4321, guard=False, precision=prec)
    assert d.dtype == np.float64 and nrm(d, g["find_delta_seed4321_N16"]) <= tol


@pytest.mark.parametrize("prec,tol", [("f64", 1e-11), ("f32", 1e-5)])
def test_delta_psi_golden(api, prec, tol):
    g = golden("ref_grf.npz")
    d, psi = api.grf_delta_psi(g["sim_Pk_grid"], 500.0, 16, seed=99, guard=True, precision=prec)
    assert nrm(d, g["sim_seed99_delta"]) <= tol
    assert nrm(np.stack(psi), g["sim_seed99_psi"]) <= tol


def test_particle_assigment_golden(api):
    g = golden("ref_grf.npz")
    psi = tuple(g["sim_seed99_psi"])
    dc, parts = api.particle_assigment(psi, 500.0, 16, precision="f64", return_particles=True)
    ref, rparts = restate.particle_assigment(psi, 500.0, 16, return_particles=True)
    assert np.array_equal(parts, rparts)                        # displaced + wrapped positions bit-exact
    assert np.abs(dc - g["sim_seed99_delta_cic"]).max() <= 1e-11
    assert nrm(api.particle_assigment(psi, 500.0, 16, precision="f32"), g["sim_seed99_delta_cic"]) <= 1e-5


def test_gather_is_exact_on_nodes_and_matches_oracle_off_nodes(api):
    rs = np.random.RandomState(2)
    N, L = 24, 300.0
    f = rs.normal(size=(3, N, N, N))
    lat = restate.lattice(L, N)
    out = api.trilinear_gather(f, lat, L, N)
    assert np.array_equal(out, f.reshape(3, -1).T)              # node values returned exactly (Q5)
    pos = rs.uniform(-0.2 * L, 1.3 * L, size=(5000, 3))
    out = api.trilinear_gather(f[0], pos, L, N)
    assert np.abs(out - restate.trilinear_gather(f[0], pos, L, N)).max() <= 1e-13
    assert api.trilinear_gather(f[0], np.zeros((0, 3)), L, N).shape == (0,)


@pytest.mark.parametrize("N", [32, 27])
def test_grf_larger_and_odd(api, N):
    L = 1000.0
    km = restate.k_mag_grid(N, L)
    km[0, 0, 0] = 1e-10
    pk = surrogate_pk(km)
    rs = np.random.RandomState(N)
    c1 = rs.normal(0, 1 / np.sqrt(2), size=(N, N, N)); c2 = rs.normal(0, 1 / np.sqrt(2), size=(N, N, N))
    for guard in (True, False):
        R = restate.hermitian_R(c1, c2, guard)
        dref, psiref = restate.grf_delta_psi(pk, R, L, N)
        d, psi = api.grf_delta_psi(pk, L, N, c1=c1, c2=c2, guard=guard, precision="f64")
        assert nrm(d, dref) <= 1e-11 and nrm(np.stack(psi), np.stack(psiref)) <= 1e-11
        d, psi = api.grf_delta_psi(pk, L, N, c1=c1, c2=c2, guard=guard, precision="f32")
        assert nrm(d, dref) <= 1e-5 and nrm(np.stack(psi), np.stack(psiref)) <= 1e-5
