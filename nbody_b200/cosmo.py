"""Background cosmology, linear theory and the kick/drift integrals of the time loop
(the caller of the force step).

The reference runs its simulations with CONCEPT and takes the linear spectrum from CLASS
(`Giant.param:70-84`: H0, Ωb, Ωcdm, a_begin, primordial_spectrum A_s / n_s; `Cosmo data create
(new).py:52-110` for the CLASS call).  Neither is part of this package; what the time loop needs
from them is restated here in closed form:

* flat ΛCDM background without radiation, H(a) = H0 sqrt(Ωm a⁻³ + 1 − Ωm);
* growth factor D₊(a) = (5/2) Ωm H0² H(a) ∫₀ᵃ da' / (a' H(a'))³ (→ a in matter domination) and
  growth rate f = dlnD/dlna;
* the Eisenstein & Hu (1998) zero-baryon transfer function in place of CLASS;
* the comoving leapfrog integrals ∫dt/a (kick) and ∫dt/a² (drift) that multiply the reference's
  `mom += m·g·Δt ; pos += mom/(m a²)·Δt` update when positions are comoving and
  `mom = a² m dx/dt` (CONCEPT's convention, the `mom` dataset of its snapshots).

Units are CONCEPT's (Mpc, Gyr, 10¹⁰ M☉) as in params.py.  Pure host code: numpy only.
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import numpy as np

#: speed of light in Mpc / Gyr
C_LIGHT = 299792.458 * (1e3 / 3.0856775814913673e22) * (1e9 * 365.25 * 86400.0)
K_PIVOT = 0.05          # 1/Mpc, CLASS default pivot of the primordial spectrum
T_CMB = 2.7255

_GL_CACHE: dict = {}


def _gauss_legendre(n):
    if n not in _GL_CACHE:
        _GL_CACHE[n] = np.polynomial.legendre.leggauss(n)
    return _GL_CACHE[n]


def _integrate(fn, lo, hi, n=48):
    x, w = _gauss_legendre(n)
    half, mid = 0.5 * (hi - lo), 0.5 * (hi + lo)
    return half * float(np.sum(w * fn(mid + half * x)))


@dataclass
class Cosmology:
    """Flat ΛCDM background.  H0 in 1/Gyr (67.36 km/s/Mpc = 0.068890 / Gyr, the `H0` attribute of
    the reference's snapshots)."""
    H0: float = 0.068889891437478
    Omega_b: float = 0.049301692328524445
    Omega_cdm: float = 0.26447041034523616

    @property
    def Omega_m(self):
        return self.Omega_b + self.Omega_cdm

    @property
    def h(self):
        return self.H0 / 0.10227121650537077     # 100 km/s/Mpc in 1/Gyr

    def H(self, a):
        a = np.asarray(a, dtype=np.float64)
        return self.H0 * np.sqrt(self.Omega_m / a ** 3 + (1.0 - self.Omega_m))

    # ---- leapfrog integrals -------------------------------------------------------------
    def kick_factor(self, a0, a1):
        """∫ dt / a = ∫ da / (a² H) between a0 and a1."""
        return _integrate(lambda u: np.exp(-u) / self.H(np.exp(u)), math.log(a0), math.log(a1))

    def drift_factor(self, a0, a1):
        """∫ dt / a² = ∫ da / (a³ H) between a0 and a1."""
        return _integrate(lambda u: np.exp(-2.0 * u) / self.H(np.exp(u)), math.log(a0), math.log(a1))

    def cosmic_time(self, a):
        """t(a) = ∫₀ᵃ da / (a H) in Gyr."""
        return _integrate(lambda s: 2.0 * s / (s * s * self.H(s * s)), 0.0, math.sqrt(a), n=96)

    # ---- linear growth ------------------------------------------------------------------
    def _growth_integral(self, a):
        # substitution a' = s²: the integrand ~ a'^(3/2) becomes polynomial near 0
        return _integrate(lambda s: 2.0 * s / (s * s * self.H(s * s)) ** 3, 0.0, math.sqrt(a), n=96)

    def growth(self, a):
        """D₊(a), normalised to a in matter domination."""
        return 2.5 * self.Omega_m * self.H0 ** 2 * float(self.H(a)) * self._growth_integral(a)

    def growth_rate(self, a):
        """f = dlnD₊/dlna."""
        Ha = float(self.H(a))
        dlnH = -1.5 * self.Omega_m / a ** 3 * self.H0 ** 2 / Ha ** 2
        return dlnH + a / ((a * Ha) ** 3 * self._growth_integral(a))

    # ---- linear matter spectrum ---------------------------------------------------------
    def transfer(self, k):
        """Eisenstein & Hu (1998) zero-baryon ('no wiggle') transfer function, k in 1/Mpc."""
        k = np.asarray(k, dtype=np.float64)
        h = self.h
        om, ob = self.Omega_m * h * h, self.Omega_b * h * h
        fb = ob / om
        theta = T_CMB / 2.7
        s = 44.5 * math.log(9.83 / om) / math.sqrt(1.0 + 10.0 * ob ** 0.75)            # Mpc
        alpha = 1.0 - 0.328 * math.log(431.0 * om) * fb + 0.38 * math.log(22.3 * om) * fb * fb
        gamma = self.Omega_m * h * (alpha + (1.0 - alpha) / (1.0 + (0.43 * k * s) ** 4))
        q = (k / h) * theta * theta / gamma
        L0 = np.log(2.0 * math.e + 1.8 * q)
        C0 = 14.2 + 731.0 / (1.0 + 62.5 * q)
        return L0 / (L0 + C0 * q * q)

    def transfer_eh98(self, k):
        """Eisenstein & Hu (1998) transfer function WITH baryons (acoustic oscillations and Silk
        damping; their eqs. 2-24, the `TFfit_onek` formulae), k in 1/Mpc.  Against the `P_linear`
        column of the reference's powerspec files (CLASS, PPPMP3M/P3M/powerspec/powerspec_a=1.00:11-26)
        the resulting spectrum is within 3 % per bin (tests/test_concept_powerspec.py)."""
        k = np.asarray(k, dtype=np.float64)
        h = self.h
        omhh, obhh = self.Omega_m * h * h, self.Omega_b * h * h
        fb = obhh / omhh
        th = T_CMB / 2.7
        z_eq = 2.50e4 * omhh / th ** 4
        k_eq = 0.0746 * omhh / th ** 2
        b1 = 0.313 * omhh ** -0.419 * (1.0 + 0.607 * omhh ** 0.674)
        b2 = 0.238 * omhh ** 0.223
        z_d = 1291.0 * omhh ** 0.251 / (1.0 + 0.659 * omhh ** 0.828) * (1.0 + b1 * obhh ** b2)
        R_d = 31.5 * obhh / th ** 4 * (1000.0 / (1.0 + z_d))
        R_eq = 31.5 * obhh / th ** 4 * (1000.0 / z_eq)
        s = 2.0 / 3.0 / k_eq * math.sqrt(6.0 / R_eq) * math.log(
            (math.sqrt(1.0 + R_d) + math.sqrt(R_d + R_eq)) / (1.0 + math.sqrt(R_eq)))
        k_silk = 1.6 * obhh ** 0.52 * omhh ** 0.73 * (1.0 + (10.4 * omhh) ** -0.95)
        a1 = (46.9 * omhh) ** 0.670 * (1.0 + (32.1 * omhh) ** -0.532)
        a2 = (12.0 * omhh) ** 0.424 * (1.0 + (45.0 * omhh) ** -0.582)
        alpha_c = a1 ** -fb * a2 ** -(fb ** 3)
        bb1 = 0.944 / (1.0 + (458.0 * omhh) ** -0.708)
        bb2 = (0.395 * omhh) ** -0.0266
        beta_c = 1.0 / (1.0 + bb1 * ((1.0 - fb) ** bb2 - 1.0))
        y = z_eq / (1.0 + z_d)
        G = y * (-6.0 * math.sqrt(1.0 + y) + (2.0 + 3.0 * y) * math.log(
            (math.sqrt(1.0 + y) + 1.0) / (math.sqrt(1.0 + y) - 1.0)))
        alpha_b = 2.07 * k_eq * s * (1.0 + R_d) ** -0.75 * G
        beta_node = 8.41 * omhh ** 0.435
        beta_b = 0.5 + fb + (3.0 - 2.0 * fb) * math.sqrt((17.2 * omhh) ** 2 + 1.0)
        q = k / (13.41 * k_eq)
        xx = k * s
        ln_b = np.log(math.e + 1.8 * beta_c * q)
        ln_nb = np.log(math.e + 1.8 * q)
        C_na = 14.2 + 386.0 / (1.0 + 69.9 * q ** 1.08)
        C_a = 14.2 / alpha_c + 386.0 / (1.0 + 69.9 * q ** 1.08)
        f = 1.0 / (1.0 + (xx / 5.4) ** 4)
        T_c = f * ln_b / (ln_b + C_na * q * q) + (1.0 - f) * ln_b / (ln_b + C_a * q * q)
        s_t = s * (1.0 + (beta_node / xx) ** 3) ** (-1.0 / 3.0)
        T0 = ln_nb / (ln_nb + C_na * q * q)
        T_b = np.sinc(k * s_t / math.pi) * (T0 / (1.0 + (xx / 5.2) ** 2)
                                            + alpha_b / (1.0 + (beta_b / xx) ** 3) * np.exp(-(k / k_silk) ** 1.4))
        return fb * T_b + (1.0 - fb) * T_c

    def linear_power(self, k, A_s=2.1e-9, n_s=0.9649, a=1.0, wiggles=True):
        """P(k, a) in Mpc³ (convention P = |δ_k|² / V): primordial curvature spectrum
        2π² A_s k⁻³ (k/k_pivot)^(n_s−1) times [(2/5) (k c/H0)² T(k) D₊(a) / Ωm]².  T is the
        Eisenstein-Hu form with baryons (wiggles=True) or their zero-baryon fit."""
        k = np.asarray(k, dtype=np.float64)
        out = np.zeros_like(k)
        ok = k > 0
        kk = k[ok]
        prim = 2.0 * math.pi ** 2 * A_s / kk ** 3 * (kk / K_PIVOT) ** (n_s - 1.0)
        tk = self.transfer_eh98(kk) if wiggles else self.transfer(kk)
        amp = 0.4 * (kk * C_LIGHT / self.H0) ** 2 * tk * self.growth(a) / self.Omega_m
        out[ok] = prim * amp * amp
        return out

    def sigma_R(self, R=None, **kw):
        """rms linear fluctuation in tophat spheres of radius R (default 8 Mpc/h)."""
        R = 8.0 / self.h if R is None else R
        lnk = np.linspace(math.log(1e-5), math.log(50.0), 4000)
        k = np.exp(lnk)
        x = k * R
        W = 3.0 * (np.sin(x) - x * np.cos(x)) / x ** 3
        integrand = k ** 3 * self.linear_power(k, **kw) * W * W / (2.0 * math.pi ** 2)
        return math.sqrt(float(np.sum(0.5 * (integrand[1:] + integrand[:-1]) * np.diff(lnk))))


def scale_factor_edges(a_begin, a_end, dlna_max=0.025):
    """Step edges uniform in ln a with Δln a ≤ dlna_max (CONCEPT limits Δt to a fraction of the
    Hubble time, which is a bound on Δln a; its particle-speed and rung criteria are not restated)."""
    if not 0.0 < a_begin < a_end:
        raise ValueError("need 0 < a_begin < a_end")
    n = max(1, int(math.ceil(math.log(a_end / a_begin) / dlna_max - 1e-12)))
    return np.exp(np.linspace(math.log(a_begin), math.log(a_end), n + 1))
