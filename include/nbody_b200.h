/* nbody_b200 — C-ABI of the B200-native gravity / density hot path.
 *
 * The reference (HusseinYoud/Bachelor-Thesis-N-Body) has no FFI: its boundary is a set of
 * numpy-in / numpy-out Python functions.  This header is what those
 * functions bind through ctypes in nbody_b200/_lib.py; each entry cites the reference
 * code it replaces (paths relative to the reference root).  INTEGRATION.md shows the
 * binding a maintainer of the reference would add.
 *
 * Conventions
 *  - every pointer is a DEVICE pointer unless its name ends in _host;
 *  - `stream` is a cudaStream_t passed as void* (0 = legacy default stream); calls are
 *    asynchronous on it, the caller synchronises;
 *  - return 0 on success, a negative NB_E* code on failure; nb_last_error() gives the
 *    message of the last failure on the calling thread;
 *  - the caller owns all data buffers; the context owns cuFFT plans and scratch;
 *  - grids are C-order [ix][iy][iz] (z fastest) like the reference's numpy arrays;
 *    `prec` is 0 for float32 and 1 for float64 storage of a grid / spectrum;
 *  - positions are (Np,3) float64 AoS, exactly the snapshot layout
 *    (`components/matter/pos`, FraP_kTilDelta.py:55-60).
 *  - there is no CPU fallback anywhere behind this interface.
 */
#ifndef NBODY_B200_H
#define NBODY_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NB_ABI_VERSION 6

#define NB_OK 0
#define NB_EINVAL (-1)  /* bad argument */
#define NB_ECUDA (-2)   /* CUDA runtime error */
#define NB_ECUFFT (-3)  /* cuFFT error */
#define NB_ENOMEM (-4)  /* scratch allocation failed */

#define NB_F32 0
#define NB_F64 1

typedef struct nb_ctx nb_ctx;

/* ---- context ------------------------------------------------------------------------ */
int nb_abi_version(void);
/* Creates the context on CUDA device `device` (plans and scratch are created lazily). */
int nb_init(int device, nb_ctx** out);
void nb_destroy(nb_ctx* ctx);
const char* nb_last_error(void);
/* Frees cached scratch buffers and FFT plans (they are re-created on demand). */
int nb_release_scratch(nb_ctx* ctx);
/* Number of kernels this library has launched through `ctx` since nb_init (bench.py's
 * gpu_launches claim). */
int64_t nb_launch_count(nb_ctx* ctx);
/* Optional device-side phase timing of the composite calls (CUDA events on the caller's
 * stream around each phase).  nb_profile_read sums the recorded intervals per phase id
 * (NB_PH_*), returns them in HOST arrays and clears the records. */
#define NB_PH_DEPOSIT 0   /* K1 inside nb_pm_force_mesh (memset + kernel)        */
#define NB_PH_FFT_FWD 1   /* cuFFT R2C                                          */
#define NB_PH_KSPACE 2    /* K2                                                 */
#define NB_PH_FFT_INV 3   /* cuFFT C2R (all components)                         */
#define NB_PH_PK_BIN 4    /* K6 shell-binning kernel inside nb_power_spectrum   */
#define NB_PH_SORT 5      /* cell sort                                          */
#define NB_PH_COUNT 6
int nb_profile_enable(nb_ctx* ctx, int on);
int nb_profile_read(nb_ctx* ctx, double* ms_sum_host, int64_t* count_host, int nphases);
/* Measured FP32 roofline denominator for the pair kernels K4/K5 ("FP32 peak computed, not
 * measured"): a pure-FMA microkernel, packed = 0 scalar FFMA, 1 packed FFMA2 (fma.rn.f32x2).  Best of
 * `reps` launches of `iters` x 64 FMAs per thread on 2 x 1024 threads per SM, in TFLOP/s (2 flop per lane). */
int nb_fp32_peak(nb_ctx* ctx, int packed, int iters, int reps, double* tflops_host, void* stream);

/* ---- K1  CIC mass deposit -------------------------------------------------------------
 * Replaces cic_density(positions, boxsize, N)   FraP_kTilDelta.py:93-123,
 *          Power_compare.py:100-132, CiC/3D startbetingelser1.py:178-227 and
 *          cic_assign(delta, coords, dx, Nmesh)  Plotter.py:128-145.
 * i = int(p/cell) % N (truncation, Python mod), d = (p % cell)/cell, both in float64;
 * 8 weights (1-d | d) products added to grid[(i+di)%N,(j+dj)%N,(k+dk)%N].
 * cell      = boxsize/N as the caller computed it in float64 (cic_assign's `dx`)
 * shift     = added to every coordinate before the index math (0; half a cell for the
 *             second interlaced mesh of the PM step)
 * accumulate= 0: the grid is zeroed first (cic_density); 1: added in place (cic_assign)
 * dbg_index = NULL or (Np,3) int32: receives (i,j,k) per particle (parity tests)       */
int nb_cic_deposit(nb_ctx* ctx, const double* pos, int64_t np, double cell, int N,
                   void* grid, int prec, double shift, int accumulate,
                   int32_t* dbg_index, void* stream);

/* ---- K8  element-wise grid operations ------------------------------------------------ */
/* out_sum[0] = sum of all n cells in float64 (density.mean() numerator,
 * FraP_kTilDelta.py:126-127).  out_sum is a device double. */
int nb_grid_sum(nb_ctx* ctx, const void* grid, int64_t n, int prec, double* out_sum, void* stream);
/* grid <- (grid / a) / b - c, the three overdensity variants:
 *   compute_delta      FraP_kTilDelta.py:126-127   a = mean, b = 1, c = 1
 *   use_fac=False      Power_compare.py:153-157    a = cell_vol, b = Np/L^3, c = 1
 *   Plotter            Plotter.py:172-173          a = Np/Nmesh^3, b = 1, c = 1
 *   delta -= 1         Power_compare.py:159        a = 1, b = 1, c = 1
 * If a_dev is not NULL, a is read from that device double divided by a_div (mean =
 * sum / n without a host round trip). */
int nb_grid_affine(nb_ctx* ctx, void* grid, int64_t n, int prec, double a, const double* a_dev,
                   double a_div, double b, double c, void* stream);
/* max over i of |v_i|^2 for (n,3) float64 vectors (device scalar out; +inf if any component is NaN).  The
 * sub-cycled time loop derives its number of short-range sub-steps from it (CONCEPT picks rungs from the
 * short-range acceleration: the params_id0200_<A_s>.log files under TheFeedingPipeline/Datafolder4/
 * Test_Data_concept, lines 42 and 55). */
int nb_vec3_max_norm2(nb_ctx* ctx, const double* v, int64_t n, double* out_max, void* stream);

/* ---- a3  CIC window deconvolution ------------------------------------------------------
 * Replaces deconvolve_cic(delta_grid, boxsize, ngrid)  FraP_kTilDelta.py:87-91 and
 * Power_compare.py:93-97: grid <- ifftn(fftn(grid) / W).real,
 * W = ((s[i]*s[j])*s[k])^2 with s = np.sinc(k_vals/(2 k_N)) (N host-computed float64
 * values uploaded by the caller).  guard=1: W==0 -> 1 (FraP variant). */
int nb_deconvolve_cic(nb_ctx* ctx, void* grid, int N, int prec, const double* sinc_tab,
                      int guard, void* stream);

/* ---- K6  |delta_k|^2 shell binning ------------------------------------------------------
 * Replaces the tail of compute_power_spectrum_from_snapshot (Power_compare.py:166-200,
 * Plotter.py:176-202, 1vs2tlptplotter.py:96-116) and compute_power_spectrum
 * (Power_compare.py:422-465, CiC/3D startbetingelser1.py:274-315, Plotter.py:294-333):
 * R2C FFT of delta, P = |delta_k|^2 * norm / W^window_power, binned over k>0.
 * k2_tab[N]      = k_vals**2 (float64, host-computed: (2*pi*fftfreq)**2); k2_tab and sinc_tab are even in the
 *                  FFT index (tab[N-m] == tab[m], exact for fftfreq) and the complex64 kernel reads tab[0..N/2] only
 * thresh[n_edges]= for every bin edge e the smallest float64 x with sqrt(x) >= e, so that
 *                  "sqrt((kx²+ky²)+kz²) >= e" is decided exactly without a device sqrt
 * closed_last    = 1: np.histogram semantics (k == last edge counted in the last bin);
 *                  0: np.digitize semantics (half-open everywhere)        
 * thresh_first/_last = thresh[0], thresh[n_edges-1] as host scalars (they bracket the
 *                  float32 bin guess); thresh_hi = smallest x with sqrt(x) > edges[-1]
 *                  ("k == edges[-1]"  <=>  thresh_last <= k² < thresh_hi)
 * window_power   = 0 none, 1 P/=W (Plotter compute_power_spectrum), 2 P/=W^2
 * clip_mode      = 0 none, 1 W==0 -> 1, 2 W <- max(W, clip_val)
 * counts[n_edges-1] int64 and sums[n_edges-1] float64 are zeroed then filled.
 * delta is not modified.  Full-grid counts are reproduced from the half spectrum with
 * mode weights 2/1. */
int nb_power_spectrum(nb_ctx* ctx, const void* delta, int N, int prec, const double* k2_tab,
                      const double* sinc_tab, const double* thresh, double thresh_first,
                      double thresh_last, double thresh_hi, int n_edges, int window_power,
                      int clip_mode, double clip_val, double norm, int closed_last,
                      int64_t* counts, double* sums, void* stream);

/* CONCEPT-style estimator front end (options `powerspec_options['interlace']`,
 * `['deconvolve']` of SnapVsPK/Uncertain.param:29-62; outputs to match: PPPMP3M/{PP,PM,P3M}/powerspec/
 * powerspec_a=1.00, SnapVsPK/powerspec/powerspec_a={0.17,1.00}).  Same as nb_power_spectrum with a
 * second real grid `delta_shifted` = the CIC deposit on the mesh shifted by half a cell in x, y, z
 * (nb_cic_deposit with shift = cell/2): both are transformed and combined as
 * 0.5 (d1_k + d2_k e^{+i k.s}) before the |.|^2 / W^window_power binning (window_power = 2 is the
 * CIC deconvolution of the power).  The caller passes bin thresholds in k^2 (any increasing edges). */
int nb_power_spectrum_interlaced(nb_ctx* ctx, const void* delta, const void* delta_shifted, int N, int prec,
                                 const double* k2_tab, const double* sinc_tab, const double* thresh,
                                 double thresh_first, double thresh_last, double thresh_hi, int n_edges,
                                 int window_power, int clip_mode, double clip_val, double norm,
                                 int closed_last, int64_t* counts, double* sums, void* stream);

/* ---- K7  Gaussian random field from P(k) ------------------------------------------------
 * nb_hermitian_R replaces createR's symmetrisation loop (Cosmo data create (new).py:35-45;
 * guarded copies CiC/3D startbetingelser1.py:82-93, Power_compare.py:301-312):
 * R = c1 + i c2, R[0,0,0] real, then `R[-i,-j,-k] = conj(R[i,j,k])` in lexicographic order.
 * c1, c2: N^3 float64 draws; R_out: N^3 complex128 (interleaved re,im).
 * guard_selfconj = 1: `if (i,j,k) != (-i,-j,-k)` variant (self-conjugate points keep their
 * imaginary part); 0: createR (imaginary part negated). */
int nb_hermitian_R(nb_ctx* ctx, const double* c1, const double* c2, int N, int guard_selfconj,
                   double* R_out, void* stream);
/* nb_grf_synthesize replaces find_delta (Cosmo data create (new).py:114-137) and the δ / ψ
 * block of CiC/3D startbetingelser1.py:98-112 (= Power_compare.py:316-331):
 *   delta = ifftn(sqrt(P)·fac·R).real,  psi_a = ifftn(1j·k_a/k_mag²·sqrt(P)·fac·R).real
 * with k_mag[0,0,0] = 1e-10.  pk_grid: N^3 float64 P(k) on the FFT grid (the caller
 * evaluates its spectrum there; CLASS is not part of this library); k_vals: N float64;
 * fac: sqrt(2π/L³)·N³ (or 1, Plotter.py:100).  delta_out (N^3) and psi_out (3·N^3) are
 * float32/float64 by `prec`; either may be NULL. */
int nb_grf_synthesize(nb_ctx* ctx, const double* pk_grid, const double* c1, const double* c2, int N,
                      const double* k_vals, double fac, int guard_selfconj, int prec,
                      void* delta_out, void* psi_out, void* stream);
/* Second-order (2LPT) displacement psi2 = grad(phi2), laplace(phi2) = sum_{a>b} [phi1,aa phi1,bb - phi1,ab^2],
 * laplace(phi1) = delta, of the lattice ICs (`realization_options['LPT'] = 2`, Different_Sizes/Giant.param:70-72;
 * the reference looks at both orders in 1LPTvs2LPT/1vs2tlptplotter.py).  float64: delta (N,N,N) in, work6 = 6 N^3
 * doubles of scratch, psi2_out (3,N,N,N).  11 cuFFT transforms + the spectral-factor and source kernels. */
int nb_lpt2_displacement(nb_ctx* ctx, const double* delta, int N, const double* k_vals, double* work6,
                         double* psi2_out, void* stream);

/* ---- K3  CIC gather, displacement, kick / drift -----------------------------------------
 * nb_cic_gather: out[p,c] (=|+=) scale · Σ_8 w · field[c][cell]  with the deposit's float64
 * index/weight math; true periodic trilinear interpolation, equal to the reference's
 * RegularGridInterpolator where the reference evaluates it (lattice nodes,
 * Power_compare.py:356-385).  field: ncomp·N^3 (`prec`), out: (Np,ncomp) f64. */
int nb_cic_gather(nb_ctx* ctx, const void* field, int ncomp, int N, int prec, const double* pos,
                  int64_t np, double cell, double shift, double scale, int accumulate, double* out,
                  void* stream);
/* pos <- (pos + disp) % L with Python-mod semantics: `particles += ψ; particles %= L`
 * (Power_compare.py:386-391). */
int nb_displace_wrap(nb_ctx* ctx, double* pos, const double* disp, int64_t np, double L, void* stream);
/* mom += kick·acc ; pos <- (pos + drift·mom) % L   (kick = m·Δt_kick, drift = Δt_drift/(m a²);
 * no reference source: CONCEPT-internal leapfrog). */
int nb_kick_drift(nb_ctx* ctx, double* pos, double* mom, const double* acc, int64_t np, double kick,
                  double drift, double L, void* stream);
/* Fused PM tail: g = gscale · gather(force3 [, averaged with force3_shifted gathered at
 * pos + cell/2]); acc_out (nullable) <- g; if mom != NULL: kick + drift + wrap as above. */
int nb_gather_kick_drift(nb_ctx* ctx, const float* force3, const float* force3_shifted, int N,
                         double cell, double* pos, double* mom, double* acc_out, int64_t np,
                         double gscale, double kick, double drift, double L, void* stream);

/* As nb_gather_kick_drift(_slab) with g += acc_add[p] (nullable) before acc_out / kick / drift:
 * the tail of a P3M step whose short-range force was computed concurrently on another stream.
 * Full mesh: x0 = 0, nx_own = N, n_ghost = 0. */
int nb_gather_kick_drift_add(nb_ctx* ctx, const float* force3, const float* force3_shifted, int N, int x0,
                             int nx_own, int n_ghost, double cell, double* pos, double* mom,
                             const double* acc_add, double* acc_out, int64_t np, double gscale, double kick,
                             double drift, double L, void* stream);

/* ---- K2  PM / P3M long-range solve --------------------------------------------------------
 * No reference source (CONCEPT-internal; configured by Different_Sizes/Giant.param:28-68,
 * TheFeedingPipeline/NN.param:53-83).  Built from the reference's window
 * (FraP_kTilDelta.py:73-84) and its i·k/k² operator (CiC/3D startbetingelser1.py:105-107):
 *   g_k = pref · i k/k² · exp(-k² r_s²) / W^deconv_order · rho_k,  g_0 = 0, Nyquist
 *   gradient zeroed; rho2 != NULL: upstream interlacing 0.5(rho1 + rho2·e^{+ik·s});
 *   h3 != NULL: second set g·e^{-ik·s} for downstream interlacing.
 * rho1/rho2: [N][N][N/2+1] complex64; g3/h3: 3 consecutive spectra of that shape. */
int nb_kspace_fused(nb_ctx* ctx, const void* rho1, const void* rho2, int N, const double* k_vals,
                    const double* sinc_tab, double r_s, int deconv_order, double pref, void* g3,
                    void* h3, void* stream);
/* deposit -> R2C -> [shifted deposit -> R2C] -> nb_kspace_fused -> 3 (6) C2R.
 * force3 / force3_shifted: 3·N^3 float32 (unnormalised C2R: fold 1/N^3 into pref). */
int nb_pm_force_mesh(nb_ctx* ctx, const double* pos, int64_t np, double cell, int N,
                     const double* k_vals, const double* sinc_tab, double r_s, int deconv_order,
                     int interlace_up, int interlace_down, double pref, float* force3,
                     float* force3_shifted, void* stream);
/* nb_pm_force_mesh from meshes the caller has deposited already (nb_cic_deposit: float32, counts; rho_shifted = the
 * deposit with shift = cell/2, NULL without upstream interlacing).  Used by the host-buffer step, which deposits each
 * chunk of positions while the next one is still being uploaded.  The meshes are read, not changed. */
int nb_pm_force_mesh_rho(nb_ctx* ctx, const float* rho, const float* rho_shifted, int N, const double* k_vals,
                         const double* sinc_tab, double r_s, int deconv_order, int interlace_down, double pref,
                         float* force3, float* force3_shifted, void* stream);

/* ---- K4  PP direct sum -------------------------------------------------------------------
 * No reference source (CONCEPT PP; only its run time is in PPPMP3M/pppmp3mplotter.py:18-19).
 * nb_pp_pack: (Np,3) f64 -> float4 (x-ox, y-oy, z-oz, w); w = masses[p] or `mass`.
 * nb_pp_accel: acc[i] (=|+=) scale · Σ_j w_j d/(|d|²+eps²)^{3/2}, d = x_j - x_i (minimum
 * image if box > 0).  FP32 FMA + rsqrt in 512-particle j-tiles, FP64 across tiles.
 * variant: 0 default, 1 scalar FFMA, 2 packed FFMA2 (4 i/thread), 3 packed (8 i/thread). */
int nb_pp_pack(nb_ctx* ctx, const double* pos, int64_t np, double ox, double oy, double oz, double mass,
               const double* masses, void* out_xyzw, void* stream);
int nb_pp_accel(nb_ctx* ctx, const void* ipos_xyzw, int64_t ni, const void* jpos_xyzw, int64_t nj,
                double eps, double box, double scale, int variant, int accumulate, double* acc,
                void* stream);
/* Near-field fix of nb_pp_accel's float32 coordinates for i == j sets (pos is the (np,3) float64 array that
 * nb_pp_pack packed with origin (ox,oy,oz) and uniform weight mass_w; perm / cell_start from nb_cell_sort of
 * the same positions on a box of side L with ncs >= 3 cells of side >= r_c): for every pair with r < r_c
 * adds scale x (float64 pair term - the float32 term nb_pp_accel summed for that pair) to acc, for the
 * i-particles [i0, i0 + ni) of pos (acc has ni rows: an i-sharded rank passes its own range).  periodic = 1
 * for the minimum-image variant (box = L), 0 for open boundaries.  Restores max|dg|/rms|g| <= 1e-5 where a
 * single pair closer than the softening length would otherwise leave 4e-4 (128^3, L = 500). */
int nb_pp_near_fix(nb_ctx* ctx, const double* pos, int64_t np, int64_t i0, int64_t ni, const int32_t* perm,
                   const int32_t* cell_start,
                   int ncs, double L, double ox, double oy, double oz, double mass_w, double eps, double r_c,
                   int periodic, double scale, double* acc, void* stream);

/* ---- K4e  periodic PP with Ewald summation ----------------------------------
 * No reference source (CONCEPT PP gravity in a periodic box: softened nearest-image force plus a
 * tabulated Ewald correction; method of Hernquist, Bouchet & Suto 1991).
 * nb_ewald_table: C(d) = F_periodic(d) - d/|d|^3 on the (g+1)^3 nodes of the octant [0, box/2]^3,
 *   splitting parameter alpha (2/box is the usual choice), images |n| <= nreal, wave vectors
 *   |m| <= nfour.  table_f64: (g+1)^3 x 3 doubles and/or table_f4: (g+1)^3 float4 (x,y,z,0).
 * nb_pp_accel_ewald: nb_pp_accel with box > 0 plus the trilinear interpolation of table_f4. */
int nb_ewald_table(nb_ctx* ctx, double box, int g, double alpha, int nreal, int nfour, double* table_f64,
                   void* table_f4, void* stream);
int nb_pp_accel_ewald(nb_ctx* ctx, const void* ipos_xyzw, int64_t ni, const void* jpos_xyzw, int64_t nj,
                      double eps, double box, double scale, const void* table_f4, int g, int accumulate,
                      double* acc, void* stream);

/* ---- K5  P3M short range -------------------------------------------------------------------
 * No reference source (CONCEPT-internal; log lines 42-46, 89-95 show tiling + reordering).
 * nb_cell_sort: counting sort into ncs^3 cells of side L/ncs; sorted_xyzw float4
 * (cell-relative x,y,z, mass), perm[slot] = original index, cell_start[ncs^3+1].
 * nb_p3m_short: for the cells [cell_lo, cell_hi) (slab of a multi-GPU run, or all):
 *   acc[perm[i]] (=|+=) scale · Σ_{r<r_cut} w_j d/(r²+eps²)^{3/2}·(1 - u³Q(t)),
 *   u = r/(2 r_s), t = 2u²/u_max² - 1, Q given by ncoef = 11 monomial coefficients (host).
 * i_limit > 0: particles with perm >= i_limit are sources only (halo copies of another slab's
 * particles in a multi-GPU run; their acc rows are not touched); 0 = every particle.
 * Only the acc rows of i-particles that live in [cell_lo, cell_hi) are written (accumulate = 0) or added
 * to (accumulate = 1): with a partial or empty cell range the caller zeroes the other rows itself.
 * sorted_xyzw.w must be the uniform mass nb_cell_sort writes. */
int nb_cell_sort(nb_ctx* ctx, const double* pos, int64_t np, double L, int ncs, double mass,
                 void* sorted_xyzw, int32_t* perm, int32_t* cell_start, void* stream);
/* nb_cell_sort without allocations: no counterpart in the reference; added
 * for callers that cannot call cudaMalloc (CUDA graph capture, serving loops, harnesses that forbid it).  Same
 * arguments and outputs as nb_cell_sort (cell_start bit-identical; the order inside a cell is decided by atomics
 * in both).  The scratch (cell id per particle, per-cell counters) is a static device array of the library, 9.0 MiB
 * on every GPU that loads it: at most 2^21 = 2 097 152 particles and 2^18 = 262 144 cells (ncs <= 64); larger
 * sizes are rejected with NB_EINVAL.  Calls no cudaMalloc / cudaFree / allocating library routine.  All calls share
 * the array: calls that overlap on one GPU (two streams, two contexts) race. */
int nb_cell_sort_static(nb_ctx* ctx, const double* pos, int64_t np, double L, int ncs, double mass,
                        void* sorted_xyzw, int32_t* perm, int32_t* cell_start, void* stream);
int nb_p3m_short(nb_ctx* ctx, const void* sorted_xyzw, const int32_t* cell_start, const int32_t* perm,
                 int ncs, int64_t cell_lo, int64_t cell_hi, double cell_side, double r_cut, double r_s,
                 double eps, const double* coef_host, int ncoef, double u_max, double scale,
                 int accumulate, int64_t i_limit, double* acc, void* stream);
/* nb_p3m_short with the close pairs in float64: the tile coordinates of the pair kernel are float32 in a frame of up
 * to 4 cells, which a pair at a few softening lengths turns into 1e-4..1e-3 of the rms force.  Here the kernel does
 * not evaluate an in-range pair with 0 < r < r_near but lists it, and a second kernel evaluates the listed pairs
 * from pos ((np,3) float64, the array nb_cell_sort sorted; box = periodic box side) with erfc/exp in double.
 * near_count_out (device int32 or NULL): number of near pairs found; beyond the list capacity (2 np entries, at
 * least 2^20) pairs keep the float32 evaluation.  No reference source (CONCEPT short range, float64 throughout). */
int nb_p3m_short_exact(nb_ctx* ctx, const void* sorted_xyzw, const int32_t* cell_start, const int32_t* perm,
                       int ncs, int64_t cell_lo, int64_t cell_hi, double cell_side, double r_cut, double r_s,
                       double eps, const double* coef_host, int ncoef, double u_max, double scale, int accumulate,
                       int64_t i_limit, double* acc, const double* pos, int64_t np, double box, double r_near,
                       int32_t* near_count_out, void* stream);
/* Resident CTAs per SM of the nb_p3m_short grid (0 = default: many waves of CTAs).  A value
 * below the kernel's natural residency (4) leaves shared memory / registers on every SM for
 * kernels of another stream (the PM chain of the same step runs concurrently). */
int nb_p3m_set_ctas_per_sm(nb_ctx* ctx, int ctas_per_sm);
/* sm_limit > 0: the next nb_p3m_short calls run the pair kernel as a persistent grid confined to the SMs
 * [0, sm_limit) (CTAs on other SMs exit at once, work units come from an atomic counter), so that the
 * memory-bound mesh chain of the same step can run beside it on the remaining SMs (another stream).  0 = off. */
int nb_p3m_set_sm_limit(nb_ctx* ctx, int sm_limit);
/* Raises the stop flag of the confined pair kernel from `stream` (the mesh-chain stream, after its last kernel):
 * the confined CTAs stop fetching units and the unconfined second launch of nb_p3m_short finishes the list. */
int nb_p3m_mesh_done(nb_ctx* ctx, void* stream);
/* One thread spinning for about `us` microseconds on `stream` (head of the mesh-chain stream of the overlapped
 * P3M step: the persistent pair kernel launched beside it gets its CTAs resident first). */
int nb_delay_us(nb_ctx* ctx, int us, void* stream);
/* Pair statistics of the same traversal: out2[0] = candidate pairs (all j in the 27
 * neighbour cells of every i), out2[1] = pairs with 0 < r < r_cut (device int64[2]).
 * perm / i_limit as in nb_p3m_short (perm may be NULL when i_limit = 0). */
int nb_p3m_count_pairs(nb_ctx* ctx, const void* sorted_xyzw, const int32_t* cell_start,
                       const int32_t* perm, int64_t i_limit, int ncs, int64_t cell_lo, int64_t cell_hi,
                       double cell_side, double r_cut, int64_t* out2, void* stream);

/* ---- slab decomposition (multi-GPU) ---------------------------------------------
 * No reference counterpart: every CONCEPT run of the reference is one process ("Domain
 * decomposition: 1x1x1", run logs line 17).  Real mesh in x-slabs (rank r owns planes
 * [r·N/G, (r+1)·N/G)), spectrum in y-slabs of [N kx][N/G ky][N/2+1 kz] after the transpose.
 * The NCCL exchanges between these calls are issued by nbody_b200/dist.py.
 *
 * nb_cic_deposit_slab / nb_gather_kick_drift_slab: K1 / K3 on a slab buffer
 *   [nx_own + n_ghost][N][N] holding planes x0.. (mod N); ghost planes follow the own planes
 *   (+x side; 1 plane, 2 with the half-cell interlace shift).  Particles outside the slab are
 *   dropped and counted in n_outside (device int64, nullable) / gather zero.
 * nb_kspace_fused_slab, nb_pk_bin_slab: K2 / K6 on the y-slab ky in [y0, y0+ny_local);
 *   nb_pk_bin_slab ADDS into counts/sums (zero them first, all-reduce afterwards).
 * nb_slab_fft2d: batched 2-D R2C (dir=+1) / C2R (dir=-1) over (y,z) of nplanes x-planes.
 * nb_slab_fft1d_x: in-place 1-D C2C along x of spec[N][rows], dir = +1 forward / -1 inverse.
 * nb_slab_pack:   [nxl][G][nyl·nzh] -> [G][nxl][nyl·nzh] (send layout of the forward transpose)
 * nb_slab_unpack: [G][nxl][nyl·nzh] -> [nxl][G·nyl][nzh] (after the inverse transpose)
 * nb_slab_owner: owner[p] = (int(x/cell) % N) / planes_per_rank with the deposit's float64
 *   index math; counts (device int64[G], nullable) receives particles per owner.
 * nb_grid_add: dst += src (float32; received ghost planes onto the first own planes). */
int nb_cic_deposit_slab(nb_ctx* ctx, const double* pos, int64_t np, double cell, int N, int x0,
                        int nx_own, int n_ghost, float* grid, double shift, int accumulate,
                        int64_t* n_outside, void* stream);
int nb_gather_kick_drift_slab(nb_ctx* ctx, const float* force3, const float* force3_shifted, int N,
                              int x0, int nx_own, int n_ghost, double cell, double* pos, double* mom,
                              double* acc_out, int64_t np, double gscale, double kick, double drift,
                              double L, void* stream);
int nb_kspace_fused_slab(nb_ctx* ctx, const void* rho1, const void* rho2, int N, int y0, int ny_local,
                         const double* k_vals, const double* sinc_tab, double r_s, int deconv_order,
                         double pref, void* g3, void* h3, void* stream);
int nb_pk_bin_slab(nb_ctx* ctx, const void* spec, int N, int y0, int ny_local, const double* k2_tab,
                   const double* sinc_tab, const double* thresh, double thresh_first,
                   double thresh_last, double thresh_hi, int n_edges, int window_power, int clip_mode,
                   double clip_val, double norm, int closed_last, int64_t* counts, double* sums,
                   void* stream);
int nb_slab_fft2d(nb_ctx* ctx, int N, int nplanes, int dir, void* real, void* spec, void* stream);
int nb_slab_fft1d_x(nb_ctx* ctx, int N, int64_t rows, int dir, void* spec, void* stream);
int nb_slab_pack(nb_ctx* ctx, const void* src, void* dst, int nxl, int G, int nyl, int nzh, void* stream);
int nb_slab_unpack(nb_ctx* ctx, const void* src, void* dst, int nxl, int G, int nyl, int nzh, void* stream);
int nb_slab_owner(nb_ctx* ctx, const double* pos, int64_t np, double cell, int N, int planes_per_rank,
                  int32_t* owner, int64_t* counts, int G, void* stream);
/* Device-side particle exchange of the slab decomposition (no reference counterpart; replaces eager mask /
 * nonzero / cat code and its host round trips).
 * nb_halo_select: positions with x < lo_edge -> out_lo, x >= hi_edge -> out_hi (each (cap,3) float64; unused rows
 *   stay NaN, which nb_cell_sort skips, so neighbours can exchange the buffers at fixed size); counts2 (device
 *   int32[2]) = true numbers (> cap means overflow: rows beyond cap were dropped).
 * nb_migrate_split: particles with owner != rank (owner from nb_slab_owner) -> send_prev / send_next ((cap,6):
 *   pos, mom), their row indices -> holes (2 cap); counts4 (device int32) = {to_prev, to_next, holes, far} where
 *   far = particles that crossed more than one slab (not moved).
 * nb_migrate_merge: na + nb arrivals fill the first holes / are appended behind row np; with more leavers than
 *   arrivals the tail is compacted into the remaining holes.  work: 3 (n_holes - na - nb) + 2 int32 of device
 *   scratch in that case.  *new_np_host = np + na + nb - n_holes. */
int nb_halo_select(nb_ctx* ctx, const double* pos, int64_t np, double lo_edge, double hi_edge, double* out_lo,
                   double* out_hi, int cap, int32_t* counts2, void* stream);
int nb_migrate_split(nb_ctx* ctx, const double* pos, const double* mom, int64_t np, const int32_t* owner, int rank,
                     int prev, int next, double* send_prev, double* send_next, int cap, int32_t* holes,
                     int32_t* counts4, void* stream);
int nb_migrate_merge(nb_ctx* ctx, double* pos, double* mom, int64_t np, const int32_t* holes, int n_holes,
                     const double* recv_a, int na, const double* recv_b, int nb, int32_t* work,
                     int64_t* new_np_host, void* stream);
int nb_grid_add(nb_ctx* ctx, float* dst, const float* src, int64_t n, void* stream);

/* ---- batched snapshot -> delta -------------------------------------------
 * Replaces the per-file loop of Feeder_pipe.run_frap_to_delta (TheFeedingPipeline/
 * Feeder_pipe.py:163-195): for each of B snapshots  cic_density -> dens/dens.mean() - 1 ->
 * ifftn(fftn(.)/W).real  [-> delta -= delta.mean() where submean[b] != 0]  (FraP_kTilDelta.py:73-127).
 * pos: concatenated (total_particles, 3) float64 positions; offsets_dev: B+1 int64 (first particle
 * of each snapshot, offsets[B] = total); cell_dev: B float64 cell sizes boxsize_b / N;
 * sinc_tabs_dev: [B][N] float64 window tables of each snapshot's box; submean_dev: B int32, nullable;
 * grids: [B][N][N][N] float32 (prec = NB_F32) or float64 (NB_F64), overwritten with delta. */
int nb_snapshot_delta_batch(nb_ctx* ctx, const double* pos, const int64_t* offsets_dev,
                            int64_t total_particles, int B, const double* cell_dev, int N,
                            const double* sinc_tabs_dev, int guard, const int32_t* submean_dev,
                            void* grids, int prec, void* stream);
/* Batched snapshot P(k) (replaces `mp.Pool(10).starmap(compute_power_spectrum_from_snapshot, ...)`,
 * Different_Sizes/Plotter.py:205-238, and the per-file loops of SnapVsPK/Power_compare.py:211-235): B particle
 * sets that share the mesh size N are deposited, turned into overdensities grid / (Np_b / N^3) - 1, transformed by
 * one batched R2C and binned like nb_power_spectrum, per snapshot b with its own tables: cell_dev[b],
 * npart_dev[b] (device doubles), k2_tabs_dev / sinc_tabs_dev [B][N], thresh_dev [B][n_edges] (device),
 * thresh_first/_last/_hi_host[b] and norm_host[b] (HOST doubles).  counts [B][n_edges-1] int64 and sums
 * [B][n_edges-1] float64 (device) are zeroed then filled.  float32 grids. */
int nb_snapshot_pk_batch(nb_ctx* ctx, const double* pos, const int64_t* offsets_dev, int64_t total_particles, int B,
                         const double* cell_dev, const double* npart_dev, int N, const double* k2_tabs_dev,
                         const double* sinc_tabs_dev, const double* thresh_dev, const double* thresh_first_host,
                         const double* thresh_last_host, const double* thresh_hi_host, int n_edges, int window_power,
                         int clip_mode, double clip_val, const double* norm_host, int closed_last, int64_t* counts,
                         double* sums, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* NBODY_B200_H */
