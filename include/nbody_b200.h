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

#define NB_ABI_VERSION 1

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
 * k2_tab[N]      = k_vals**2 (float64, host-computed: (2*pi*fftfreq)**2)
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

#ifdef __cplusplus
}
#endif
#endif /* NBODY_B200_H */
