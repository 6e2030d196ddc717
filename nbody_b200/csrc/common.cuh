// Shared internals of libnbody_b200.so (sm_100a only).  Not part of the C-ABI.
#pragma once
#include <cuda_runtime.h>
#include <cufft.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <stdio.h>
#include <string.h>
#include <map>
#include <tuple>
#include <vector>

#include "../../include/nbody_b200.h"

#define NB_NUM_SMS_FALLBACK 148

void nb_set_error(const char* fmt, ...);

#define NB_CHECK_CUDA(expr)                                                              \
  do {                                                                                   \
    cudaError_t _e = (expr);                                                             \
    if (_e != cudaSuccess) {                                                             \
      nb_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
      return NB_ECUDA;                                                                   \
    }                                                                                    \
  } while (0)

#define NB_CHECK_CUFFT(expr)                                                   \
  do {                                                                         \
    cufftResult _r = (expr);                                                   \
    if (_r != CUFFT_SUCCESS) {                                                 \
      nb_set_error("%s:%d: %s -> cufft error %d", __FILE__, __LINE__, #expr, (int)_r); \
      return NB_ECUFFT;                                                        \
    }                                                                          \
  } while (0)

#define NB_REQUIRE(cond, msg)                                        \
  do {                                                               \
    if (!(cond)) {                                                   \
      nb_set_error("%s:%d: invalid argument: %s", __FILE__, __LINE__, msg); \
      return NB_EINVAL;                                              \
    }                                                                \
  } while (0)

// after a kernel launch: catch launch-configuration errors and count the launch
#define NB_LAUNCHED(ctx)                 \
  do {                                   \
    NB_CHECK_CUDA(cudaGetLastError());   \
    (ctx)->launches++;                   \
  } while (0)

enum NbScratchSlot {
  NB_SCR_SPEC0 = 0,   // complex spectrum
  NB_SCR_SPEC1,
  NB_SCR_SPEC2,
  NB_SCR_SPEC3,
  NB_SCR_REAL0,       // real work grids
  NB_SCR_REAL1,
  NB_SCR_MISC0,       // small reductions / counters
  NB_SCR_MISC1,
  NB_SCR_SORT0,
  NB_SCR_SORT1,
  NB_SCR_SORT2,
  NB_SCR_SORT3,
  NB_SCR_NEAR,        // near-pair list of nb_p3m_short_exact
  NB_SCR_FLAGS,       // a few ints that outlive one call (unit counter / stop flag of the overlapped P3M step)
  NB_SCR_NSLOTS
};

struct nb_ctx {
  int device = 0;
  int num_sms = NB_NUM_SMS_FALLBACK;
  int64_t launches = 0;
  int p3m_ctas_per_sm = 0;   // nb_p3m_set_ctas_per_sm
  int p3m_sm_limit = 0;      // nb_p3m_set_sm_limit: pair kernel confined to the first SMs (overlapped step)
  void* scratch[NB_SCR_NSLOTS] = {};
  size_t scratch_bytes[NB_SCR_NSLOTS] = {};
  // key: (cufftType, n0, n1, n2, batch, layout) ; layout 0 = plain 3-D
  std::map<std::tuple<int, int, int, int, int, int>, cufftHandle> plans;
  // optional per-phase device timing (nb_profile_enable): (phase, start, end) event pairs
  bool prof_on = false;
  std::vector<std::tuple<int, cudaEvent_t, cudaEvent_t>> prof_events;
};

// RAII phase marker: records a CUDA event pair on `st` around a phase of a composite call
// (deposit / FFT / k-space ... inside nb_pm_force_mesh) when profiling is enabled.
struct NbPhase {
  nb_ctx* ctx;
  cudaStream_t st;
  cudaEvent_t e1 = nullptr;
  NbPhase(nb_ctx* c, int phase, cudaStream_t s) : ctx(c), st(s) {
    if (!ctx->prof_on) return;
    cudaEvent_t e0;
    if (cudaEventCreate(&e0) != cudaSuccess || cudaEventCreate(&e1) != cudaSuccess) { e1 = nullptr; return; }
    cudaEventRecord(e0, st);
    ctx->prof_events.emplace_back(phase, e0, e1);
  }
  ~NbPhase() {
    if (e1) cudaEventRecord(e1, st);
  }
};

// grow-only scratch
int nb_scratch(nb_ctx* ctx, int slot, size_t bytes, void** out);
// cached 3-D plan (type = CUFFT_R2C / C2R / D2Z / Z2D / C2C / Z2Z), bound to `stream`
int nb_plan3d(nb_ctx* ctx, int N, cufftType type, cudaStream_t stream, cufftHandle* out);
// cached batched plan: rank-`rank` transform of dims n[] over `batch` contiguous items
int nb_plan_many(nb_ctx* ctx, int rank, const int* n, int batch, cufftType type, int layout_tag,
                 const int* inembed, int istride, int idist, const int* onembed, int ostride,
                 int odist, cudaStream_t stream, cufftHandle* out);

// C2R / Z2D into `dst`; if dst is not 16-byte aligned (odd N, component c of a packed
// 3·N^3 array) the transform goes through scratch and is copied (cuFFT needs aligned I/O).
int nb_exec_c2r(nb_ctx* ctx, cufftHandle plan, int prec, void* spec, void* dst, size_t real_bytes,
                cudaStream_t st);

// Traversal order of the particle kernels (deposit, gather).  Snapshots and the Zel'dovich /
// glass initial conditions are in lattice order (ix, iy, iz; iz fastest), so the particles that
// share mesh sectors with particle p sit 1, n and n² entries away; the n² neighbours are ~100 MB
// of mesh footprint apart in a linear sweep and their shared sectors fall out of L2 in between.
// When Np = n³ the kernels therefore sweep y-blocks of NB_YBLOCK lattice rows (for each block:
// all ix, rows of the block, all iz), which brings the x-neighbours within n·NB_YBLOCK entries.
// It is only a processing ORDER: every particle is still visited exactly once and inputs that
// are not lattice ordered are neither helped nor harmed.  Returns the block height (0 = linear).
// NB_LINEAR_ORDER=1 switches the blocked sweep off, NB_YBLOCK=<rows> overrides its height (A/B runs)
static inline bool nb_blocked_enabled() {
  static const bool on = [] { const char* e = getenv("NB_LINEAR_ORDER"); return !(e && atoi(e) == 1); }();
  return on;
}
static inline int nb_lattice_yblock(int64_t np, int* n_out) {
  *n_out = 0;
  if (np < 16 * 16 * 16 || np >= ((int64_t)1 << 31)) return 0;
  int64_t n = (int64_t)(cbrt((double)np) + 0.5);
  if (n * n * n != np) return 0;
  *n_out = (int)n;
  static const int forced = [] { const char* e = getenv("NB_YBLOCK"); return e ? atoi(e) : 0; }();
  if (forced > 0 && n % forced == 0) return forced;
  if (n % 32 == 0 && n >= 64) return 32;
  if (n % 16 == 0 && n >= 32) return 16;
  if (n % 8 == 0) return 8;      // small sets: no benefit, but keeps the path under the parity tests
  return 0;
}

static inline int nb_grid_for(const nb_ctx* ctx, int64_t work_items, int block, int max_waves = 8) {
  int64_t want = (work_items + block - 1) / block;
  int64_t cap = (int64_t)ctx->num_sms * max_waves;
  if (want < 1) want = 1;
  return (int)(want < cap ? want : cap);
}

// ---------------------------------------------------------------------------------------
// device helpers: float64 index math that must match numpy/numba bit for bit
// ---------------------------------------------------------------------------------------
#ifdef __CUDACC__
// q-th particle of the y-blocked sweep (see nb_lattice_yblock); yb = 0 -> linear order
__device__ __forceinline__ int64_t nb_blocked_order(int64_t q, int n, int yb) {
  if (yb == 0) return q;
  const unsigned per_x = (unsigned)(yb * n), per_block = per_x * (unsigned)n;
  const unsigned uq = (unsigned)q;
  const unsigned b = uq / per_block, r = uq - b * per_block;
  const unsigned ix = r / per_x, r2 = r - ix * per_x;
  const unsigned iyl = r2 / (unsigned)n, iz = r2 - iyl * (unsigned)n;
  return ((int64_t)ix * n + (b * (unsigned)yb + iyl)) * n + iz;
}

// Python/numpy float modulo: result carries the sign of the divisor.
__device__ __forceinline__ double nb_pymod(double a, double b) {
  double r = fmod(a, b);
  if (r != 0.0) {
    if ((b < 0.0) != (r < 0.0)) r += b;
  } else {
    r = copysign(0.0, b);
  }
  return r;
}

// i = int(p/cell) % N : IEEE division, truncation toward zero, Python integer mod.
__device__ __forceinline__ int nb_cell_index(double p, double cell, int N) {
  double q = __ddiv_rn(p, cell);
  long long t = (long long)q;  // cvt.rzi
  long long m = t % (long long)N;
  if (m < 0) m += N;
  return (int)m;
}

// d = (p % cell)/cell
__device__ __forceinline__ double nb_cell_frac(double p, double cell) {
  return __ddiv_rn(nb_pymod(p, cell), cell);
}

__device__ __forceinline__ int nb_wrap1(int i, int N) { return i + 1 == N ? 0 : i + 1; }

// Cell index and in-cell fraction of one coordinate.
//   EXACT = true : idx = int(p/cell) % N and frac = (p % cell)/cell exactly as numpy evaluates
//                  them (two IEEE divisions and an fmod per coordinate; float64-grid API).
//   EXACT = false: the SAME idx, bit for bit, without the divisions: q~ = p * (1/cell) is within
//                  1.5 * 2^-52 * Q of both the exact quotient Q and the rounded one RN(Q) that
//                  numpy truncates; for 0 <= Q < 2^20 that is < 3.5e-10, so whenever q~ is
//                  further than 2^-29 from an integer, trunc(q~) == trunc(RN(Q)) == floor(Q).
//                  Anything else (negative p, huge p, q~ near an integer) takes the exact path.
//                  p % cell = p - floor(Q) * cell is exactly representable, so one FMA returns
//                  it exactly; frac = (p % cell) * (1/cell) is within 1 ulp of the reference's
//                  division (weights carry a 1e-5 tolerance, indices none).
template <bool EXACT>
__device__ __forceinline__ void nb_cell_split(double p, double cell, double inv_cell, int N, int& idx,
                                              double& frac) {
  if constexpr (!EXACT) {
    const double MAGIC = 6755399441055744.0;   // 1.5 * 2^52: (q + MAGIC) - MAGIC = rint(q)
    const double qt = __dmul_rn(p, inv_cell);
    const double a = __dadd_rn(qt, MAGIC);
    const double r = __dsub_rn(a, MAGIC);
    const double f = __dsub_rn(qt, r);         // exact, in [-0.5, 0.5]
    if (p >= 0.0 && qt < 1048576.0 && fabs(f) > 1.862645149230957e-09) {
      const bool below = f < 0.0;
      const int t = __double2loint(a) - (below ? 1 : 0);   // low word of q + MAGIC = rint(q)
      const double td = below ? __dsub_rn(r, 1.0) : r;
      frac = __dmul_rn(__fma_rn(-td, cell, p), inv_cell);
      idx = t >= N ? t % N : t;
      return;
    }
  }
  idx = nb_cell_index(p, cell, N);
  frac = nb_cell_frac(p, cell);
}

// pymod(v, L) for L > 0 without fmod when v is within one box length of [0, L): v - L is exact
// for L <= v < 2L (Sterbenz), and for -L <= v < 0 Python returns RN(fmod(v, L) + L) = RN(v + L).
__device__ __forceinline__ double nb_pymod_near(double v, double L) {
  if (v >= 0.0 && v < L) return v;
  if (v >= L && v < __dadd_rn(L, L)) return __dsub_rn(v, L);
  if (v < 0.0 && v >= -L) return __dadd_rn(v, L);
  return nb_pymod(v, L);
}

__device__ __forceinline__ float nb_warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double nb_warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
#endif
