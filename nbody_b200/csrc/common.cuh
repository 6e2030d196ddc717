// Shared internals of libnbody_b200.so (sm_100a only).  Not part of the C-ABI.
#pragma once
#include <cuda_runtime.h>
#include <cufft.h>
#include <stdint.h>
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
  NB_SCR_NSLOTS
};

struct nb_ctx {
  int device = 0;
  int num_sms = NB_NUM_SMS_FALLBACK;
  int64_t launches = 0;
  int p3m_ctas_per_sm = 0;   // nb_p3m_set_ctas_per_sm
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
