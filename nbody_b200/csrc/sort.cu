// Cell sort for the P3M short-range force (CONCEPT's "Reordering matter particles in memory
// according to the gravity tiling", log lines 89-90 of the reference's run logs): counting
// sort of the particles into ncs³ cubic cells of side L/ncs >= r_cut.
//
// Output layout (device):
//   sorted_xyzw[slot] = (x - cx*cs, y - cy*cs, z - cz*cs, mass) in float32 — CELL-RELATIVE
//                       coordinates (computed in float64, then rounded), so pair separations
//                       keep ~1e-7·cs absolute accuracy independent of the box size;
//   perm[slot]        = original particle index;
//   cell_start[c]     = first slot of cell c = (cx*ncs + cy)*ncs + cz ; cell_start[ncs³] = number of
//                       valid particles (rows whose x is NaN are skipped: unused halo-buffer slots).
#include <cub/device/device_scan.cuh>

#include <mutex>

#include "common.cuh"

__device__ __forceinline__ int sort_cell_coord(double x, double L, double inv_cs, int ncs, double* rel, double cs) {
  double xw = x;
  if (xw < 0.0 || xw >= L) xw = nb_pymod(xw, L);
  int c = (int)(xw * inv_cs);
  if (c >= ncs) c = ncs - 1;
  if (c < 0) c = 0;
  *rel = xw - (double)c * cs;
  return c;
}

__global__ void __launch_bounds__(256)
sort_count_kernel(const double* __restrict__ pos, int64_t np, double L, double cs, double inv_cs, int ncs,
                  int* __restrict__ cell_of, int* __restrict__ counts) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < np; p += stride) {
    double r;
    if (!(pos[3 * p] == pos[3 * p])) {      // NaN row: an unused slot of a fixed-capacity halo buffer (nb_halo_select)
      cell_of[p] = -1;
      continue;
    }
    const int cx = sort_cell_coord(pos[3 * p], L, inv_cs, ncs, &r, cs);
    const int cy = sort_cell_coord(pos[3 * p + 1], L, inv_cs, ncs, &r, cs);
    const int cz = sort_cell_coord(pos[3 * p + 2], L, inv_cs, ncs, &r, cs);
    const int c = (cx * ncs + cy) * ncs + cz;
    cell_of[p] = c;
    atomicAdd(&counts[c], 1);
  }
}

// (A warp-aggregated variant — match.any on the cell id, one returning atomic per (warp, cell) group — measured
// 5.21 ms against 5.04 ms for the whole sort at C5: the kernel is bound by its scattered 16-byte stores, not by
// the atomics.)
__global__ void __launch_bounds__(256)
sort_scatter_kernel(const double* __restrict__ pos, int64_t np, double L, double cs, double inv_cs, int ncs,
                    float mass, const int* __restrict__ cell_of, const int* __restrict__ cell_start,
                    int* __restrict__ fill, float4* __restrict__ sorted, int* __restrict__ perm) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < np; p += stride) {
    const int c = cell_of[p];
    if (c < 0) continue;
    const int slot = cell_start[c] + atomicAdd(&fill[c], 1);
    double rx, ry, rz;
    sort_cell_coord(pos[3 * p], L, inv_cs, ncs, &rx, cs);
    sort_cell_coord(pos[3 * p + 1], L, inv_cs, ncs, &ry, cs);
    sort_cell_coord(pos[3 * p + 2], L, inv_cs, ncs, &rz, cs);
    sorted[slot] = make_float4((float)rx, (float)ry, (float)rz, mass);
    perm[slot] = (int)p;
  }
}

extern "C" int nb_cell_sort(nb_ctx* ctx, const double* pos, int64_t np, double L, int ncs, double mass,
                            void* sorted_xyzw, int32_t* perm, int32_t* cell_start, void* stream) {
  NB_REQUIRE(ctx && sorted_xyzw && perm && cell_start && (np == 0 || pos), "NULL argument");
  NB_REQUIRE(ncs >= 1 && (int64_t)ncs * ncs * ncs < ((int64_t)1 << 31) - 2, "bad ncs");
  NB_REQUIRE(np < ((int64_t)1 << 31), "at most 2^31-1 particles per device");
  NB_REQUIRE(L > 0, "L must be positive");
  cudaStream_t st = (cudaStream_t)stream;
  NB_CHECK_CUDA(cudaSetDevice(ctx->device));
  const int64_t nc = (int64_t)ncs * ncs * ncs;
  void *cell_of = nullptr, *counts = nullptr, *tmp = nullptr;
  int rc;
  if ((rc = nb_scratch(ctx, NB_SCR_SORT0, sizeof(int) * (size_t)(np > 0 ? np : 1), &cell_of))) return rc;
  if ((rc = nb_scratch(ctx, NB_SCR_SORT1, sizeof(int) * (size_t)(nc + 1), &counts))) return rc;
  size_t tmp_bytes = 0;
  cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, (int*)counts, cell_start, (int)(nc + 1), st);
  if ((rc = nb_scratch(ctx, NB_SCR_SORT2, tmp_bytes + 16, &tmp))) return rc;
  NB_CHECK_CUDA(cudaMemsetAsync(counts, 0, sizeof(int) * (size_t)(nc + 1), st));
  const double cs = L / ncs, inv_cs = (double)ncs / L;
  const int nblk = nb_grid_for(ctx, np > 0 ? np : 1, 256, 16);
  if (np > 0) {
    sort_count_kernel<<<nblk, 256, 0, st>>>(pos, np, L, cs, inv_cs, ncs, (int*)cell_of, (int*)counts);
    NB_LAUNCHED(ctx);
  }
  NB_CHECK_CUDA(cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, (int*)counts, cell_start, (int)(nc + 1), st));
  ctx->launches += 2;  // cub scan: upsweep + downsweep
  if (np > 0) {
    NB_CHECK_CUDA(cudaMemsetAsync(counts, 0, sizeof(int) * (size_t)(nc + 1), st));
    sort_scatter_kernel<<<nblk, 256, 0, st>>>(pos, np, L, cs, inv_cs, ncs, (float)mass, (const int*)cell_of,
                                             cell_start, (int*)counts, (float4*)sorted_xyzw, perm);
    NB_LAUNCHED(ctx);
  }
  return NB_OK;
}

// ---------------------------------------------------------------------------------------------------
// nb_cell_sort_static: the same counting sort for callers that must not allocate (CUDA graph capture, serving
// loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// Its scratch — the cell id of every particle and the per-cell counters — lives in a file-scope
// device array instead of nb_scratch's cudaMalloc'ed slots, and the prefix sum is a kernel of this file instead
// of cub's (no temporary storage to size).  Every GPU that loads the library holds the array until exit:
//   NB_SORT_STATIC_MAX_NP ints (cell ids) + NB_SORT_STATIC_MAX_NC + 1 ints (counters) = 9.0 MiB.
// All calls share the array and it keeps the last call's contents: the counters are cleared at the start of each
// call; two calls that overlap on the GPU (different streams, or different contexts on one device) race.
static constexpr int64_t NB_SORT_STATIC_MAX_NP = (int64_t)1 << 21;    // 2 097 152 particles (128^3)
static constexpr int64_t NB_SORT_STATIC_MAX_NC = (int64_t)1 << 18;    // 262 144 cells (64^3)
static __device__ __align__(256) int sort_static_scratch[NB_SORT_STATIC_MAX_NP + NB_SORT_STATIC_MAX_NC + 64];

// exclusive prefix sum of n ints by ONE CTA of 1024 threads (n <= 2^18 + 1: at most 257 rounds)
__global__ void __launch_bounds__(1024) sort_scan_one_cta_kernel(const int* __restrict__ in, int* __restrict__ out, int n) {
  __shared__ int warp_tot[32];
  __shared__ int s_base;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (threadIdx.x == 0) s_base = 0;
  __syncthreads();
  for (int t0 = 0; t0 < n; t0 += 1024) {
    const int t = t0 + threadIdx.x;
    const int v = t < n ? in[t] : 0;
    int incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int up = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += up;
    }
    if (lane == 31) warp_tot[w] = incl;
    __syncthreads();
    if (w == 0) {
      int x = warp_tot[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int up = __shfl_up_sync(0xffffffffu, x, o);
        if (lane >= o) x += up;
      }
      warp_tot[lane] = x;          // inclusive over the warps
    }
    __syncthreads();
    const int base = s_base + (w ? warp_tot[w - 1] : 0);
    if (t < n) out[t] = base + incl - v;
    __syncthreads();
    if (threadIdx.x == 0) s_base += warp_tot[31];
    __syncthreads();
  }
}

// device address of the array, looked up once per device with that device current (in host code the array's name
// is only a host-side stand-in for the symbol, not a pointer a kernel can use)
static int sort_static_address(int device, int** out) {
  constexpr int MAX_DEV = 64;
  static int* addr[MAX_DEV] = {};
  static std::mutex mu;
  NB_REQUIRE(device >= 0 && device < MAX_DEV, "device ordinal out of range");
  std::lock_guard<std::mutex> lock(mu);
  if (!addr[device]) {
    int* p = nullptr;
    NB_CHECK_CUDA(cudaGetSymbolAddress((void**)&p, sort_static_scratch));
    NB_REQUIRE(p != nullptr, "cudaGetSymbolAddress returned NULL");
    addr[device] = p;
  }
  *out = addr[device];
  return NB_OK;
}

extern "C" int nb_cell_sort_static(nb_ctx* ctx, const double* pos, int64_t np, double L, int ncs, double mass,
                                   void* sorted_xyzw, int32_t* perm, int32_t* cell_start, void* stream) {
  NB_REQUIRE(ctx && sorted_xyzw && perm && cell_start && (np == 0 || pos), "NULL argument");
  NB_REQUIRE(np >= 0 && ncs >= 1 && L > 0, "bad arguments");
  const int64_t nc = (int64_t)ncs * ncs * ncs;
  if (np > NB_SORT_STATIC_MAX_NP) {
    nb_set_error("nb_cell_sort_static: %lld particles exceed the static scratch (at most %lld); use nb_cell_sort",
                 (long long)np, (long long)NB_SORT_STATIC_MAX_NP);
    return NB_EINVAL;
  }
  if (nc > NB_SORT_STATIC_MAX_NC) {
    nb_set_error("nb_cell_sort_static: %d^3 cells exceed the static scratch (at most %lld cells); use nb_cell_sort",
                 ncs, (long long)NB_SORT_STATIC_MAX_NC);
    return NB_EINVAL;
  }
  cudaStream_t st = (cudaStream_t)stream;
  NB_CHECK_CUDA(cudaSetDevice(ctx->device));
  int* scratch = nullptr;
  int rc = sort_static_address(ctx->device, &scratch);
  if (rc) return rc;
  int* cell_of = scratch;
  int* counts = scratch + NB_SORT_STATIC_MAX_NP;
  NB_CHECK_CUDA(cudaMemsetAsync(counts, 0, sizeof(int) * (size_t)(nc + 1), st));
  const double cs = L / ncs, inv_cs = (double)ncs / L;
  const int nblk = nb_grid_for(ctx, np > 0 ? np : 1, 256, 16);
  if (np > 0) {
    sort_count_kernel<<<nblk, 256, 0, st>>>(pos, np, L, cs, inv_cs, ncs, cell_of, counts);
    NB_LAUNCHED(ctx);
  }
  sort_scan_one_cta_kernel<<<1, 1024, 0, st>>>(counts, cell_start, (int)(nc + 1));
  NB_LAUNCHED(ctx);
  if (np > 0) {
    NB_CHECK_CUDA(cudaMemsetAsync(counts, 0, sizeof(int) * (size_t)(nc + 1), st));
    sort_scatter_kernel<<<nblk, 256, 0, st>>>(pos, np, L, cs, inv_cs, ncs, (float)mass, cell_of, cell_start, counts,
                                             (float4*)sorted_xyzw, perm);
    NB_LAUNCHED(ctx);
  }
  return NB_OK;
}
