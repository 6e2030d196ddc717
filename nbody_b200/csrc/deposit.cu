// K1 — CIC mass deposit (reference FraP_kTilDelta.py:93-123).
//
// Index and weight math is float64 and written with explicit _rn intrinsics so that no FMA
// contraction can change a cell index or a weight relative to numpy/numba.
//
// cic_deposit_direct: one thread per particle, reductions (REDG) straight to L2, correct for
// any particle order.  The mesh has 8 cells per particle at the BASELINE configs, so every
// cell is touched about once: there is nothing to aggregate in shared memory, and the
// measured shared-memory atomic rate (2 cyc/lane) is below the global reduction rate
// (0.85-1.3 cyc/lane) — the kernel is bounded by the reduction-lane rate, see DESIGN.md.
#include <stdlib.h>

#include "common.cuh"

// z-pair reduction: the two cells (k0, k0+1) of one (i,j) row are adjacent in memory; when k0
// is even (8-byte aligned, no wrap) they go out as ONE vector reduction red.global.add.v2.f32
// (sm_90+), otherwise as two scalar ones.  The LSU/L2 reduction rate is per lane, so this
// removes a quarter of the reduction lanes on average.
__device__ __forceinline__ void red_pair(float* __restrict__ row, int k0, int k1, float a, float b) {
  if ((k0 & 1) == 0 && k1 == k0 + 1) {
    asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(row + k0), "f"(a), "f"(b) : "memory");
  } else {
    atomicAdd(row + k0, a);
    atomicAdd(row + k1, b);
  }
}
__device__ __forceinline__ void red_pair(double* __restrict__ row, int k0, int k1, double a, double b) {
  atomicAdd(row + k0, a);
  atomicAdd(row + k1, b);
}

template <typename T, bool VEC2>
__global__ void __launch_bounds__(256)
cic_deposit_direct(const double* __restrict__ pos, int64_t np, double cell, int N, int x0, int nxp,
                   T* __restrict__ grid, double shift, int32_t* __restrict__ dbg,
                   unsigned long long* __restrict__ n_outside, int lat_n, int lat_yb) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const double inv_cell = __ddiv_rn(1.0, cell);
  for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < np; q += stride) {
    const int64_t p = nb_blocked_order(q, lat_n, lat_yb);
    double x = pos[3 * p + 0], y = pos[3 * p + 1], z = pos[3 * p + 2];
    if (shift != 0.0) {
      x = __dadd_rn(x, shift);
      y = __dadd_rn(y, shift);
      z = __dadd_rn(z, shift);
    }
    // float64 grids keep numpy's divisions; float32 grids use the division-free form (same
    // indices bit for bit, weights within 1 ulp of float64 — see nb_cell_split)
    constexpr bool EXACT = sizeof(T) == 8;
    int i0, j0, k0;
    double dx, dy, dz;
    nb_cell_split<EXACT>(x, cell, inv_cell, N, i0, dx);
    nb_cell_split<EXACT>(y, cell, inv_cell, N, j0, dy);
    nb_cell_split<EXACT>(z, cell, inv_cell, N, k0, dz);
    if (dbg) {
      dbg[3 * p + 0] = i0;
      dbg[3 * p + 1] = j0;
      dbg[3 * p + 2] = k0;
    }
    const int j1 = nb_wrap1(j0, N), k1 = nb_wrap1(k0, N);
    // x planes relative to the first plane x0 of this grid ([nxp][N][N]; the full mesh has
    // x0 = 0, nxp = N; a slab has nxp = own planes + ghost planes on the +x side)
    int pa = i0 - x0, pb = nb_wrap1(i0, N) - x0;
    pa = pa < 0 ? pa + N : pa;
    pb = pb < 0 ? pb + N : pb;
    if (pa >= nxp || pb >= nxp) {   // not a particle of this slab: dropped and reported
      if (n_outside) atomicAdd(n_outside, 1ull);
      continue;
    }
    const double wx0 = __dsub_rn(1.0, dx), wy0 = __dsub_rn(1.0, dy), wz0 = __dsub_rn(1.0, dz);
    // weight = (wx * wy) * wz, left to right like the reference expression
    const double w00 = __dmul_rn(wx0, wy0), w01 = __dmul_rn(wx0, dy);
    const double w10 = __dmul_rn(dx, wy0), w11 = __dmul_rn(dx, dy);
    const int64_t r00 = ((int64_t)pa * N + j0) * N, r01 = ((int64_t)pa * N + j1) * N;
    const int64_t r10 = ((int64_t)pb * N + j0) * N, r11 = ((int64_t)pb * N + j1) * N;
    if constexpr (VEC2) {
      red_pair(grid + r00, k0, k1, (T)__dmul_rn(w00, wz0), (T)__dmul_rn(w00, dz));
      red_pair(grid + r01, k0, k1, (T)__dmul_rn(w01, wz0), (T)__dmul_rn(w01, dz));
      red_pair(grid + r10, k0, k1, (T)__dmul_rn(w10, wz0), (T)__dmul_rn(w10, dz));
      red_pair(grid + r11, k0, k1, (T)__dmul_rn(w11, wz0), (T)__dmul_rn(w11, dz));
    } else {
      atomicAdd(&grid[r00 + k0], (T)__dmul_rn(w00, wz0));
      atomicAdd(&grid[r00 + k1], (T)__dmul_rn(w00, dz));
      atomicAdd(&grid[r01 + k0], (T)__dmul_rn(w01, wz0));
      atomicAdd(&grid[r01 + k1], (T)__dmul_rn(w01, dz));
      atomicAdd(&grid[r10 + k0], (T)__dmul_rn(w10, wz0));
      atomicAdd(&grid[r10 + k1], (T)__dmul_rn(w10, dz));
      atomicAdd(&grid[r11 + k0], (T)__dmul_rn(w11, wz0));
      atomicAdd(&grid[r11 + k1], (T)__dmul_rn(w11, dz));
    }
  }
}

static int deposit_launch(nb_ctx* ctx, const double* pos, int64_t np, double cell, int N, int x0, int nxp,
                          void* grid, int prec, double shift, int accumulate, int32_t* dbg_index,
                          unsigned long long* n_outside, cudaStream_t st) {
  const size_t esz = prec == NB_F64 ? 8 : 4;
  if (!accumulate) NB_CHECK_CUDA(cudaMemsetAsync(grid, 0, esz * (size_t)nxp * N * N, st));
  if (n_outside) NB_CHECK_CUDA(cudaMemsetAsync(n_outside, 0, sizeof(unsigned long long), st));
  if (np == 0) return NB_OK;
  const int block = 256;
  const int nblk = nb_grid_for(ctx, np, block, 16);
  int lat_n = 0;
  const int lat_yb = nb_blocked_enabled() ? nb_lattice_yblock(np, &lat_n) : 0;
  static const int scalar_red = [] { const char* e = getenv("NB_DEPOSIT_SCALAR"); return e && atoi(e) == 1; }();
  if (prec == NB_F64)
    cic_deposit_direct<double, false><<<nblk, block, 0, st>>>(pos, np, cell, N, x0, nxp, (double*)grid, shift,
                                                              dbg_index, n_outside, lat_n, lat_yb);
  else if (scalar_red || (N & 1))   // odd N: rows are not 8-byte aligned
    cic_deposit_direct<float, false><<<nblk, block, 0, st>>>(pos, np, cell, N, x0, nxp, (float*)grid, shift,
                                                             dbg_index, n_outside, lat_n, lat_yb);
  else
    cic_deposit_direct<float, true><<<nblk, block, 0, st>>>(pos, np, cell, N, x0, nxp, (float*)grid, shift,
                                                            dbg_index, n_outside, lat_n, lat_yb);
  NB_LAUNCHED(ctx);
  return NB_OK;
}

extern "C" int nb_cic_deposit(nb_ctx* ctx, const double* pos, int64_t np, double cell, int N,
                              void* grid, int prec, double shift, int accumulate,
                              int32_t* dbg_index, void* stream) {
  NB_REQUIRE(ctx && grid, "ctx/grid is NULL");
  NB_REQUIRE(np >= 0 && (np == 0 || pos), "pos is NULL");
  NB_REQUIRE(N >= 1 && (int64_t)N * N * N < ((int64_t)1 << 40), "bad mesh size");
  NB_REQUIRE(cell > 0.0, "cell size must be positive");
  NB_REQUIRE(prec == NB_F32 || prec == NB_F64, "prec must be NB_F32 or NB_F64");
  NB_CHECK_CUDA(cudaSetDevice(ctx->device));
  return deposit_launch(ctx, pos, np, cell, N, 0, N, grid, prec, shift, accumulate, dbg_index, nullptr,
                        (cudaStream_t)stream);
}

// Slab form: grid is [nx_own + n_ghost][N][N] float32 and holds the mesh planes
// x0 .. x0 + nx_own + n_ghost - 1 (mod N); ghost planes collect what the slab's particles
// deposit beyond its last own plane (1 plane, 2 with the half-cell interlace shift).
extern "C" int nb_cic_deposit_slab(nb_ctx* ctx, const double* pos, int64_t np, double cell, int N, int x0,
                                   int nx_own, int n_ghost, float* grid, double shift, int accumulate,
                                   int64_t* n_outside, void* stream) {
  NB_REQUIRE(ctx && grid, "ctx/grid is NULL");
  NB_REQUIRE(np >= 0 && (np == 0 || pos), "pos is NULL");
  NB_REQUIRE(N >= 1 && cell > 0.0, "bad N / cell");
  NB_REQUIRE(x0 >= 0 && x0 < N && nx_own >= 1 && nx_own <= N && n_ghost >= 0, "bad slab");
  NB_CHECK_CUDA(cudaSetDevice(ctx->device));
  return deposit_launch(ctx, pos, np, cell, N, x0, nx_own + n_ghost, grid, NB_F32, shift, accumulate, nullptr,
                        (unsigned long long*)n_outside, (cudaStream_t)stream);
}
