// K1 — CIC mass deposit (reference FraP_kTilDelta.py:93-123).
//
// Index and weight math is float64 and written with explicit _rn intrinsics so that no FMA
// contraction can change a cell index or a weight relative to numpy/numba.
//
// Two kernels:
//   cic_deposit_direct  one thread per particle, 8 reductions (RED.ADD) straight to L2.
//                       Correct for any particle order; used for small N and as fallback.
//   cic_deposit_tiled   (deposit_tiled.cu) shared-memory tile accumulation for spatially
//                       coherent particle streams.
#include "common.cuh"

template <typename T>
__global__ void __launch_bounds__(256)
cic_deposit_direct(const double* __restrict__ pos, int64_t np, double cell, int N,
                   T* __restrict__ grid, double shift, int32_t* __restrict__ dbg) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < np; p += stride) {
    double x = pos[3 * p + 0], y = pos[3 * p + 1], z = pos[3 * p + 2];
    if (shift != 0.0) {
      x = __dadd_rn(x, shift);
      y = __dadd_rn(y, shift);
      z = __dadd_rn(z, shift);
    }
    const int i0 = nb_cell_index(x, cell, N);
    const int j0 = nb_cell_index(y, cell, N);
    const int k0 = nb_cell_index(z, cell, N);
    const double dx = nb_cell_frac(x, cell);
    const double dy = nb_cell_frac(y, cell);
    const double dz = nb_cell_frac(z, cell);
    if (dbg) {
      dbg[3 * p + 0] = i0;
      dbg[3 * p + 1] = j0;
      dbg[3 * p + 2] = k0;
    }
    const int i1 = nb_wrap1(i0, N), j1 = nb_wrap1(j0, N), k1 = nb_wrap1(k0, N);
    const double wx0 = __dsub_rn(1.0, dx), wy0 = __dsub_rn(1.0, dy), wz0 = __dsub_rn(1.0, dz);
    // weight = (wx * wy) * wz, left to right like the reference expression
    const double w00 = __dmul_rn(wx0, wy0), w01 = __dmul_rn(wx0, dy);
    const double w10 = __dmul_rn(dx, wy0), w11 = __dmul_rn(dx, dy);
    const int64_t r00 = ((int64_t)i0 * N + j0) * N, r01 = ((int64_t)i0 * N + j1) * N;
    const int64_t r10 = ((int64_t)i1 * N + j0) * N, r11 = ((int64_t)i1 * N + j1) * N;
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

int nb_deposit_tiled_try(nb_ctx* ctx, const double* pos, int64_t np, double cell, int N, float* grid,
                         double shift, cudaStream_t st, int* used);

extern "C" int nb_cic_deposit(nb_ctx* ctx, const double* pos, int64_t np, double cell, int N,
                              void* grid, int prec, double shift, int accumulate,
                              int32_t* dbg_index, void* stream) {
  NB_REQUIRE(ctx && grid, "ctx/grid is NULL");
  NB_REQUIRE(np >= 0 && (np == 0 || pos), "pos is NULL");
  NB_REQUIRE(N >= 1 && (int64_t)N * N * N < ((int64_t)1 << 40), "bad mesh size");
  NB_REQUIRE(cell > 0.0, "cell size must be positive");
  NB_REQUIRE(prec == NB_F32 || prec == NB_F64, "prec must be NB_F32 or NB_F64");
  cudaStream_t st = (cudaStream_t)stream;
  NB_CHECK_CUDA(cudaSetDevice(ctx->device));
  const size_t esz = prec == NB_F64 ? 8 : 4;
  if (!accumulate) NB_CHECK_CUDA(cudaMemsetAsync(grid, 0, esz * (size_t)N * N * N, st));
  if (np == 0) return NB_OK;
  if (prec == NB_F32 && !dbg_index) {
    int used = 0;
    int rc = nb_deposit_tiled_try(ctx, pos, np, cell, N, (float*)grid, shift, st, &used);
    if (rc != NB_OK) return rc;
    if (used) return NB_OK;
  }
  const int block = 256;
  const int nblk = nb_grid_for(ctx, np, block, 16);
  if (prec == NB_F64)
    cic_deposit_direct<double><<<nblk, block, 0, st>>>(pos, np, cell, N, (double*)grid, shift, dbg_index);
  else
    cic_deposit_direct<float><<<nblk, block, 0, st>>>(pos, np, cell, N, (float*)grid, shift, dbg_index);
  NB_LAUNCHED(ctx);
  return NB_OK;
}
