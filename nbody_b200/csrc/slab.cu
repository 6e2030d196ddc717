// Slab-decomposed mesh support: the pieces of the distributed FFT that run on
// one GPU between the NCCL exchanges, plus particle-ownership and ghost-plane helpers.
//
//   real mesh   x-slabs:  rank r owns planes [r·nxl, (r+1)·nxl) of [N][N][N], nxl = N/G
//   spectrum    y-slabs:  rank r owns ky in [r·nyl, (r+1)·nyl) of [N kx][nyl][N/2+1]
//
//   forward:  2-D R2C over (y,z) of each own x-plane        nb_slab_fft2d(dir=+1)
//             pack   [nxl][G][nyl][nzh] -> [G][nxl][nyl][nzh] nb_slab_pack
//             all-to-all (NCCL, host side: dist.py)          recv = [N][nyl][nzh]
//             1-D C2C along x, stride nyl·nzh                nb_slab_fft1d_x(dir=+1)
//   inverse:  1-D C2C inverse along x; the result IS the send layout [G][nxl][nyl][nzh]
//             all-to-all; unpack [G][nxl][nyl][nzh] -> [nxl][N][nzh]; 2-D C2R.
// The reference has no counterpart (its only FFTs are single-process np.fft calls, e.g.
// FraP_kTilDelta.py:88-90); cuFFTMp is not in the image, so the transposes are ours.
#include "common.cuh"

// dst[a][b][row] = src[b][a][row]  for a < na, b < nb, rows of `row` float2 elements.
// One warp copies one row segment with 16-byte accesses when the row length is even.
__global__ void __launch_bounds__(256)
slab_swap01_kernel(const float2* __restrict__ src, float2* __restrict__ dst, int na, int nb, int64_t row) {
  const int64_t nrows = (int64_t)na * nb;
  const int64_t total = nrows * row;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  if ((row & 1) == 0) {
    const int64_t row2 = row >> 1, total2 = total >> 1;
    const float4* s4 = reinterpret_cast<const float4*>(src);
    float4* d4 = reinterpret_cast<float4*>(dst);
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total2; t += stride) {
      const int64_t r = t / row2, e = t - r * row2;      // destination row r = a*nb + b
      const int64_t a = r / nb, b = r - a * nb;
      d4[t] = s4[(b * na + a) * row2 + e];
    }
  } else {
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += stride) {
      const int64_t r = t / row, e = t - r * row;
      const int64_t a = r / nb, b = r - a * nb;
      dst[t] = src[(b * na + a) * row + e];
    }
  }
}

static int swap01(nb_ctx* ctx, const void* src, void* dst, int na, int nb, int64_t row, cudaStream_t st) {
  NB_REQUIRE(ctx && src && dst && src != dst, "NULL or aliased buffers");
  NB_REQUIRE(na >= 1 && nb >= 1 && row >= 1, "bad shape");
  NB_REQUIRE((((uintptr_t)src | (uintptr_t)dst) & 15u) == 0, "buffers must be 16-byte aligned");
  NB_CHECK_CUDA(cudaSetDevice(ctx->device));
  const int64_t total = (int64_t)na * nb * row;
  slab_swap01_kernel<<<nb_grid_for(ctx, total / 2 + 1, 256, 16), 256, 0, st>>>((const float2*)src, (float2*)dst, na,
                                                                              nb, row);
  NB_LAUNCHED(ctx);
  return NB_OK;
}

// [nxl][G][nyl·nzh] -> [G][nxl][nyl·nzh]   (complex64)
extern "C" int nb_slab_pack(nb_ctx* ctx, const void* src, void* dst, int nxl, int G, int nyl, int nzh,
                            void* stream) {
  return swap01(ctx, src, dst, G, nxl, (int64_t)nyl * nzh, (cudaStream_t)stream);
}
// [G][nxl][nyl·nzh] -> [nxl][G][nyl·nzh] = [nxl][N][nzh]
extern "C" int nb_slab_unpack(nb_ctx* ctx, const void* src, void* dst, int nxl, int G, int nyl, int nzh,
                              void* stream) {
  return swap01(ctx, src, dst, nxl, G, (int64_t)nyl * nzh, (cudaStream_t)stream);
}

// Batched 2-D transforms over (y,z) of `nplanes` x-planes.  dir = +1: R2C real[nplanes][N][N]
// -> spec[nplanes][N][N/2+1]; dir = -1: C2R the other way (unnormalised, spec is destroyed).
extern "C" int nb_slab_fft2d(nb_ctx* ctx, int N, int nplanes, int dir, void* real, void* spec, void* stream) {
  NB_REQUIRE(ctx && real && spec, "NULL argument");
  NB_REQUIRE(N >= 2 && nplanes >= 1 && (dir == 1 || dir == -1), "bad N / nplanes / dir");
  cudaStream_t st = (cudaStream_t)stream;
  NB_CHECK_CUDA(cudaSetDevice(ctx->device));
  const int n[2] = {N, N};
  const int nzh = N / 2 + 1;
  const int rembed[2] = {N, N}, cembed[2] = {N, nzh};
  cufftHandle h;
  int rc;
  if (dir == 1) {
    if ((rc = nb_plan_many(ctx, 2, n, nplanes, CUFFT_R2C, 1, rembed, 1, N * N, cembed, 1, N * nzh, st, &h))) return rc;
    NB_CHECK_CUFFT(cufftExecR2C(h, (cufftReal*)real, (cufftComplex*)spec));
  } else {
    if ((rc = nb_plan_many(ctx, 2, n, nplanes, CUFFT_C2R, 1, cembed, 1, N * nzh, rembed, 1, N * N, st, &h))) return rc;
    NB_CHECK_CUFFT(cufftExecC2R(h, (cufftComplex*)spec, (cufftReal*)real));
  }
  return NB_OK;
}

// In-place 1-D C2C along x of spec[N][rows] (stride `rows`, `rows` independent transforms).
extern "C" int nb_slab_fft1d_x(nb_ctx* ctx, int N, int64_t rows, int dir, void* spec, void* stream) {
  NB_REQUIRE(ctx && spec, "NULL argument");
  NB_REQUIRE(N >= 2 && rows >= 1 && rows < ((int64_t)1 << 31) && (dir == 1 || dir == -1), "bad N / rows / dir");
  cudaStream_t st = (cudaStream_t)stream;
  NB_CHECK_CUDA(cudaSetDevice(ctx->device));
  const int n[1] = {N};
  const int embed[1] = {N};
  cufftHandle h;
  int rc;
  if ((rc = nb_plan_many(ctx, 1, n, (int)rows, CUFFT_C2C, 2, embed, (int)rows, 1, embed,
                         (int)rows, 1, st, &h)))
    return rc;
  NB_CHECK_CUFFT(cufftExecC2C(h, (cufftComplex*)spec, (cufftComplex*)spec, dir == 1 ? CUFFT_FORWARD : CUFFT_INVERSE));
  return NB_OK;
}

// owner[p] = (int(x/cell) % N) / planes_per_rank : the rank whose x-slab holds the particle,
// with the deposit's float64 index math (bit-exact cell index).
__global__ void __launch_bounds__(256)
slab_owner_kernel(const double* __restrict__ pos, int64_t np, double cell, int N, int ppr, int* __restrict__ owner,
                  unsigned long long* __restrict__ counts, int G) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < np; p += stride) {
    const int o = nb_cell_index(pos[3 * p], cell, N) / ppr;
    owner[p] = o;
    if (counts) {
      // warp-aggregated count: almost all lanes of a warp share one owner
      const unsigned m = __match_any_sync(__activemask(), o);
      if ((int)(__ffs(m) - 1) == (int)(threadIdx.x & 31)) atomicAdd(&counts[o < G ? o : G - 1], (unsigned long long)__popc(m));
    }
  }
}

extern "C" int nb_slab_owner(nb_ctx* ctx, const double* pos, int64_t np, double cell, int N, int planes_per_rank,
                             int32_t* owner, int64_t* counts, int G, void* stream) {
  NB_REQUIRE(ctx && owner && (np == 0 || pos), "NULL argument");
  NB_REQUIRE(N >= 1 && cell > 0 && planes_per_rank >= 1 && G >= 1, "bad N / cell / planes_per_rank / G");
  cudaStream_t st = (cudaStream_t)stream;
  NB_CHECK_CUDA(cudaSetDevice(ctx->device));
  if (counts) NB_CHECK_CUDA(cudaMemsetAsync(counts, 0, sizeof(int64_t) * G, st));
  if (np == 0) return NB_OK;
  slab_owner_kernel<<<nb_grid_for(ctx, np, 256, 16), 256, 0, st>>>(pos, np, cell, N, planes_per_rank, owner,
                                                                   (unsigned long long*)counts, G);
  NB_LAUNCHED(ctx);
  return NB_OK;
}

// dst[i] += src[i]  (ghost planes received from the previous slab are added to the first
// own planes of the density mesh)
__global__ void __launch_bounds__(256)
slab_add_kernel(float* __restrict__ dst, const float* __restrict__ src, int64_t n) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += stride) dst[t] += src[t];
}

extern "C" int nb_grid_add(nb_ctx* ctx, float* dst, const float* src, int64_t n, void* stream) {
  NB_REQUIRE(ctx && (n == 0 || (dst && src)), "NULL argument");
  if (n <= 0) return NB_OK;
  NB_CHECK_CUDA(cudaSetDevice(ctx->device));
  slab_add_kernel<<<nb_grid_for(ctx, n, 256, 16), 256, 0, (cudaStream_t)stream>>>(dst, src, n);
  NB_LAUNCHED(ctx);
  return NB_OK;
}
