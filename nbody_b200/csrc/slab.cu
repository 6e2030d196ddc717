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

// ---------------------------------------------------------------------------------------------------
// Device-side particle exchange of the slab decomposition (the halo selection and
// the migration were eager boolean masks / nonzero / cat with two host round trips per step).
//
// Halo: the positions within r_cut of the two slab faces are appended to two fixed-capacity send
// buffers (warp-aggregated atomics); unused rows stay NaN (the caller memsets the buffers to 0xFF), and
// nb_cell_sort skips NaN rows — so the neighbours exchange FIXED-size buffers and nothing about the
// counts has to reach the host.  counts[0..1] still receive the true numbers (overflow check, folded
// into the one host sync of the step).
__device__ __forceinline__ int slab_warp_append(bool take, int* counter) {
  const unsigned m = __ballot_sync(__activemask(), take);
  if (!take) return -1;
  const int lane = threadIdx.x & 31, leader = __ffs(m) - 1;
  int base = 0;
  if (lane == leader) base = atomicAdd(counter, __popc(m));
  base = __shfl_sync(m, base, leader);
  return base + __popc(m & ((1u << lane) - 1u));
}

__global__ void __launch_bounds__(256)
halo_select_kernel(const double* __restrict__ pos, int64_t np, double lo_edge, double hi_edge,
                   double* __restrict__ out_lo, double* __restrict__ out_hi, int cap, int* __restrict__ counts) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t nround = (np + stride - 1) / stride;
  for (int64_t r = 0; r < nround; ++r) {                 // whole warps stay in the loop (ballots)
    const int64_t p = r * stride + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const bool in = p < np;
    double x = 0.0, y = 0.0, z = 0.0;
    if (in) { x = pos[3 * p]; y = pos[3 * p + 1]; z = pos[3 * p + 2]; }
    const int slo = slab_warp_append(in && x < lo_edge, counts);
    if (slo >= 0 && slo < cap) { out_lo[3 * slo] = x; out_lo[3 * slo + 1] = y; out_lo[3 * slo + 2] = z; }
    const int shi = slab_warp_append(in && x >= hi_edge, counts + 1);
    if (shi >= 0 && shi < cap) { out_hi[3 * shi] = x; out_hi[3 * shi + 1] = y; out_hi[3 * shi + 2] = z; }
  }
}

extern "C" int nb_halo_select(nb_ctx* ctx, const double* pos, int64_t np, double lo_edge, double hi_edge,
                              double* out_lo, double* out_hi, int cap, int32_t* counts2, void* stream) {
  NB_REQUIRE(ctx && out_lo && out_hi && counts2 && (np == 0 || pos) && cap >= 1, "bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  NB_CHECK_CUDA(cudaSetDevice(ctx->device));
  NB_CHECK_CUDA(cudaMemsetAsync(out_lo, 0xFF, sizeof(double) * 3 * (size_t)cap, st));   // all-ones = NaN
  NB_CHECK_CUDA(cudaMemsetAsync(out_hi, 0xFF, sizeof(double) * 3 * (size_t)cap, st));
  NB_CHECK_CUDA(cudaMemsetAsync(counts2, 0, 2 * sizeof(int32_t), st));
  if (np == 0) return NB_OK;
  halo_select_kernel<<<nb_grid_for(ctx, np, 256, 16), 256, 0, st>>>(pos, np, lo_edge, hi_edge, out_lo, out_hi, cap,
                                                                   counts2);
  NB_LAUNCHED(ctx);
  return NB_OK;
}

// Migration, step 1: the particles whose owner (nb_slab_owner) is not `rank` are copied — position and
// momentum, 6 doubles — to the send buffer of the previous (owner == prev) or next (owner == next) slab,
// and their rows are recorded as holes.  counts4 = {to_prev, to_next, holes, far}: `far` counts particles
// that crossed more than one slab (left in place; the caller falls back to its general exchange).
__global__ void __launch_bounds__(256)
migrate_split_kernel(const double* __restrict__ pos, const double* __restrict__ mom, int64_t np,
                     const int* __restrict__ owner, int rank, int prev, int next, double* __restrict__ send_prev,
                     double* __restrict__ send_next, int cap, int* __restrict__ holes, int* __restrict__ counts) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t nround = (np + stride - 1) / stride;
  for (int64_t r = 0; r < nround; ++r) {
    const int64_t p = r * stride + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int o = p < np ? owner[p] : rank;
    const bool leaves = o != rank;
    const bool to_prev = leaves && o == prev;                  // world == 2: prev == next, everything goes "prev"
    const bool to_next = leaves && !to_prev && o == next;
    const bool far = leaves && !to_prev && !to_next;
    const int sp = slab_warp_append(to_prev, counts);
    const int sn = slab_warp_append(to_next, counts + 1);
    const int sh = slab_warp_append(to_prev || to_next, counts + 2);
    if (far) atomicAdd(counts + 3, 1);
    double* dst = nullptr;
    if (sp >= 0 && sp < cap) dst = send_prev + 6 * (int64_t)sp;
    if (sn >= 0 && sn < cap) dst = send_next + 6 * (int64_t)sn;
    if (dst) {
#pragma unroll
      for (int c = 0; c < 3; ++c) { dst[c] = pos[3 * p + c]; dst[3 + c] = mom[3 * p + c]; }
    }
    if (sh >= 0 && sh < 2 * cap) holes[sh] = (int)p;
  }
}

extern "C" int nb_migrate_split(nb_ctx* ctx, const double* pos, const double* mom, int64_t np,
                                const int32_t* owner, int rank, int prev, int next, double* send_prev,
                                double* send_next, int cap, int32_t* holes, int32_t* counts4, void* stream) {
  NB_REQUIRE(ctx && owner && send_prev && send_next && holes && counts4 && (np == 0 || (pos && mom)) && cap >= 1,
             "bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  NB_CHECK_CUDA(cudaSetDevice(ctx->device));
  NB_CHECK_CUDA(cudaMemsetAsync(counts4, 0, 4 * sizeof(int32_t), st));
  if (np == 0) return NB_OK;
  migrate_split_kernel<<<nb_grid_for(ctx, np, 256, 16), 256, 0, st>>>(pos, mom, np, owner, rank, prev, next,
                                                                     send_prev, send_next, cap, holes, counts4);
  NB_LAUNCHED(ctx);
  return NB_OK;
}

// Migration, step 2 (all counts known on the host after the step's one synchronisation): the n_in arrivals
// (rows of recv_a then recv_b, 6 doubles each) fill the first holes, a surplus is appended behind row np;
// if more particles left than arrived, the remaining holes below the new end are filled with the non-hole
// rows of the tail.  pos / mom must have room for np + max(0, n_in - n_holes) rows.
__global__ void __launch_bounds__(256)
migrate_fill_kernel(double* __restrict__ pos, double* __restrict__ mom, int64_t np, const int* __restrict__ holes,
                    int n_holes, const double* __restrict__ recv_a, int na, const double* __restrict__ recv_b, int nb) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= na + nb) return;
  const double* src = k < na ? recv_a + 6 * (int64_t)k : recv_b + 6 * (int64_t)(k - na);
  const int64_t d = k < n_holes ? (int64_t)holes[k] : np + (k - n_holes);
#pragma unroll
  for (int c = 0; c < 3; ++c) { pos[3 * d + c] = src[c]; mom[3 * d + c] = src[3 + c]; }
}

__global__ void __launch_bounds__(256)
migrate_tailmark_kernel(const int* __restrict__ holes, int first, int n_rem, int64_t new_n, int* __restrict__ tail_is_hole) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n_rem) return;
  const int h = holes[first + t];
  if (h >= new_n) tail_is_hole[h - new_n] = 1;
}

__global__ void __launch_bounds__(256)
migrate_pair_kernel(const int* __restrict__ holes, int first, int n_rem, int64_t new_n,
                    const int* __restrict__ tail_is_hole, int* __restrict__ src, int* __restrict__ dst,
                    int* __restrict__ cnt2) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n_rem) return;
  if (!tail_is_hole[t]) src[atomicAdd(cnt2, 1)] = (int)(new_n + t);       // a live particle in the tail
  const int h = holes[first + t];
  if (h < new_n) dst[atomicAdd(cnt2 + 1, 1)] = h;                          // an open slot below the new end
}

__global__ void __launch_bounds__(256)
migrate_move_kernel(double* __restrict__ pos, double* __restrict__ mom, const int* __restrict__ src,
                    const int* __restrict__ dst, const int* __restrict__ cnt2, int n_rem) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n_rem || k >= cnt2[0]) return;
  const int64_t s = src[k], d = dst[k];
#pragma unroll
  for (int c = 0; c < 3; ++c) { pos[3 * d + c] = pos[3 * s + c]; mom[3 * d + c] = mom[3 * s + c]; }
}

// work: 3 * n_rem + 2 int32 of scratch (n_rem = max(0, n_holes - n_in)); returns the new row count in *new_np_host
extern "C" int nb_migrate_merge(nb_ctx* ctx, double* pos, double* mom, int64_t np, const int32_t* holes,
                                int n_holes, const double* recv_a, int na, const double* recv_b, int nb,
                                int32_t* work, int64_t* new_np_host, void* stream) {
  NB_REQUIRE(ctx && new_np_host && n_holes >= 0 && na >= 0 && nb >= 0, "bad arguments");
  NB_REQUIRE((n_holes == 0 || holes) && (na == 0 || recv_a) && (nb == 0 || recv_b), "NULL buffer");
  NB_REQUIRE((na + nb == 0 && n_holes == 0) || (pos && mom), "NULL particle arrays");
  cudaStream_t st = (cudaStream_t)stream;
  NB_CHECK_CUDA(cudaSetDevice(ctx->device));
  const int n_in = na + nb;
  if (n_in > 0) {
    migrate_fill_kernel<<<(n_in + 255) / 256, 256, 0, st>>>(pos, mom, np, holes, n_holes, recv_a, na, recv_b, nb);
    NB_LAUNCHED(ctx);
  }
  const int n_rem = n_holes - n_in;
  *new_np_host = np + n_in - n_holes;
  if (n_rem > 0) {
    NB_REQUIRE(work, "work buffer needed when more particles leave than arrive");
    const int64_t new_n = np - n_rem;
    int* tail = work;
    int* src = work + n_rem;
    int* dst = src + n_rem;
    int* cnt2 = dst + n_rem;
    NB_CHECK_CUDA(cudaMemsetAsync(tail, 0, sizeof(int) * (size_t)n_rem, st));
    NB_CHECK_CUDA(cudaMemsetAsync(cnt2, 0, 2 * sizeof(int), st));
    const int g = (n_rem + 255) / 256;
    migrate_tailmark_kernel<<<g, 256, 0, st>>>(holes, n_in, n_rem, new_n, tail);
    migrate_pair_kernel<<<g, 256, 0, st>>>(holes, n_in, n_rem, new_n, tail, src, dst, cnt2);
    migrate_move_kernel<<<g, 256, 0, st>>>(pos, mom, src, dst, cnt2, n_rem);
    ctx->launches += 3;
  }
  return NB_OK;
}
