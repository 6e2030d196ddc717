// Batched snapshot -> delta service.
//
// Reference: Feeder_pipe.run_frap_to_delta (TheFeedingPipeline/Feeder_pipe.py:163-195) loops over
// the 300 CONCEPT snapshots of the training corpus and runs, per file,
//     dens  = cic_density(pos, box, ngrid)           FraP_kTilDelta.py:93-123
//     delta = dens / dens.mean() - 1                 FraP_kTilDelta.py:126-127
//     delta = ifftn(fftn(delta) / W).real            FraP_kTilDelta.py:73-91   (W == 0 -> 1)
//     delta -= delta.mean()                          Feeder_pipe.py:193 (training/validation set only)
// Each problem is tiny (32^3 particles on a 32^3 mesh: 0.8 MB of positions, 128 KB of mesh), so
// one problem per launch is pure launch latency.  Here B problems go through the same arithmetic
// in seven launches + two batched cuFFT calls: the meshes are [B][N][N][N], the particle sets are
// concatenated with a (B+1)-entry offset table (ragged sets allowed), and every per-problem
// scalar (cell size, mean, window table, subtract-mean flag) is indexed by the batch element.
// Index / weight math is nb_cell_split (bit-exact cell indices; exact divisions for float64).
#include "common.cuh"

// batch element of concatenated particle p: largest b with offsets[b] <= p
__device__ __forceinline__ int batch_of(const int64_t* __restrict__ offsets, int B, int64_t p) {
  int lo = 0, hi = B;   // offsets[lo] <= p < offsets[hi]
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (__ldg(offsets + mid) <= p) lo = mid; else hi = mid;
  }
  return lo;
}

template <typename T>
__global__ void __launch_bounds__(256)
batch_deposit_kernel(const double* __restrict__ pos, const int64_t* __restrict__ offsets, int B,
                     const double* __restrict__ cell_of, int N, T* __restrict__ grids) {
  constexpr bool EXACT = sizeof(T) == 8;
  const int64_t total = __ldg(offsets + B);
  const int64_t n3 = (int64_t)N * N * N;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < total; p += stride) {
    const int b = batch_of(offsets, B, p);
    const double cell = __ldg(cell_of + b);
    const double inv_cell = __ddiv_rn(1.0, cell);
    int i0, j0, k0;
    double dx, dy, dz;
    nb_cell_split<EXACT>(pos[3 * p + 0], cell, inv_cell, N, i0, dx);
    nb_cell_split<EXACT>(pos[3 * p + 1], cell, inv_cell, N, j0, dy);
    nb_cell_split<EXACT>(pos[3 * p + 2], cell, inv_cell, N, k0, dz);
    const int i1 = nb_wrap1(i0, N), j1 = nb_wrap1(j0, N), k1 = nb_wrap1(k0, N);
    const double wx0 = __dsub_rn(1.0, dx), wy0 = __dsub_rn(1.0, dy), wz0 = __dsub_rn(1.0, dz);
    const double w00 = __dmul_rn(wx0, wy0), w01 = __dmul_rn(wx0, dy);
    const double w10 = __dmul_rn(dx, wy0), w11 = __dmul_rn(dx, dy);
    T* g = grids + (int64_t)b * n3;
    const int64_t r00 = ((int64_t)i0 * N + j0) * N, r01 = ((int64_t)i0 * N + j1) * N;
    const int64_t r10 = ((int64_t)i1 * N + j0) * N, r11 = ((int64_t)i1 * N + j1) * N;
    atomicAdd(&g[r00 + k0], (T)__dmul_rn(w00, wz0));
    atomicAdd(&g[r00 + k1], (T)__dmul_rn(w00, dz));
    atomicAdd(&g[r01 + k0], (T)__dmul_rn(w01, wz0));
    atomicAdd(&g[r01 + k1], (T)__dmul_rn(w01, dz));
    atomicAdd(&g[r10 + k0], (T)__dmul_rn(w10, wz0));
    atomicAdd(&g[r10 + k1], (T)__dmul_rn(w10, dz));
    atomicAdd(&g[r11 + k0], (T)__dmul_rn(w11, wz0));
    atomicAdd(&g[r11 + k1], (T)__dmul_rn(w11, dz));
  }
}

// sums[b] += sum of grid b (float64); grid = (chunks, B)
template <typename T>
__global__ void __launch_bounds__(256)
batch_sum_kernel(const T* __restrict__ grids, int64_t n3, double* __restrict__ sums) {
  const int b = blockIdx.y;
  const T* g = grids + (int64_t)b * n3;
  double acc = 0.0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n3; i += (int64_t)gridDim.x * blockDim.x)
    acc += (double)g[i];
  acc = nb_warp_sum(acc);
  __shared__ double part[8];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (lane == 0) part[w] = acc;
  __syncthreads();
  if (w == 0) {
    double v = lane < (blockDim.x >> 5) ? part[lane] : 0.0;
    v = nb_warp_sum(v);
    if (lane == 0) atomicAdd(&sums[b], v);
  }
}

// mode 0: grid = grid / (sums[b]/n3) - 1   (compute_delta)
// mode 1: grid = grid - sums[b]/n3 where flags[b] != 0   (delta -= delta.mean())
template <typename T>
__global__ void __launch_bounds__(256)
batch_affine_kernel(T* __restrict__ grids, int64_t n3, const double* __restrict__ sums, int mode,
                    const int32_t* __restrict__ flags) {
  const int b = blockIdx.y;
  if (mode == 1 && !(flags && flags[b])) return;
  const double mean = __ddiv_rn(sums[b], (double)n3);
  T* g = grids + (int64_t)b * n3;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n3; i += (int64_t)gridDim.x * blockDim.x) {
    const double v = (double)g[i];
    g[i] = mode == 0 ? (T)__dsub_rn(__ddiv_rn(v, mean), 1.0) : (T)__dsub_rn(v, mean);
  }
}

// spectrum [B][N][N][N/2+1]: divide by W = ((s[i]*s[j])*s[k])^2 (W == 0 -> 1 when guard) and by N^3
template <typename C, typename T>
__global__ void __launch_bounds__(256)
batch_deconv_kernel(C* __restrict__ spec, int N, int B, const double* __restrict__ sinc_tabs, int guard,
                    double inv_n3) {
  const int nzh = N / 2 + 1;
  const int64_t per = (int64_t)N * N * nzh, total = per * B;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
    const int b = (int)(t / per);
    const int64_t u = t - (int64_t)b * per;
    const int k = (int)(u % nzh);
    const int64_t r = u / nzh;
    const int j = (int)(r % N), i = (int)(r / N);
    const double* s = sinc_tabs + (int64_t)b * N;
    double w = __dmul_rn(__dmul_rn(s[i], s[j]), s[k]);
    w = __dmul_rn(w, w);
    if (guard && w == 0.0) w = 1.0;
    C v = spec[t];
    v.x = (T)(__ddiv_rn((double)v.x, w) * inv_n3);
    v.y = (T)(__ddiv_rn((double)v.y, w) * inv_n3);
    spec[t] = v;
  }
}

template <typename T, typename C>
static int batch_run(nb_ctx* ctx, const double* pos, const int64_t* offsets, int64_t total, int B,
                     const double* cell_of, int N, const double* sinc_tabs, int guard, const int32_t* submean,
                     T* grids, cufftType fwd_t, cufftType inv_t, cudaStream_t st) {
  const int64_t n3 = (int64_t)N * N * N;
  const int64_t nspec = (int64_t)N * N * (N / 2 + 1);
  void *spec = nullptr, *sums_v = nullptr;
  int rc;
  if ((rc = nb_scratch(ctx, NB_SCR_SPEC0, (size_t)nspec * B * sizeof(C), &spec))) return rc;
  if ((rc = nb_scratch(ctx, NB_SCR_MISC0, sizeof(double) * 2 * (size_t)B, &sums_v))) return rc;
  double* sums = (double*)sums_v;
  NB_CHECK_CUDA(cudaMemsetAsync(grids, 0, sizeof(T) * (size_t)n3 * B, st));
  NB_CHECK_CUDA(cudaMemsetAsync(sums, 0, sizeof(double) * 2 * (size_t)B, st));
  if (total > 0) {
    batch_deposit_kernel<T><<<nb_grid_for(ctx, total, 256, 16), 256, 0, st>>>(pos, offsets, B, cell_of, N, grids);
    NB_LAUNCHED(ctx);
  }
  int chunks = (int)((n3 + 256 * 8 - 1) / (256 * 8));
  const int max_chunks = (ctx->num_sms * 8 + B - 1) / B;
  if (chunks > max_chunks) chunks = max_chunks;
  if (chunks < 1) chunks = 1;
  const dim3 g2(chunks, B);
  batch_sum_kernel<T><<<g2, 256, 0, st>>>(grids, n3, sums);
  NB_LAUNCHED(ctx);
  batch_affine_kernel<T><<<g2, 256, 0, st>>>(grids, n3, sums, 0, nullptr);
  NB_LAUNCHED(ctx);
  int dims[3] = {N, N, N};
  cufftHandle fwd, inv;
  if ((rc = nb_plan_many(ctx, 3, dims, B, fwd_t, 7001, nullptr, 1, 0, nullptr, 1, 0, st, &fwd))) return rc;
  if ((rc = nb_plan_many(ctx, 3, dims, B, inv_t, 7002, nullptr, 1, 0, nullptr, 1, 0, st, &inv))) return rc;
  if (sizeof(T) == 8)
    NB_CHECK_CUFFT(cufftExecD2Z(fwd, (cufftDoubleReal*)grids, (cufftDoubleComplex*)spec));
  else
    NB_CHECK_CUFFT(cufftExecR2C(fwd, (cufftReal*)grids, (cufftComplex*)spec));
  batch_deconv_kernel<C, T><<<nb_grid_for(ctx, nspec * B, 256, 8), 256, 0, st>>>((C*)spec, N, B, sinc_tabs, guard,
                                                                               1.0 / (double)n3);
  NB_LAUNCHED(ctx);
  if (sizeof(T) == 8)
    NB_CHECK_CUFFT(cufftExecZ2D(inv, (cufftDoubleComplex*)spec, (cufftDoubleReal*)grids));
  else
    NB_CHECK_CUFFT(cufftExecC2R(inv, (cufftComplex*)spec, (cufftReal*)grids));
  if (submean) {
    batch_sum_kernel<T><<<g2, 256, 0, st>>>(grids, n3, sums + B);
    NB_LAUNCHED(ctx);
    batch_affine_kernel<T><<<g2, 256, 0, st>>>(grids, n3, sums + B, 1, submean);
    NB_LAUNCHED(ctx);
  }
  return NB_OK;
}

extern "C" int nb_snapshot_delta_batch(nb_ctx* ctx, const double* pos, const int64_t* offsets_dev,
                                       int64_t total_particles, int B, const double* cell_dev, int N,
                                       const double* sinc_tabs_dev, int guard, const int32_t* submean_dev,
                                       void* grids, int prec, void* stream) {
  NB_REQUIRE(ctx && offsets_dev && cell_dev && sinc_tabs_dev && grids, "NULL argument");
  NB_REQUIRE(B >= 1 && B <= 65535, "batch size must be in [1, 65535]");
  NB_REQUIRE(N >= 2 && (int64_t)N * N * N * B < ((int64_t)1 << 40), "bad mesh size");
  NB_REQUIRE(total_particles >= 0 && (total_particles == 0 || pos), "pos is NULL");
  NB_REQUIRE(prec == NB_F32 || prec == NB_F64, "prec must be NB_F32 or NB_F64");
  NB_CHECK_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = (cudaStream_t)stream;
  if (prec == NB_F64)
    return batch_run<double, double2>(ctx, pos, offsets_dev, total_particles, B, cell_dev, N, sinc_tabs_dev, guard,
                                      submean_dev, (double*)grids, CUFFT_D2Z, CUFFT_Z2D, st);
  return batch_run<float, float2>(ctx, pos, offsets_dev, total_particles, B, cell_dev, N, sinc_tabs_dev, guard,
                                  submean_dev, (float*)grids, CUFFT_R2C, CUFFT_C2R, st);
}

// ---------------------------------------------------------------------------------------------------
// Batched snapshot P(k): replaces the process pool of Different_Sizes/Plotter.py:205-238
// (`mp.Pool(10).starmap(compute_power_spectrum_from_snapshot, arg_list)`, one snapshot per worker) and the
// per-file loops of SnapVsPK/Power_compare.py:211-235 for snapshots that share a mesh size: B particle sets
// are deposited into [B][N][N][N] by one kernel, turned into overdensities (grid / (Np_b / N³) - 1,
// Plotter.py:172-173 = Power_compare.py:153-157), transformed by ONE batched cuFFT call and binned by the K6
// kernel once per snapshot, all on the caller's stream without a host round trip in between.
__global__ void __launch_bounds__(256)
batch_overdensity_kernel(float* __restrict__ grids, int64_t n3, const double* __restrict__ npart) {
  const int b = blockIdx.y;
  const double inv_mean = (double)n3 / npart[b];
  float* g = grids + (int64_t)b * n3;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n3; i += (int64_t)gridDim.x * blockDim.x)
    g[i] = (float)((double)g[i] * inv_mean - 1.0);
}

int nb_pk_bin_spectrum(nb_ctx* ctx, const void* spec, int N, int y0, int nyl, int prec, const double* k2_tab,
                       const double* sinc_tab, const double* thresh, double thresh_hi, int n_edges,
                       int window_power, int clip_mode, double clip_val, double norm, int closed_last,
                       double k2min_host, double k2max_host, int64_t* counts, double* sums, cudaStream_t st);

extern "C" int nb_snapshot_pk_batch(nb_ctx* ctx, const double* pos, const int64_t* offsets_dev,
                                    int64_t total_particles, int B, const double* cell_dev,
                                    const double* npart_dev, int N, const double* k2_tabs_dev,
                                    const double* sinc_tabs_dev, const double* thresh_dev,
                                    const double* thresh_first_host, const double* thresh_last_host,
                                    const double* thresh_hi_host, int n_edges, int window_power, int clip_mode,
                                    double clip_val, const double* norm_host, int closed_last, int64_t* counts,
                                    double* sums, void* stream) {
  NB_REQUIRE(ctx && offsets_dev && cell_dev && npart_dev && k2_tabs_dev && sinc_tabs_dev && thresh_dev &&
                 thresh_first_host && thresh_last_host && thresh_hi_host && norm_host && counts && sums,
             "NULL argument");
  NB_REQUIRE(B >= 1 && B <= 65535, "batch size must be in [1, 65535]");
  NB_REQUIRE(N >= 2 && n_edges >= 2 && (int64_t)N * N * N * B < ((int64_t)1 << 36), "bad mesh size / n_edges");
  NB_REQUIRE(total_particles >= 0 && (total_particles == 0 || pos), "pos is NULL");
  NB_REQUIRE(window_power >= 0 && window_power <= 2 && clip_mode >= 0 && clip_mode <= 2, "bad window options");
  NB_CHECK_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t n3 = (int64_t)N * N * N, nspec = (int64_t)N * N * (N / 2 + 1);
  void *grids_v = nullptr, *spec = nullptr;
  int rc;
  if ((rc = nb_scratch(ctx, NB_SCR_REAL0, sizeof(float) * (size_t)n3 * B, &grids_v))) return rc;
  if ((rc = nb_scratch(ctx, NB_SCR_SPEC0, sizeof(float2) * (size_t)nspec * B, &spec))) return rc;
  float* grids = (float*)grids_v;
  NB_CHECK_CUDA(cudaMemsetAsync(grids, 0, sizeof(float) * (size_t)n3 * B, st));
  NB_CHECK_CUDA(cudaMemsetAsync(counts, 0, sizeof(int64_t) * (size_t)(n_edges - 1) * B, st));
  NB_CHECK_CUDA(cudaMemsetAsync(sums, 0, sizeof(double) * (size_t)(n_edges - 1) * B, st));
  if (total_particles > 0) {
    batch_deposit_kernel<float><<<nb_grid_for(ctx, total_particles, 256, 16), 256, 0, st>>>(pos, offsets_dev, B,
                                                                                         cell_dev, N, grids);
    NB_LAUNCHED(ctx);
  }
  int chunks = (int)((n3 + 256 * 8 - 1) / (256 * 8));
  const int max_chunks = (ctx->num_sms * 8 + B - 1) / B;
  if (chunks > max_chunks) chunks = max_chunks;
  if (chunks < 1) chunks = 1;
  batch_overdensity_kernel<<<dim3(chunks, B), 256, 0, st>>>(grids, n3, npart_dev);
  NB_LAUNCHED(ctx);
  int dims[3] = {N, N, N};
  cufftHandle fwd;
  if ((rc = nb_plan_many(ctx, 3, dims, B, CUFFT_R2C, 7001, nullptr, 1, 0, nullptr, 1, 0, st, &fwd))) return rc;
  NB_CHECK_CUFFT(cufftExecR2C(fwd, (cufftReal*)grids, (cufftComplex*)spec));
  for (int b = 0; b < B; ++b) {
    NB_REQUIRE(thresh_first_host[b] > 0 && thresh_last_host[b] > thresh_first_host[b], "bad bin thresholds");
    if ((rc = nb_pk_bin_spectrum(ctx, (const float2*)spec + (int64_t)b * nspec, N, 0, N, NB_F32,
                                 k2_tabs_dev + (int64_t)b * N, sinc_tabs_dev + (int64_t)b * N,
                                 thresh_dev + (int64_t)b * n_edges, thresh_hi_host[b], n_edges, window_power, clip_mode,
                                 clip_val, norm_host[b], closed_last, thresh_first_host[b], thresh_last_host[b],
                                 counts + (int64_t)b * (n_edges - 1), sums + (int64_t)b * (n_edges - 1), st)))
      return rc;
  }
  return NB_OK;
}
