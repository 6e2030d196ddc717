// K8 — element-wise grid operations and the CIC-window deconvolution (a2, a3).
#include "common.cuh"

// ------------------------------------------------------------------ sum of a grid (float64)
template <typename T>
__global__ void __launch_bounds__(256) grid_sum_kernel(const T* __restrict__ g, int64_t n, double* out) {
  double acc = 0.0;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) acc += (double)g[i];
  acc = nb_warp_sum(acc);
  __shared__ double part[8];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (lane == 0) part[w] = acc;
  __syncthreads();
  if (w == 0) {
    double v = lane < (blockDim.x >> 5) ? part[lane] : 0.0;
    v = nb_warp_sum(v);
    if (lane == 0) atomicAdd(out, v);
  }
}

extern "C" int nb_grid_sum(nb_ctx* ctx, const void* grid, int64_t n, int prec, double* out_sum, void* stream) {
  NB_REQUIRE(ctx && grid && out_sum, "NULL argument");
  NB_REQUIRE(prec == NB_F32 || prec == NB_F64, "bad prec");
  cudaStream_t st = (cudaStream_t)stream;
  NB_CHECK_CUDA(cudaSetDevice(ctx->device));
  NB_CHECK_CUDA(cudaMemsetAsync(out_sum, 0, sizeof(double), st));
  const int nblk = nb_grid_for(ctx, n / 4 + 1, 256, 4);
  if (prec == NB_F64)
    grid_sum_kernel<double><<<nblk, 256, 0, st>>>((const double*)grid, n, out_sum);
  else
    grid_sum_kernel<float><<<nblk, 256, 0, st>>>((const float*)grid, n, out_sum);
  NB_LAUNCHED(ctx);
  return NB_OK;
}

// ------------------------------------------------------------------ grid <- (grid/a)/b - c
template <typename T>
__global__ void __launch_bounds__(256)
grid_affine_kernel(T* __restrict__ g, int64_t n, double a, const double* __restrict__ a_dev, double a_div,
                   double b, double c) {
  if (a_dev) a = __ddiv_rn(*a_dev, a_div);
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    double v = (double)g[i];
    if (a != 1.0) v = __ddiv_rn(v, a);
    if (b != 1.0) v = __ddiv_rn(v, b);
    g[i] = (T)__dsub_rn(v, c);
  }
}

extern "C" int nb_grid_affine(nb_ctx* ctx, void* grid, int64_t n, int prec, double a, const double* a_dev,
                              double a_div, double b, double c, void* stream) {
  NB_REQUIRE(ctx && grid, "NULL argument");
  NB_REQUIRE(prec == NB_F32 || prec == NB_F64, "bad prec");
  cudaStream_t st = (cudaStream_t)stream;
  NB_CHECK_CUDA(cudaSetDevice(ctx->device));
  const int nblk = nb_grid_for(ctx, n, 256, 8);
  if (prec == NB_F64)
    grid_affine_kernel<double><<<nblk, 256, 0, st>>>((double*)grid, n, a, a_dev, a_div, b, c);
  else
    grid_affine_kernel<float><<<nblk, 256, 0, st>>>((float*)grid, n, a, a_dev, a_div, b, c);
  NB_LAUNCHED(ctx);
  return NB_OK;
}

// ------------------------------------------------------------------ deconvolution
// spectrum layout: [N][N][N/2+1] complex (cuFFT R2C); W = ((s[i]*s[j])*s[k])^2.
template <typename C, typename T>
__global__ void __launch_bounds__(256)
deconv_kernel(C* __restrict__ spec, int N, const double* __restrict__ s, int guard, double inv_n3) {
  const int nzh = N / 2 + 1;
  const int64_t total = (int64_t)N * N * nzh;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += stride) {
    const int k = (int)(t % nzh);
    const int64_t r = t / nzh;
    const int j = (int)(r % N), i = (int)(r / N);
    double w = __dmul_rn(__dmul_rn(s[i], s[j]), s[k]);
    w = __dmul_rn(w, w);
    if (guard && w == 0.0) w = 1.0;
    C v = spec[t];
    v.x = (T)(__ddiv_rn((double)v.x, w) * inv_n3);
    v.y = (T)(__ddiv_rn((double)v.y, w) * inv_n3);
    spec[t] = v;
  }
}

extern "C" int nb_deconvolve_cic(nb_ctx* ctx, void* grid, int N, int prec, const double* sinc_tab,
                                 int guard, void* stream) {
  NB_REQUIRE(ctx && grid && sinc_tab, "NULL argument");
  NB_REQUIRE(N >= 2, "N must be >= 2");
  NB_REQUIRE(prec == NB_F32 || prec == NB_F64, "bad prec");
  cudaStream_t st = (cudaStream_t)stream;
  NB_CHECK_CUDA(cudaSetDevice(ctx->device));
  const int64_t nspec = (int64_t)N * N * (N / 2 + 1);
  void* spec = nullptr;
  int rc = nb_scratch(ctx, NB_SCR_SPEC0, (size_t)nspec * (prec == NB_F64 ? 16 : 8), &spec);
  if (rc) return rc;
  const double inv_n3 = 1.0 / ((double)N * N * N);
  const int nblk = nb_grid_for(ctx, nspec, 256, 8);
  cufftHandle fwd, inv;
  if (prec == NB_F64) {
    if ((rc = nb_plan3d(ctx, N, CUFFT_D2Z, st, &fwd))) return rc;
    if ((rc = nb_plan3d(ctx, N, CUFFT_Z2D, st, &inv))) return rc;
    NB_CHECK_CUFFT(cufftExecD2Z(fwd, (cufftDoubleReal*)grid, (cufftDoubleComplex*)spec));
    deconv_kernel<double2, double><<<nblk, 256, 0, st>>>((double2*)spec, N, sinc_tab, guard, inv_n3);
    NB_LAUNCHED(ctx);
    NB_CHECK_CUFFT(cufftExecZ2D(inv, (cufftDoubleComplex*)spec, (cufftDoubleReal*)grid));
  } else {
    if ((rc = nb_plan3d(ctx, N, CUFFT_R2C, st, &fwd))) return rc;
    if ((rc = nb_plan3d(ctx, N, CUFFT_C2R, st, &inv))) return rc;
    NB_CHECK_CUFFT(cufftExecR2C(fwd, (cufftReal*)grid, (cufftComplex*)spec));
    deconv_kernel<float2, float><<<nblk, 256, 0, st>>>((float2*)spec, N, sinc_tab, guard, inv_n3);
    NB_LAUNCHED(ctx);
    NB_CHECK_CUFFT(cufftExecC2R(inv, (cufftComplex*)spec, (cufftReal*)grid));
  }
  return NB_OK;
}
