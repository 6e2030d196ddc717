// K8 — element-wise grid operations and the CIC-window deconvolution (a2, a3).
#include "common.cuh"

// ------------------------------------------------------------------ sum of a grid (float64)
// 16-byte loads, four independent float64 partial sums per thread (the grid is streamed once).
template <typename T>
__global__ void __launch_bounds__(256) grid_sum_kernel(const T* __restrict__ g, int64_t n, double* out) {
  constexpr int V = 16 / sizeof(T);                 // elements per 16-byte vector
  double acc[4] = {0.0, 0.0, 0.0, 0.0};
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t nvec = 0;
  if ((reinterpret_cast<uintptr_t>(g) & 15) == 0) {
    nvec = n / V;
    const int4* gv = reinterpret_cast<const int4*>(g);
    for (int64_t i = tid; i < nvec; i += stride) {
      const int4 raw = __ldg(gv + i);
      const T* e = reinterpret_cast<const T*>(&raw);
#pragma unroll
      for (int k = 0; k < V; ++k) acc[k & 3] += (double)e[k];
    }
  }
  for (int64_t i = nvec * V + tid; i < n; i += stride) acc[0] += (double)g[i];
  double a = nb_warp_sum((acc[0] + acc[1]) + (acc[2] + acc[3]));
  __shared__ double part[8];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (lane == 0) part[w] = a;
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
  const int nblk = nb_grid_for(ctx, n / 4 + 1, 256, 8);
  if (prec == NB_F64)
    grid_sum_kernel<double><<<nblk, 256, 0, st>>>((const double*)grid, n, out_sum);
  else
    grid_sum_kernel<float><<<nblk, 256, 0, st>>>((const float*)grid, n, out_sum);
  NB_LAUNCHED(ctx);
  return NB_OK;
}

// ------------------------------------------------------------------ grid <- (grid/a)/b - c
// float64 grids keep the reference's two IEEE divisions (compute_delta, FraP_kTilDelta.py:126;
// Power_compare.py:153-159).  float32 grids multiply by the float64 reciprocals instead: the
// float64 result differs by <= 2 ulp(double) before it is rounded to float32, far inside the
// 1e-5 tolerance of the float32 path, and the kernel becomes a 16-byte-vector stream.
template <typename T>
__device__ __forceinline__ T grid_affine_one(T x, double a, double b, double ra, double rb, double c) {
  double v = (double)x;
  if constexpr (sizeof(T) == 8) {
    if (a != 1.0) v = __ddiv_rn(v, a);
    if (b != 1.0) v = __ddiv_rn(v, b);
  } else {
    if (a != 1.0) v = __dmul_rn(v, ra);
    if (b != 1.0) v = __dmul_rn(v, rb);
  }
  return (T)__dsub_rn(v, c);
}

template <typename T>
__global__ void __launch_bounds__(256)
grid_affine_kernel(T* __restrict__ g, int64_t n, double a, const double* __restrict__ a_dev, double a_div,
                   double b, double c) {
  if (a_dev) a = __ddiv_rn(*a_dev, a_div);
  const double ra = __ddiv_rn(1.0, a), rb = __ddiv_rn(1.0, b);
  constexpr int V = 16 / sizeof(T);
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t nvec = 0;
  if ((reinterpret_cast<uintptr_t>(g) & 15) == 0) {
    nvec = n / V;
    int4* gv = reinterpret_cast<int4*>(g);
    for (int64_t i = tid; i < nvec; i += stride) {
      int4 raw = gv[i];
      T* e = reinterpret_cast<T*>(&raw);
#pragma unroll
      for (int k = 0; k < V; ++k) e[k] = grid_affine_one<T>(e[k], a, b, ra, rb, c);
      gv[i] = raw;
    }
  }
  for (int64_t i = nvec * V + tid; i < n; i += stride) g[i] = grid_affine_one<T>(g[i], a, b, ra, rb, c);
}

extern "C" int nb_grid_affine(nb_ctx* ctx, void* grid, int64_t n, int prec, double a, const double* a_dev,
                              double a_div, double b, double c, void* stream) {
  NB_REQUIRE(ctx && grid, "NULL argument");
  NB_REQUIRE(prec == NB_F32 || prec == NB_F64, "bad prec");
  cudaStream_t st = (cudaStream_t)stream;
  NB_CHECK_CUDA(cudaSetDevice(ctx->device));
  const int nblk = nb_grid_for(ctx, n / 4 + 1, 256, 8);
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

// ------------------------------------------------------------------ max |a|² over (n,3) float64 vectors
// Step-size criterion of the sub-cycled time loop (timeloop.evolve, SURVEY §8f row 1: CONCEPT assigns
// particles to rungs by their short-range acceleration, log "Determining rung population" :42,55):
// the largest squared acceleration of the set, one atomicMax per CTA on the bit pattern of a
// non-negative double.
__global__ void __launch_bounds__(256) vec3_max_norm2_kernel(const double* __restrict__ v, int64_t n,
                                                              unsigned long long* out) {
  double m = 0.0;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const double x = v[3 * i], y = v[3 * i + 1], z = v[3 * i + 2];
    const double s = x * x + y * y + z * z;
    m = s > m ? s : m;            // NaN compares false: it is reported through the count below
    if (!(s == s)) m = __longlong_as_double(0x7ff0000000000000LL);
  }
  for (int o = 16; o > 0; o >>= 1) {
    const double t = __shfl_xor_sync(0xffffffffu, m, o);
    m = t > m ? t : m;
  }
  __shared__ double part[8];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (lane == 0) part[w] = m;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int k = 1; k < (int)(blockDim.x >> 5); ++k) m = part[k] > m ? part[k] : m;
    atomicMax(out, (unsigned long long)__double_as_longlong(m));   // order-preserving for doubles >= 0
  }
}

extern "C" int nb_vec3_max_norm2(nb_ctx* ctx, const double* v, int64_t n, double* out_max, void* stream) {
  NB_REQUIRE(ctx && v && out_max && n >= 0, "bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  NB_CHECK_CUDA(cudaSetDevice(ctx->device));
  NB_CHECK_CUDA(cudaMemsetAsync(out_max, 0, sizeof(double), st));
  if (n == 0) return NB_OK;
  const int nblk = nb_grid_for(ctx, n, 256, 8);
  vec3_max_norm2_kernel<<<nblk, 256, 0, st>>>(v, n, (unsigned long long*)out_max);
  NB_LAUNCHED(ctx);
  return NB_OK;
}
