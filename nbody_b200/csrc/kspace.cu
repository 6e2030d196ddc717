// K2 — fused k-space kernel of the PM / P3M long-range solve:
//   g_k = pref · (i k / k²) · exp(-k² r_s²) / W^p · rho_k      (g_0 = 0)
// Green's function, Gaussian long-range filter, CIC deconvolution (reference window
// FraP_kTilDelta.py:73-84, power p = deconv_order, `deconvolve (True,True)` => p = 2),
// Fourier gradient (the reference's ψ operator i·k/k², CiC/3D startbetingelser1.py:105-107),
// upstream interlacing combine and optional downstream interlacing outputs — one pass:
// 8 B (16 B interlaced) read and 24 B (48 B) written per retained R2C mode.
//
// All factors are separable per axis and live in shared-memory float32 tables built by
// each CTA from the host float64 tables (k_vals, sinc).  One warp per (ix,iy) row, lanes
// along kz (coalesced 8-byte accesses).
#include "common.cuh"

#define KS_WARPS 8

__global__ void __launch_bounds__(KS_WARPS * 32)
kspace_fused_kernel(const float2* __restrict__ rho1, const float2* __restrict__ rho2, int N, int y0, int nyl,
                    const double* __restrict__ kv, const double* __restrict__ sinc, double r_s,
                    int deconv_order, float pref, float2* __restrict__ gx, float2* __restrict__ gy,
                    float2* __restrict__ gz, float2* __restrict__ hx, float2* __restrict__ hy,
                    float2* __restrict__ hz) {
  extern __shared__ float s_tab[];
  float* s_kg = s_tab;          // gradient k (Nyquist zeroed)
  float* s_k2 = s_kg + N;       // k²
  float* s_t = s_k2 + N;        // exp(-k² r_s²) / sinc^(2p)
  float* s_pc = s_t + N;        // cos(pi m / N)  (0 at Nyquist)
  float* s_ps = s_pc + N;       // sin(pi m / N)  (0 at Nyquist)
  const int nyq = (N % 2 == 0) ? N / 2 : -1;
  for (int m = threadIdx.x; m < N; m += blockDim.x) {
    const double k = kv[m];
    const double s = sinc[m];
    double t = exp(-(k * k) * (r_s * r_s));
    for (int q = 0; q < deconv_order; ++q) t /= (s * s);
    s_kg[m] = (m == nyq) ? 0.f : (float)k;
    s_k2[m] = (float)(k * k);
    s_t[m] = (float)t;
    const int ms = (m <= N / 2 && m != nyq) ? m : m - N;   // signed frequency (fftfreq)
    double sn, cs;
    sincospi((double)ms / (double)N, &sn, &cs);
    s_pc[m] = (m == nyq) ? 0.f : (float)cs;
    s_ps[m] = (m == nyq) ? 0.f : (float)sn;
  }
  __syncthreads();

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nzh = N / 2 + 1;
  // spectra are [N (kx)][nyl (ky = y0 .. y0+nyl-1)][N/2+1 (kz)]: the whole half spectrum
  // (y0 = 0, nyl = N) or the y-slab a rank holds after the slab-FFT transpose
  const int64_t nrows = (int64_t)N * nyl;
  const int64_t wstride = (int64_t)gridDim.x * KS_WARPS;
  for (int64_t row = (int64_t)blockIdx.x * KS_WARPS + warp; row < nrows; row += wstride) {
    const int j = y0 + (int)(row % nyl), i = (int)(row / nyl);
    const float kgx = s_kg[i], kgy = s_kg[j];
    const float k2xy = s_k2[i] + s_k2[j];
    const float txy = s_t[i] * s_t[j] * pref;
    // phase of the (x,y) part: (cx + i sx)(cy + i sy)
    const float pcxy = s_pc[i] * s_pc[j] - s_ps[i] * s_ps[j];
    const float psxy = s_pc[i] * s_ps[j] + s_ps[i] * s_pc[j];
    const int64_t base = row * nzh;
#pragma unroll 2
    for (int kz = lane; kz < nzh; kz += 32) {
      const int64_t o = base + kz;
      float2 r = rho1[o];
      float pc = 1.f, ps = 0.f;
      if (rho2 || hx) {
        pc = pcxy * s_pc[kz] - psxy * s_ps[kz];
        ps = pcxy * s_ps[kz] + psxy * s_pc[kz];
      }
      if (rho2) {
        const float2 q = rho2[o];   // 0.5 (rho1 + rho2 * e^{+i k.s})
        r.x = 0.5f * (r.x + q.x * pc - q.y * ps);
        r.y = 0.5f * (r.y + q.x * ps + q.y * pc);
      }
      const float k2 = k2xy + s_k2[kz];
      const float f = (k2 > 0.f) ? txy * s_t[kz] / k2 : 0.f;
      // i * k * f * r  = (-k f r.y, k f r.x)
      const float ax = -f * r.y, ay = f * r.x;
      const float kgz = s_kg[kz];
      gx[o] = make_float2(kgx * ax, kgx * ay);
      gy[o] = make_float2(kgy * ax, kgy * ay);
      gz[o] = make_float2(kgz * ax, kgz * ay);
      if (hx) {                      // downstream interlacing: g * e^{-i k.s}
        const float bx = ax * pc + ay * ps, by = ay * pc - ax * ps;
        hx[o] = make_float2(kgx * bx, kgx * by);
        hy[o] = make_float2(kgy * bx, kgy * by);
        hz[o] = make_float2(kgz * bx, kgz * by);
      }
    }
  }
}

// rho_k (half spectrum, cuFFT R2C layout) -> 3 (or 6) force spectra.  Exposed for the
// slab-decomposed path and for tests; nb_pm_force_mesh is the single-GPU composition.
static int kspace_launch(nb_ctx* ctx, const void* rho1, const void* rho2, int N, int y0, int nyl,
                         const double* k_vals, const double* sinc_tab, double r_s, int deconv_order, double pref,
                         void* g3, void* h3, cudaStream_t st) {
  NB_REQUIRE(ctx && rho1 && k_vals && sinc_tab && g3, "NULL argument");
  NB_REQUIRE(N >= 2 && N <= 8192, "bad N");
  NB_REQUIRE(y0 >= 0 && nyl >= 1 && y0 + nyl <= N, "bad y-slab");
  NB_REQUIRE(deconv_order >= 0 && deconv_order <= 2, "deconv_order must be 0, 1 or 2");
  NB_CHECK_CUDA(cudaSetDevice(ctx->device));
  const int64_t nspec = (int64_t)N * nyl * (N / 2 + 1);
  const size_t smem = 5 * sizeof(float) * (size_t)N;
  NB_CHECK_CUDA(cudaFuncSetAttribute(kspace_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  static const int ks_ctas = [] { const char* e = getenv("NB_KS_CTAS"); const int v = e ? atoi(e) : 8; return v >= 1 && v <= 16 ? v : 8; }();
  int nblk = ctx->num_sms * ks_ctas;
  const int64_t need = ((int64_t)N * nyl + KS_WARPS - 1) / KS_WARPS;
  if (nblk > need) nblk = (int)need;
  float2* g = (float2*)g3;
  float2* h = (float2*)h3;
  kspace_fused_kernel<<<nblk, KS_WARPS * 32, smem, st>>>(
      (const float2*)rho1, (const float2*)rho2, N, y0, nyl, k_vals, sinc_tab, r_s, deconv_order, (float)pref, g,
      g + nspec, g + 2 * nspec, h, h ? h + nspec : nullptr, h ? h + 2 * nspec : nullptr);
  NB_LAUNCHED(ctx);
  return NB_OK;
}

extern "C" int nb_kspace_fused(nb_ctx* ctx, const void* rho1, const void* rho2, int N, const double* k_vals,
                               const double* sinc_tab, double r_s, int deconv_order, double pref,
                               void* g3, void* h3, void* stream) {
  return kspace_launch(ctx, rho1, rho2, N, 0, N, k_vals, sinc_tab, r_s, deconv_order, pref, g3, h3,
                       (cudaStream_t)stream);
}

// Slab form: spectra are [N][ny_local][N/2+1] with ky = y0 + local index (transposed layout
// of the slab FFT); g3 / h3 are 3 consecutive spectra of that shape.
extern "C" int nb_kspace_fused_slab(nb_ctx* ctx, const void* rho1, const void* rho2, int N, int y0, int ny_local,
                                    const double* k_vals, const double* sinc_tab, double r_s, int deconv_order,
                                    double pref, void* g3, void* h3, void* stream) {
  return kspace_launch(ctx, rho1, rho2, N, y0, ny_local, k_vals, sinc_tab, r_s, deconv_order, pref, g3, h3,
                       (cudaStream_t)stream);
}

// Single-GPU long-range solve: deposit -> R2C -> [shifted deposit -> R2C] -> K2 -> 3 (6) C2R.
// force3 (and force3_shifted when interlace_down) receive 3·N³ float32 each, in units of
// pref · counts (the caller folds 4πG·m/cell³/N³ into pref).
// rho1_in / rho2_in: meshes the caller has deposited already (nb_pm_force_mesh_rho), else deposited here from pos.
static int pm_force_mesh_impl(nb_ctx* ctx, const double* pos, int64_t np, double cell, int N,
                              const double* k_vals, const double* sinc_tab, double r_s, int deconv_order,
                              int interlace_up, int interlace_down, double pref, float* force3,
                              float* force3_shifted, const float* rho1_in, const float* rho2_in, void* stream) {
  NB_REQUIRE(ctx && (pos || rho1_in) && k_vals && sinc_tab && force3, "NULL argument");
  NB_REQUIRE(!interlace_down || force3_shifted, "interlace_down needs force3_shifted");
  NB_REQUIRE(N >= 2, "bad N");
  cudaStream_t st = (cudaStream_t)stream;
  NB_CHECK_CUDA(cudaSetDevice(ctx->device));
  const int64_t n3 = (int64_t)N * N * N;
  const int64_t nspec = (int64_t)N * N * (N / 2 + 1);
  void *rho = nullptr, *s1 = nullptr, *s2 = nullptr, *g = nullptr, *h = nullptr;
  int rc;
  if ((rc = nb_scratch(ctx, NB_SCR_REAL0, n3 * sizeof(float), &rho))) return rc;
  if ((rc = nb_scratch(ctx, NB_SCR_SPEC0, nspec * sizeof(float2), &s1))) return rc;
  if (interlace_up && (rc = nb_scratch(ctx, NB_SCR_SPEC1, nspec * sizeof(float2), &s2))) return rc;
  if ((rc = nb_scratch(ctx, NB_SCR_SPEC2, 3 * nspec * sizeof(float2), &g))) return rc;
  if (interlace_down && (rc = nb_scratch(ctx, NB_SCR_SPEC3, 3 * nspec * sizeof(float2), &h))) return rc;
  cufftHandle fwd, inv;
  if ((rc = nb_plan3d(ctx, N, CUFFT_R2C, st, &fwd))) return rc;
  if ((rc = nb_plan3d(ctx, N, CUFFT_C2R, st, &inv))) return rc;
  if (!rho1_in) {
    NbPhase ph(ctx, NB_PH_DEPOSIT, st);
    if ((rc = nb_cic_deposit(ctx, pos, np, cell, N, rho, NB_F32, 0.0, 0, nullptr, stream))) return rc;
  }
  {
    NbPhase ph(ctx, NB_PH_FFT_FWD, st);
    NB_CHECK_CUFFT(cufftExecR2C(fwd, (cufftReal*)(rho1_in ? const_cast<float*>(rho1_in) : (float*)rho), (cufftComplex*)s1));
  }
  if (interlace_up) {
    if (!rho2_in) {
      NbPhase ph(ctx, NB_PH_DEPOSIT, st);
      if ((rc = nb_cic_deposit(ctx, pos, np, cell, N, rho, NB_F32, 0.5 * cell, 0, nullptr, stream))) return rc;
    }
    NbPhase ph(ctx, NB_PH_FFT_FWD, st);
    NB_CHECK_CUFFT(cufftExecR2C(fwd, (cufftReal*)(rho2_in ? const_cast<float*>(rho2_in) : (float*)rho), (cufftComplex*)s2));
  }
  {
    NbPhase ph(ctx, NB_PH_KSPACE, st);
    if ((rc = nb_kspace_fused(ctx, s1, interlace_up ? s2 : nullptr, N, k_vals, sinc_tab, r_s, deconv_order, pref, g,
                              interlace_down ? h : nullptr, stream)))
      return rc;
  }
  NbPhase ph_inv(ctx, NB_PH_FFT_INV, st);
  for (int c = 0; c < 3; ++c)
    if ((rc = nb_exec_c2r(ctx, inv, NB_F32, (cufftComplex*)g + c * nspec, force3 + c * n3, n3 * sizeof(float), st)))
      return rc;
  if (interlace_down)
    for (int c = 0; c < 3; ++c)
      if ((rc = nb_exec_c2r(ctx, inv, NB_F32, (cufftComplex*)h + c * nspec, force3_shifted + c * n3,
                            n3 * sizeof(float), st)))
        return rc;
  return NB_OK;
}

extern "C" int nb_pm_force_mesh(nb_ctx* ctx, const double* pos, int64_t np, double cell, int N,
                                const double* k_vals, const double* sinc_tab, double r_s, int deconv_order,
                                int interlace_up, int interlace_down, double pref, float* force3,
                                float* force3_shifted, void* stream) {
  NB_REQUIRE(pos != nullptr, "NULL argument");
  return pm_force_mesh_impl(ctx, pos, np, cell, N, k_vals, sinc_tab, r_s, deconv_order, interlace_up, interlace_down,
                            pref, force3, force3_shifted, nullptr, nullptr, stream);
}

// The same solve from meshes the caller has deposited (nb_cic_deposit, float32, counts; rho_shifted = the deposit
// with shift = cell/2, or NULL without upstream interlacing): a host-buffer step deposits each chunk of positions
// as it arrives, while the rest is still on the link.  The meshes are read, not changed.
extern "C" int nb_pm_force_mesh_rho(nb_ctx* ctx, const float* rho, const float* rho_shifted, int N,
                                    const double* k_vals, const double* sinc_tab, double r_s, int deconv_order,
                                    int interlace_down, double pref, float* force3, float* force3_shifted,
                                    void* stream) {
  NB_REQUIRE(rho != nullptr, "NULL argument");
  return pm_force_mesh_impl(ctx, nullptr, 0, 0.0, N, k_vals, sinc_tab, r_s, deconv_order, rho_shifted != nullptr,
                            interlace_down, pref, force3, force3_shifted, rho, rho_shifted, stream);
}
