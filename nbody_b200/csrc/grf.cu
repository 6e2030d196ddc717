// K7 — Gaussian-random-field synthesis from P(k).
// Reference: createR  Cosmo data create (new).py:24-47 (no `if` guard),
//            inline   CiC/3D startbetingelser1.py:82-93, Power_compare.py:301-312 (guarded);
//            find_delta  Cosmo data create (new).py:114-137;
//            δ and ψ     CiC/3D startbetingelser1.py:98-112.
//
// The reference builds the full N³ complex array F(k) = op(k)·sqrt(P(k))·fac·R(k) and takes
// ifftn(F).real.  ifftn(F).real == C2R(H) with H(k) = (F(k) + conj(F(-k)))/2, the Hermitian
// part, for ANY F — so one kernel evaluates F at k and at -k on the R2C half spectrum,
// forms H and a single C2R per field replaces the complex transform.  This also covers the
// places where F is not Hermitian in the reference (self-conjugate points left complex by
// the guarded loop; the Nyquist planes of ψ).
#include "common.cuh"

struct Cplx { double x, y; };

// R after the reference's lexicographic loop R[-i,-j,-k] = conj(R[i,j,k]).
__device__ __forceinline__ Cplx grf_R(const double* __restrict__ c1, const double* __restrict__ c2,
                                      int64_t lin, int64_t lins, int guard) {
  Cplx r;
  if (lin < lins) {            // visited first: keeps its draw
    r.x = c1[lin];
    r.y = c2[lin];
  } else if (lin > lins) {     // overwritten by the conjugate of its partner's draw
    r.x = c1[lins];
    r.y = -c2[lins];
  } else {                     // self-conjugate point
    r.x = c1[lin];
    r.y = (lin == 0) ? 0.0 : (guard ? c2[lin] : -c2[lin]);
  }
  return r;
}

__global__ void __launch_bounds__(256)
hermitian_R_kernel(const double* __restrict__ c1, const double* __restrict__ c2, int N, int guard,
                   double2* __restrict__ out) {
  const int64_t total = (int64_t)N * N * N;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += stride) {
    const int k = (int)(t % N);
    const int64_t r = t / N;
    const int j = (int)(r % N), i = (int)(r / N);
    const int64_t lins = ((int64_t)((N - i) % N) * N + (N - j) % N) * N + (N - k) % N;
    const Cplx v = grf_R(c1, c2, t, lins, guard);
    out[t] = make_double2(v.x, v.y);
  }
}

template <typename C, typename T>
__global__ void __launch_bounds__(256)
grf_kernel(const double* __restrict__ pk, const double* __restrict__ c1, const double* __restrict__ c2, int N,
           const double* __restrict__ kv, double fac, int guard, double inv_n3, C* __restrict__ dspec,
           C* __restrict__ p0, C* __restrict__ p1, C* __restrict__ p2) {
  const int nzh = N / 2 + 1;
  const int64_t total = (int64_t)N * N * nzh;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += stride) {
    const int k = (int)(t % nzh);
    const int64_t r = t / nzh;
    const int j = (int)(r % N), i = (int)(r / N);
    const int is = (N - i) % N, js = (N - j) % N, ks = (N - k) % N;
    const int64_t la = ((int64_t)i * N + j) * N + k;
    const int64_t lb = ((int64_t)is * N + js) * N + ks;
    const Cplx Ra = grf_R(c1, c2, la, lb, guard);
    const Cplx Rb = grf_R(c1, c2, lb, la, guard);
    const double da = __dmul_rn(sqrt(pk[la]), fac);   // delta_k = sqrt(P) * fac
    const double db = __dmul_rn(sqrt(pk[lb]), fac);
    const int64_t o = t;
    if (dspec) {
      // H = (F(a) + conj(F(b)))/2 with F = delta_k * R
      const double hx = 0.5 * (da * Ra.x + db * Rb.x) * inv_n3;
      const double hy = 0.5 * (da * Ra.y - db * Rb.y) * inv_n3;
      dspec[o].x = (T)hx;
      dspec[o].y = (T)hy;
    }
    if (p0) {
      const double kxa = kv[i], kya = kv[j], kza = kv[k];
      double km = sqrt(__dadd_rn(__dadd_rn(__dmul_rn(kxa, kxa), __dmul_rn(kya, kya)), __dmul_rn(kza, kza)));
      if (la == 0) km = 1e-10;                          // k_mag[0,0,0] = 1e-10 (CiC/...:72)
      const double km2 = __dmul_rn(km, km);
      const double kxb = kv[is], kyb = kv[js], kzb = kv[ks];
      C* outs[3] = {p0, p1, p2};
      const double kas[3] = {kxa, kya, kza};
      const double kbs[3] = {kxb, kyb, kzb};
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        // F = (1j*k_c/k_mag**2*delta_k) * R  ->  (-q*R.y, q*R.x)
        const double qa = __dmul_rn(__ddiv_rn(kas[c], km2), da);
        const double qb = __dmul_rn(__ddiv_rn(kbs[c], km2), db);
        const double fax = -qa * Ra.y, fay = qa * Ra.x;
        const double fbx = -qb * Rb.y, fby = qb * Rb.x;
        outs[c][o].x = (T)(0.5 * (fax + fbx) * inv_n3);
        outs[c][o].y = (T)(0.5 * (fay - fby) * inv_n3);
      }
    }
  }
}

extern "C" int nb_hermitian_R(nb_ctx* ctx, const double* c1, const double* c2, int N, int guard_selfconj,
                              double* R_out, void* stream) {
  NB_REQUIRE(ctx && c1 && c2 && R_out, "NULL argument");
  NB_REQUIRE(N >= 1, "bad N");
  cudaStream_t st = (cudaStream_t)stream;
  NB_CHECK_CUDA(cudaSetDevice(ctx->device));
  const int nblk = nb_grid_for(ctx, (int64_t)N * N * N, 256, 8);
  hermitian_R_kernel<<<nblk, 256, 0, st>>>(c1, c2, N, guard_selfconj, (double2*)R_out);
  NB_LAUNCHED(ctx);
  return NB_OK;
}

extern "C" int nb_grf_synthesize(nb_ctx* ctx, const double* pk_grid, const double* c1, const double* c2, int N,
                                 const double* k_vals, double fac, int guard_selfconj, int prec,
                                 void* delta_out, void* psi_out, void* stream) {
  NB_REQUIRE(ctx && pk_grid && c1 && c2 && k_vals, "NULL argument");
  NB_REQUIRE(delta_out || psi_out, "nothing to compute");
  NB_REQUIRE(N >= 2, "bad N");
  NB_REQUIRE(prec == NB_F32 || prec == NB_F64, "bad prec");
  cudaStream_t st = (cudaStream_t)stream;
  NB_CHECK_CUDA(cudaSetDevice(ctx->device));
  const int64_t nspec = (int64_t)N * N * (N / 2 + 1);
  const size_t cb = prec == NB_F64 ? 16 : 8, rb = cb / 2;
  void* spec[4] = {nullptr, nullptr, nullptr, nullptr};
  int rc;
  if (delta_out && (rc = nb_scratch(ctx, NB_SCR_SPEC0, nspec * cb, &spec[0]))) return rc;
  if (psi_out) {
    for (int c = 0; c < 3; ++c)
      if ((rc = nb_scratch(ctx, NB_SCR_SPEC1 + c, nspec * cb, &spec[1 + c]))) return rc;
  }
  const double inv_n3 = 1.0 / ((double)N * N * N);
  const int nblk = nb_grid_for(ctx, nspec, 256, 8);
  cufftHandle inv;
  if (prec == NB_F64) {
    grf_kernel<double2, double><<<nblk, 256, 0, st>>>(pk_grid, c1, c2, N, k_vals, fac, guard_selfconj, inv_n3,
                                                     (double2*)spec[0], (double2*)spec[1], (double2*)spec[2],
                                                     (double2*)spec[3]);
    NB_LAUNCHED(ctx);
    if ((rc = nb_plan3d(ctx, N, CUFFT_Z2D, st, &inv))) return rc;
    if (delta_out && (rc = nb_exec_c2r(ctx, inv, prec, spec[0], delta_out, (size_t)N * N * N * rb, st))) return rc;
    if (psi_out)
      for (int c = 0; c < 3; ++c)
        if ((rc = nb_exec_c2r(ctx, inv, prec, spec[1 + c], (char*)psi_out + (size_t)c * N * N * N * rb,
                              (size_t)N * N * N * rb, st)))
          return rc;
  } else {
    grf_kernel<float2, float><<<nblk, 256, 0, st>>>(pk_grid, c1, c2, N, k_vals, fac, guard_selfconj, inv_n3,
                                                   (float2*)spec[0], (float2*)spec[1], (float2*)spec[2],
                                                   (float2*)spec[3]);
    NB_LAUNCHED(ctx);
    if ((rc = nb_plan3d(ctx, N, CUFFT_C2R, st, &inv))) return rc;
    if (delta_out && (rc = nb_exec_c2r(ctx, inv, prec, spec[0], delta_out, (size_t)N * N * N * rb, st))) return rc;
    if (psi_out)
      for (int c = 0; c < 3; ++c)
        if ((rc = nb_exec_c2r(ctx, inv, prec, spec[1 + c], (char*)psi_out + (size_t)c * N * N * N * rb,
                              (size_t)N * N * N * rb, st)))
          return rc;
  }
  return NB_OK;
}

// ---------------------------------------------------------------------------------------------------
// Second-order (2LPT) displacement of the lattice ICs (`realization_options['LPT'] = 2`,
// Different_Sizes/Giant.param:70-72; the reference compares both orders in 1LPTvs2LPT/1vs2tlptplotter.py):
//   ∇²φ1 = δ,   ∇²φ2 = Σ_{a>b} [φ1,aa φ1,bb − (φ1,ab)²],   ψ2 = ∇φ2
// float64 throughout: one D2Z of δ, six Z2D of the second derivatives, one element-wise source kernel, one
// D2Z of the source and three Z2D of the gradient.  Spectral factors per mode: k_a k_b / k² for φ1,ab of δ_k
// (a mixed derivative with exactly ONE index on its Nyquist plane is odd under k -> -k there, so the real part
// the numpy form takes of the complex inverse transform annihilates it — like the first-order gradient; with both
// indices on Nyquist planes the factor k_N² is even and stays) and −i k_a / k² for ψ2 of the source.
__global__ void __launch_bounds__(256)
lpt2_spectral_kernel(const double2* __restrict__ in, int N, const double* __restrict__ kv, int a, int b, int grad,
                     double inv_n3, double2* __restrict__ out) {
  const int nzh = N / 2 + 1, nyq = (N % 2 == 0) ? N / 2 : -1;
  const int64_t n = (int64_t)N * N * nzh, stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += stride) {
    const int iz = (int)(t % nzh), iy = (int)((t / nzh) % N), ix = (int)(t / ((int64_t)nzh * N));
    const int idx[3] = {ix, iy, iz};
    const double k[3] = {kv[ix], kv[iy], kv[iz]};
    const double k2 = (k[0] * k[0] + k[1] * k[1]) + k[2] * k[2];
    const double2 v = in[t];
    double2 r = make_double2(0.0, 0.0);
    if (k2 > 0.0) {
      if (grad) {                       // psi2_a = IFFT(i k_a phi2_k), phi2_k = -src_k / k²
        const double ka = idx[a] == nyq ? 0.0 : k[a];
        const double f = -ka / k2 * inv_n3;
        r = make_double2(-f * v.y, f * v.x);
      } else {                          // phi1,ab = IFFT(-k_a k_b phi1_k), phi1_k = -delta_k / k²
        double f;
        if (a == b) f = k[a] * k[a];
        else f = ((idx[a] == nyq) != (idx[b] == nyq)) ? 0.0 : k[a] * k[b];   // one Nyquist index: odd under k -> -k
        f = f / k2 * inv_n3;
        r = make_double2(f * v.x, f * v.y);
      }
    }
    out[t] = r;
  }
}

__global__ void __launch_bounds__(256)
lpt2_source_kernel(const double* d, int64_t n3, double* src) {   // src may alias the first field
  // d: six fields xx, yy, zz, xy, xz, yz
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < n3; t += stride) {
    const double xx = d[t], yy = d[n3 + t], zz = d[2 * n3 + t], xy = d[3 * n3 + t], xz = d[4 * n3 + t],
                 yz = d[5 * n3 + t];
    src[t] = xx * yy + xx * zz + yy * zz - xy * xy - xz * xz - yz * yz;
  }
}

// delta: (N,N,N) float64 (not modified); work: 6 N³ float64 of caller scratch; psi2_out: (3,N,N,N) float64
extern "C" int nb_lpt2_displacement(nb_ctx* ctx, const double* delta, int N, const double* k_vals, double* work6,
                                    double* psi2_out, void* stream) {
  NB_REQUIRE(ctx && delta && k_vals && work6 && psi2_out, "NULL argument");
  NB_REQUIRE(N >= 2, "bad N");
  cudaStream_t st = (cudaStream_t)stream;
  NB_CHECK_CUDA(cudaSetDevice(ctx->device));
  const int64_t n3 = (int64_t)N * N * N, nspec = (int64_t)N * N * (N / 2 + 1);
  void *dk = nullptr, *tmp = nullptr;
  int rc;
  if ((rc = nb_scratch(ctx, NB_SCR_SPEC0, nspec * 16, &dk))) return rc;
  if ((rc = nb_scratch(ctx, NB_SCR_SPEC1, nspec * 16, &tmp))) return rc;
  cufftHandle fwd, inv;
  if ((rc = nb_plan3d(ctx, N, CUFFT_D2Z, st, &fwd))) return rc;
  if ((rc = nb_plan3d(ctx, N, CUFFT_Z2D, st, &inv))) return rc;
  const double inv_n3 = 1.0 / (double)n3;
  const int nblk = nb_grid_for(ctx, nspec, 256, 8);
  NB_CHECK_CUFFT(cufftExecD2Z(fwd, (cufftDoubleReal*)const_cast<double*>(delta), (cufftDoubleComplex*)dk));
  const int pa[6] = {0, 1, 2, 0, 0, 1}, pb[6] = {0, 1, 2, 1, 2, 2};
  for (int c = 0; c < 6; ++c) {
    lpt2_spectral_kernel<<<nblk, 256, 0, st>>>((const double2*)dk, N, k_vals, pa[c], pb[c], 0, inv_n3, (double2*)tmp);
    NB_LAUNCHED(ctx);
    if ((rc = nb_exec_c2r(ctx, inv, NB_F64, tmp, work6 + c * n3, n3 * sizeof(double), st))) return rc;
  }
  // the source overwrites the first work field (each cell reads its six inputs before it writes)
  lpt2_source_kernel<<<nb_grid_for(ctx, n3, 256, 8), 256, 0, st>>>(work6, n3, work6);
  NB_LAUNCHED(ctx);
  NB_CHECK_CUFFT(cufftExecD2Z(fwd, (cufftDoubleReal*)work6, (cufftDoubleComplex*)dk));
  for (int c = 0; c < 3; ++c) {
    lpt2_spectral_kernel<<<nblk, 256, 0, st>>>((const double2*)dk, N, k_vals, c, c, 1, inv_n3, (double2*)tmp);
    NB_LAUNCHED(ctx);
    if ((rc = nb_exec_c2r(ctx, inv, NB_F64, tmp, psi2_out + c * n3, n3 * sizeof(double), st))) return rc;
  }
  return NB_OK;
}
