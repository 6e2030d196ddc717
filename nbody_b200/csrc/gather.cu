// K3 — trilinear (CIC) gather, displacement + wrap, and the fused gather + kick + drift.
// The gather is the adjoint of K1 and uses the same float64 index / weight math
// (FraP_kTilDelta.py:99-107), so deposit and gather address identical cells.
// Reference call sites: RegularGridInterpolator at lattice nodes + `particles += ψ;
// particles %= L` (Power_compare.py:356-391, CiC/3D startbetingelser1.py:137-173).
#include "common.cuh"

#ifndef GKD_MIN_CTAS
#define GKD_MIN_CTAS 3   // fused gather+kick+drift tail at C5: no bound 14.0 ms, 4 CTAs (64 regs, 56 B spills) 10.4 ms, 3 CTAs (80 regs) 8.4 ms, 5 CTAs 15.2 ms
#endif

struct CicStencil {
  int64_t r00, r01, r10, r11;
  int k0, k1;
  double w000, w001, w010, w011, w100, w101, w110, w111;
};

// x0 / nxp: first mesh plane and number of planes of the field buffer ([nxp][N][N]); the full
// mesh has x0 = 0, nxp = N, a slab has nxp = own + ghost planes.  A stencil that does not fit
// the buffer (a particle this slab does not own) gets zero weights on plane 0: no out-of-range
// access, contributes nothing.
template <bool EXACT>
__device__ __forceinline__ CicStencil cic_stencil(double x, double y, double z, double cell, double inv_cell, int N,
                                                  int x0 = 0, int nxp = 0x7fffffff) {
  CicStencil s;
  int i0, j0;
  double dx, dy, dz;
  nb_cell_split<EXACT>(x, cell, inv_cell, N, i0, dx);
  nb_cell_split<EXACT>(y, cell, inv_cell, N, j0, dy);
  nb_cell_split<EXACT>(z, cell, inv_cell, N, s.k0, dz);
  const int j1 = nb_wrap1(j0, N);
  s.k1 = nb_wrap1(s.k0, N);
  int pa = i0 - x0, pb = nb_wrap1(i0, N) - x0;
  pa = pa < 0 ? pa + N : pa;
  pb = pb < 0 ? pb + N : pb;
  const bool inside = pa < nxp && pb < nxp;
  if (!inside) pa = pb = 0;
  const double wx0 = __dsub_rn(1.0, dx), wy0 = __dsub_rn(1.0, dy), wz0 = __dsub_rn(1.0, dz);
  const double w00 = __dmul_rn(wx0, wy0), w01 = __dmul_rn(wx0, dy);
  const double w10 = __dmul_rn(dx, wy0), w11 = __dmul_rn(dx, dy);
  s.w000 = __dmul_rn(w00, wz0); s.w001 = __dmul_rn(w00, dz);
  s.w010 = __dmul_rn(w01, wz0); s.w011 = __dmul_rn(w01, dz);
  s.w100 = __dmul_rn(w10, wz0); s.w101 = __dmul_rn(w10, dz);
  s.w110 = __dmul_rn(w11, wz0); s.w111 = __dmul_rn(w11, dz);
  s.r00 = ((int64_t)pa * N + j0) * N; s.r01 = ((int64_t)pa * N + j1) * N;
  s.r10 = ((int64_t)pb * N + j0) * N; s.r11 = ((int64_t)pb * N + j1) * N;
  if (!inside) s.w000 = s.w001 = s.w010 = s.w011 = s.w100 = s.w101 = s.w110 = s.w111 = 0.0;
  return s;
}

template <typename T>
__device__ __forceinline__ double cic_interp(const T* __restrict__ f, const CicStencil& s) {
  // same accumulation order as the oracle's trilinear_gather (di, dj, dk lexicographic)
  double v = s.w000 * (double)__ldg(f + s.r00 + s.k0);
  v += s.w001 * (double)__ldg(f + s.r00 + s.k1);
  v += s.w010 * (double)__ldg(f + s.r01 + s.k0);
  v += s.w011 * (double)__ldg(f + s.r01 + s.k1);
  v += s.w100 * (double)__ldg(f + s.r10 + s.k0);
  v += s.w101 * (double)__ldg(f + s.r10 + s.k1);
  v += s.w110 * (double)__ldg(f + s.r11 + s.k0);
  v += s.w111 * (double)__ldg(f + s.r11 + s.k1);
  return v;
}

template <typename T>
__global__ void __launch_bounds__(256)
cic_gather_kernel(const T* __restrict__ field, int ncomp, int N, const double* __restrict__ pos, int64_t np,
                  double cell, double shift, double* __restrict__ out, int accumulate, double scale, int lat_n,
                  int lat_yb) {
  const int64_t n3 = (int64_t)N * N * N;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const double inv_cell = __ddiv_rn(1.0, cell);
  for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < np; q += stride) {
    const int64_t p = nb_blocked_order(q, lat_n, lat_yb);
    double x = pos[3 * p], y = pos[3 * p + 1], z = pos[3 * p + 2];
    if (shift != 0.0) { x = __dadd_rn(x, shift); y = __dadd_rn(y, shift); z = __dadd_rn(z, shift); }
    const CicStencil s = cic_stencil<sizeof(T) == 8>(x, y, z, cell, inv_cell, N);
    for (int c = 0; c < ncomp; ++c) {
      const double v = cic_interp(field + c * n3, s) * scale;
      if (accumulate) out[p * ncomp + c] += v; else out[p * ncomp + c] = v;
    }
  }
}

// pos <- pymod(pos + disp, L)     (`particles += ψ ; particles %= L`)
__global__ void __launch_bounds__(256)
displace_wrap_kernel(double* __restrict__ pos, const double* __restrict__ disp, int64_t n3, double L) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < n3; t += stride)
    pos[t] = nb_pymod_near(__dadd_rn(pos[t], disp[t]), L);
}

// mom += kick * acc ; pos = pymod(pos + drift * mom, L)
__global__ void __launch_bounds__(256)
kick_drift_kernel(double* __restrict__ pos, double* __restrict__ mom, const double* __restrict__ acc, int64_t n3,
                  double kick, double drift, double L) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < n3; t += stride) {
    const double m = mom[t] + kick * acc[t];
    mom[t] = m;
    pos[t] = nb_pymod_near(pos[t] + drift * m, L);
  }
}

// Fused PM tail: gather the 3 force meshes (optionally a second, interlace-shifted set,
// averaged), kick, drift, wrap.  One pass over the particle stream.
__global__ void __launch_bounds__(256, GKD_MIN_CTAS)
gather_kick_drift_kernel(const float* __restrict__ f1, const float* __restrict__ f2, int N, int x0, int nxp,
                         double cell, double shift2, double* __restrict__ pos, double* __restrict__ mom,
                         const double* __restrict__ acc_add, double* __restrict__ acc_out, int64_t np,
                         double gscale, double kick, double drift, double L, int lat_n, int lat_yb) {
  const int64_t n3 = (int64_t)nxp * N * N;   // component stride of the field buffers
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const double inv_cell = __ddiv_rn(1.0, cell);
  for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < np; q += stride) {
    const int64_t p = nb_blocked_order(q, lat_n, lat_yb);
    const double x = pos[3 * p], y = pos[3 * p + 1], z = pos[3 * p + 2];
    const CicStencil s = cic_stencil<false>(x, y, z, cell, inv_cell, N, x0, nxp);
    double g[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) g[c] = cic_interp(f1 + c * n3, s);
    if (f2) {
      const CicStencil s2 = cic_stencil<false>(__dadd_rn(x, shift2), __dadd_rn(y, shift2), __dadd_rn(z, shift2),
                                               cell, inv_cell, N, x0, nxp);
#pragma unroll
      for (int c = 0; c < 3; ++c) g[c] = 0.5 * (g[c] + cic_interp(f2 + c * n3, s2));
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) g[c] *= gscale;
    if (acc_add) {   // e.g. the P3M short-range part computed concurrently on another stream
      g[0] += acc_add[3 * p]; g[1] += acc_add[3 * p + 1]; g[2] += acc_add[3 * p + 2];
    }
    if (acc_out) {
      acc_out[3 * p] = g[0]; acc_out[3 * p + 1] = g[1]; acc_out[3 * p + 2] = g[2];
    }
    if (mom) {
      const double px[3] = {x, y, z};
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const double m = mom[3 * p + c] + kick * g[c];
        mom[3 * p + c] = m;
        pos[3 * p + c] = nb_pymod_near(px[c] + drift * m, L);
      }
    }
  }
}

extern "C" int nb_cic_gather(nb_ctx* ctx, const void* field, int ncomp, int N, int prec, const double* pos,
                             int64_t np, double cell, double shift, double scale, int accumulate, double* out,
                             void* stream) {
  NB_REQUIRE(ctx, "ctx is NULL");
  NB_REQUIRE(ncomp >= 1 && N >= 1 && cell > 0, "bad ncomp / N / cell");
  NB_REQUIRE(prec == NB_F32 || prec == NB_F64, "bad prec");
  if (np == 0) return NB_OK;
  NB_REQUIRE(field && out && pos, "NULL argument");
  cudaStream_t st = (cudaStream_t)stream;
  NB_CHECK_CUDA(cudaSetDevice(ctx->device));
  const int nblk = nb_grid_for(ctx, np, 256, 16);
  int lat_n = 0;
  const int lat_yb = nb_blocked_enabled() ? nb_lattice_yblock(np, &lat_n) : 0;
  if (prec == NB_F64)
    cic_gather_kernel<double><<<nblk, 256, 0, st>>>((const double*)field, ncomp, N, pos, np, cell, shift, out,
                                                    accumulate, scale, lat_n, lat_yb);
  else
    cic_gather_kernel<float><<<nblk, 256, 0, st>>>((const float*)field, ncomp, N, pos, np, cell, shift, out,
                                                   accumulate, scale, lat_n, lat_yb);
  NB_LAUNCHED(ctx);
  return NB_OK;
}

extern "C" int nb_displace_wrap(nb_ctx* ctx, double* pos, const double* disp, int64_t np, double L, void* stream) {
  NB_REQUIRE(ctx && (np == 0 || (pos && disp)), "NULL argument");
  NB_REQUIRE(L > 0, "L must be positive");
  if (np == 0) return NB_OK;
  cudaStream_t st = (cudaStream_t)stream;
  NB_CHECK_CUDA(cudaSetDevice(ctx->device));
  displace_wrap_kernel<<<nb_grid_for(ctx, 3 * np, 256, 16), 256, 0, st>>>(pos, disp, 3 * np, L);
  NB_LAUNCHED(ctx);
  return NB_OK;
}

extern "C" int nb_kick_drift(nb_ctx* ctx, double* pos, double* mom, const double* acc, int64_t np, double kick,
                             double drift, double L, void* stream) {
  NB_REQUIRE(ctx && (np == 0 || (pos && mom && acc)), "NULL argument");
  NB_REQUIRE(L > 0, "L must be positive");
  if (np == 0) return NB_OK;
  cudaStream_t st = (cudaStream_t)stream;
  NB_CHECK_CUDA(cudaSetDevice(ctx->device));
  kick_drift_kernel<<<nb_grid_for(ctx, 3 * np, 256, 16), 256, 0, st>>>(pos, mom, acc, 3 * np, kick, drift, L);
  NB_LAUNCHED(ctx);
  return NB_OK;
}

static int gkd_launch(nb_ctx* ctx, const float* force3, const float* force3_shifted, int N, int x0, int nxp,
                      double cell, double* pos, double* mom, const double* acc_add, double* acc_out, int64_t np,
                      double gscale, double kick, double drift, double L, cudaStream_t st) {
  NB_REQUIRE(ctx && force3 && (np == 0 || pos), "NULL argument");
  NB_REQUIRE(N >= 1 && cell > 0 && L > 0, "bad N / cell / L");
  NB_REQUIRE(mom || acc_out, "nothing to do: give mom (kick+drift) and/or acc_out");
  if (np == 0) return NB_OK;
  NB_CHECK_CUDA(cudaSetDevice(ctx->device));
  int lat_n = 0;
  const int lat_yb = nb_blocked_enabled() ? nb_lattice_yblock(np, &lat_n) : 0;
  gather_kick_drift_kernel<<<nb_grid_for(ctx, np, 256, 16), 256, 0, st>>>(
      force3, force3_shifted, N, x0, nxp, cell, 0.5 * cell, pos, mom, acc_add, acc_out, np, gscale, kick, drift, L,
      lat_n, lat_yb);
  NB_LAUNCHED(ctx);
  return NB_OK;
}

extern "C" int nb_gather_kick_drift(nb_ctx* ctx, const float* force3, const float* force3_shifted, int N,
                                    double cell, double* pos, double* mom, double* acc_out, int64_t np,
                                    double gscale, double kick, double drift, double L, void* stream) {
  return gkd_launch(ctx, force3, force3_shifted, N, 0, N, cell, pos, mom, nullptr, acc_out, np, gscale, kick, drift,
                    L, (cudaStream_t)stream);
}

// Slab form: force3 / force3_shifted are [3][nx_own + n_ghost][N][N] and hold the mesh planes
// x0 .. x0 + nx_own + n_ghost - 1 (mod N); the ghost planes are copies of the next slab's
// first planes (1 needed, 2 with the half-cell interlace shift).
extern "C" int nb_gather_kick_drift_slab(nb_ctx* ctx, const float* force3, const float* force3_shifted, int N,
                                         int x0, int nx_own, int n_ghost, double cell, double* pos, double* mom,
                                         double* acc_out, int64_t np, double gscale, double kick, double drift,
                                         double L, void* stream) {
  NB_REQUIRE(x0 >= 0 && x0 < N && nx_own >= 1 && nx_own <= N && n_ghost >= 0, "bad slab");
  return gkd_launch(ctx, force3, force3_shifted, N, x0, nx_own + n_ghost, cell, pos, mom, nullptr, acc_out, np,
                    gscale, kick, drift, L, (cudaStream_t)stream);
}

// g = gscale · gather + acc_add[p]: the fused tail of a P3M step whose short-range part was
// computed concurrently (another stream) into acc_add.  Full mesh: x0 = 0, nx_own = N, n_ghost = 0.
extern "C" int nb_gather_kick_drift_add(nb_ctx* ctx, const float* force3, const float* force3_shifted, int N,
                                        int x0, int nx_own, int n_ghost, double cell, double* pos, double* mom,
                                        const double* acc_add, double* acc_out, int64_t np, double gscale,
                                        double kick, double drift, double L, void* stream) {
  NB_REQUIRE(x0 >= 0 && x0 < N && nx_own >= 1 && nx_own <= N && n_ghost >= 0, "bad slab");
  return gkd_launch(ctx, force3, force3_shifted, N, x0, nx_own + n_ghost, cell, pos, mom, acc_add, acc_out, np,
                    gscale, kick, drift, L, (cudaStream_t)stream);
}
