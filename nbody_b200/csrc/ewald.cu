// Periodic PP with Ewald summation: CONCEPT's PP gravity in a periodic box is
// the softened nearest-image pair force plus a tabulated Ewald correction (the reference holds no
// source for it — only its run time, PPPMP3M/pppmp3mplotter.py:18-19 — so the published method
// is restated: Hernquist, Bouchet & Suto 1991, ApJS 75, 231).
//
//   F_per(d) =  Σ_n  (d - nL)/|d - nL|³ · [erfc(α|d - nL|) + 2α|d - nL|/√π · exp(-α²|d - nL|²)]
//             + 4π/L³ Σ_{h≠0} h/|h|² · exp(-|h|²/4α²) · sin(h·d),          h = 2π m / L
//   C(d)     =  F_per(d) - d/|d|³          (d = nearest image; C is smooth, C ≈ -4π d / 3L³ at 0)
//
// nb_ewald_table tabulates C on the octant [0, L/2]³ ((g+1)³ nodes; C_x is odd in d_x and even in
// d_y, d_z, likewise for the other components).  nb_pp_accel_ewald evaluates, per pair, the Plummer
// nearest-image force in FP32 plus the trilinear interpolation of the table (8 float4 loads that hit
// L1/L2: the 65³ table is 4.4 MB), FP32 within a j-tile and FP64 across tiles like nb_pp_accel.
#include "common.cuh"

#define EW_THREADS 128
#define EW_TJ 256

__global__ void __launch_bounds__(128)
ewald_table_kernel(double box, int g, double alpha, int nreal, int nfour, double* __restrict__ tab64,
                   float4* __restrict__ tab4) {
  const int g1 = g + 1;
  const int64_t total = (int64_t)g1 * g1 * g1;
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= total) return;
  const int iz = (int)(t % g1), iy = (int)((t / g1) % g1), ix = (int)(t / ((int64_t)g1 * g1));
  const double h = 0.5 * box / g;
  const double dx = ix * h, dy = iy * h, dz = iz * h;
  const double two_a_sqrtpi = 2.0 * alpha * 0.56418958354775628695;   // 2α/√π
  double fx = 0.0, fy = 0.0, fz = 0.0;
  for (int nx = -nreal; nx <= nreal; ++nx)
    for (int ny = -nreal; ny <= nreal; ++ny)
      for (int nz = -nreal; nz <= nreal; ++nz) {
        const double rx = dx - nx * box, ry = dy - ny * box, rz = dz - nz * box;
        const double r2 = rx * rx + ry * ry + rz * rz;
        if (r2 == 0.0) continue;
        const double r = sqrt(r2);
        double s = (erfc(alpha * r) + two_a_sqrtpi * r * exp(-alpha * alpha * r2)) / (r2 * r);
        if (nx == 0 && ny == 0 && nz == 0) s -= 1.0 / (r2 * r);          // minus the Newtonian nearest image
        fx += s * rx; fy += s * ry; fz += s * rz;
      }
  const double two_pi_L = 6.283185307179586476925 / box;
  const double pref = 2.0 * two_pi_L / (box * box);                    // 4π/L³
  const double inv4a2 = 0.25 / (alpha * alpha);
  for (int mx = -nfour; mx <= nfour; ++mx)
    for (int my = -nfour; my <= nfour; ++my)
      for (int mz = -nfour; mz <= nfour; ++mz) {
        if (mx == 0 && my == 0 && mz == 0) continue;
        const double hx = mx * two_pi_L, hy = my * two_pi_L, hz = mz * two_pi_L;
        const double h2 = hx * hx + hy * hy + hz * hz;
        const double s = pref * exp(-h2 * inv4a2) / h2 * sin(hx * dx + hy * dy + hz * dz);
        fx += s * hx; fy += s * hy; fz += s * hz;
      }
  // exact zeros the symmetry demands: the odd component on the plane through the origin (on the far
  // face d = L/2 the PERIODIC force is zero, so the correction there is minus the nearest-image term)
  if (ix == 0) fx = 0.0;
  if (iy == 0) fy = 0.0;
  if (iz == 0) fz = 0.0;
  if (tab64) { tab64[3 * t] = fx; tab64[3 * t + 1] = fy; tab64[3 * t + 2] = fz; }
  if (tab4) tab4[t] = make_float4((float)fx, (float)fy, (float)fz, 0.f);
}

__device__ __forceinline__ float4 ew_lerp(const float4 a, const float4 b, float t) {
  return make_float4(fmaf(t, b.x - a.x, a.x), fmaf(t, b.y - a.y, a.y), fmaf(t, b.z - a.z, a.z), 0.f);
}

__global__ void __launch_bounds__(EW_THREADS)
pp_ewald_kernel(const float4* __restrict__ ipos, int64_t ni, const float4* __restrict__ jpos, int64_t nj,
                int64_t j_chunk, float eps2, float box, float inv_box, float to_grid, int g,
                const float4* __restrict__ tab, double scale, double* __restrict__ acc) {
  __shared__ float4 sj[EW_TJ];
  const int64_t i = (int64_t)blockIdx.x * EW_THREADS + threadIdx.x;
  const bool live = i < ni;
  const float4 pi = live ? ipos[i] : make_float4(0.f, 0.f, 0.f, 0.f);
  const int64_t j_lo = (int64_t)blockIdx.y * j_chunk;
  const int64_t j_hi = min(nj, j_lo + j_chunk);
  const int g1 = g + 1;
  double ax = 0.0, ay = 0.0, az = 0.0;
  for (int64_t j0 = j_lo; j0 < j_hi; j0 += EW_TJ) {
    const int cnt = (int)min((int64_t)EW_TJ, j_hi - j0);
    __syncthreads();
    for (int t = threadIdx.x; t < cnt; t += EW_THREADS) sj[t] = jpos[j0 + t];
    __syncthreads();
    if (!live) continue;
    float fx = 0.f, fy = 0.f, fz = 0.f;
    for (int t = 0; t < cnt; ++t) {
      const float4 pj = sj[t];
      float dx = pj.x - pi.x, dy = pj.y - pi.y, dz = pj.z - pi.z;
      dx = fmaf(-box, rintf(dx * inv_box), dx);
      dy = fmaf(-box, rintf(dy * inv_box), dy);
      dz = fmaf(-box, rintf(dz * inv_box), dz);
      const float r2 = fmaf(dx, dx, fmaf(dy, dy, fmaf(dz, dz, eps2)));
      float inv = r2 > 0.f ? rsqrtf(r2) : 0.f;
      float s = inv * inv * inv;
      // Ewald correction: trilinear in (|dx|, |dy|, |dz|) on the octant table, signs restored
      const float ux = fminf(fabsf(dx) * to_grid, (float)g), uy = fminf(fabsf(dy) * to_grid, (float)g),
                  uz = fminf(fabsf(dz) * to_grid, (float)g);
      const int cx = min((int)ux, g - 1), cy = min((int)uy, g - 1), cz = min((int)uz, g - 1);
      const float tx = ux - (float)cx, ty = uy - (float)cy, tz = uz - (float)cz;
      const float4* b = tab + ((int64_t)cx * g1 + cy) * g1 + cz;
      const int sy = g1, sx = g1 * g1;
      const float4 c00 = ew_lerp(__ldg(b), __ldg(b + 1), tz);
      const float4 c01 = ew_lerp(__ldg(b + sy), __ldg(b + sy + 1), tz);
      const float4 c10 = ew_lerp(__ldg(b + sx), __ldg(b + sx + 1), tz);
      const float4 c11 = ew_lerp(__ldg(b + sx + sy), __ldg(b + sx + sy + 1), tz);
      const float4 c = ew_lerp(ew_lerp(c00, c01, ty), ew_lerp(c10, c11, ty), tx);
      fx = fmaf(pj.w, fmaf(s, dx, dx < 0.f ? -c.x : c.x), fx);
      fy = fmaf(pj.w, fmaf(s, dy, dy < 0.f ? -c.y : c.y), fy);
      fz = fmaf(pj.w, fmaf(s, dz, dz < 0.f ? -c.z : c.z), fz);
    }
    ax += (double)fx; ay += (double)fy; az += (double)fz;
  }
  if (live) {
    if (gridDim.y == 1) {
      acc[3 * i] += scale * ax; acc[3 * i + 1] += scale * ay; acc[3 * i + 2] += scale * az;
    } else {
      atomicAdd(&acc[3 * i], scale * ax); atomicAdd(&acc[3 * i + 1], scale * ay); atomicAdd(&acc[3 * i + 2], scale * az);
    }
  }
}

extern "C" int nb_ewald_table(nb_ctx* ctx, double box, int g, double alpha, int nreal, int nfour,
                              double* table_f64, void* table_f4, void* stream) {
  NB_REQUIRE(ctx, "ctx is NULL");
  NB_REQUIRE(box > 0 && g >= 2 && g <= 512, "box > 0 and 2 <= g <= 512");
  NB_REQUIRE(alpha > 0 && nreal >= 1 && nreal <= 8 && nfour >= 1 && nfour <= 12, "alpha > 0, 1 <= nreal <= 8, 1 <= nfour <= 12");
  NB_REQUIRE(table_f64 || table_f4, "no output table");
  cudaStream_t st = (cudaStream_t)stream;
  NB_CHECK_CUDA(cudaSetDevice(ctx->device));
  const int64_t total = (int64_t)(g + 1) * (g + 1) * (g + 1);
  ewald_table_kernel<<<(unsigned)((total + 127) / 128), 128, 0, st>>>(box, g, alpha, nreal, nfour, table_f64,
                                                                      (float4*)table_f4);
  NB_LAUNCHED(ctx);
  return NB_OK;
}

extern "C" int nb_pp_accel_ewald(nb_ctx* ctx, const void* ipos_xyzw, int64_t ni, const void* jpos_xyzw, int64_t nj,
                                 double eps, double box, double scale, const void* table_f4, int g,
                                 int accumulate, double* acc, void* stream) {
  NB_REQUIRE(ctx, "ctx is NULL");
  if (ni == 0) return NB_OK;
  NB_REQUIRE(acc && ipos_xyzw && (nj == 0 || jpos_xyzw) && table_f4, "NULL argument");
  NB_REQUIRE(eps >= 0 && box > 0 && g >= 2, "eps >= 0, box > 0, g >= 2");
  cudaStream_t st = (cudaStream_t)stream;
  NB_CHECK_CUDA(cudaSetDevice(ctx->device));
  if (!accumulate) NB_CHECK_CUDA(cudaMemsetAsync(acc, 0, sizeof(double) * 3 * ni, st));
  if (nj == 0) return NB_OK;
  const int64_t nbi = (ni + EW_THREADS - 1) / EW_THREADS;
  int64_t jsplit = 1;
  const int64_t want = (int64_t)ctx->num_sms * 8;
  if (nbi < want) jsplit = (want + nbi - 1) / nbi;
  int64_t max_split = (nj + 4 * EW_TJ - 1) / (4 * EW_TJ);
  if (max_split < 1) max_split = 1;
  if (jsplit > max_split) jsplit = max_split;
  if (jsplit > 65535) jsplit = 65535;
  int64_t j_chunk = (nj + jsplit - 1) / jsplit;
  j_chunk = ((j_chunk + EW_TJ - 1) / EW_TJ) * EW_TJ;
  jsplit = (nj + j_chunk - 1) / j_chunk;
  dim3 grid((unsigned)nbi, (unsigned)jsplit);
  pp_ewald_kernel<<<grid, EW_THREADS, 0, st>>>((const float4*)ipos_xyzw, ni, (const float4*)jpos_xyzw, nj, j_chunk,
                                               (float)(eps * eps), (float)box, (float)(1.0 / box),
                                               (float)(2.0 * g / box), g, (const float4*)table_f4, scale, acc);
  NB_LAUNCHED(ctx);
  return NB_OK;
}
