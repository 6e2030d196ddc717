// K5 — P3M short-range pair force on the cell-sorted particle set.
// No reference source (CONCEPT-internal); our specification, oracle/force.py:p3m_short_accel:
//   a_i = G Σ_{j: r<r_cut} m_j d / (r²+ε²)^{3/2} · [erfc(u) + 2u/√π e^{-u²}],  u = r/(2 r_s),
//   d = x_j - x_i (minimum image), r_s = 1.25 cells, r_cut = 4.5 r_s by default
//   (NN.param:81-83 selects 'p3m'; the 5x5x5 tiling in the run logs corroborates the range).
//
// Mapping: one warp owns one cell (side >= r_cut); lanes are the cell's i-particles.  The 27
// neighbour cells are streamed through a per-warp shared-memory buffer, 32 j-particles at a
// time, and read back as LDS.128 broadcasts.  Coordinates are cell-relative float32
// (sort.cu), the neighbour-cell offset is folded into the i-coordinates once per cell, so
// periodic wrapping costs nothing in the pair loop.
// Split factor: 1 - u³·Q(u²) with Q a degree-10 polynomial in t = 2u²/u_max² - 1 (host-fitted
// Chebyshev interpolant; |error| < 3e-7 absolute, the float32 floor).
// FP32 in the pair loop, float64 accumulation across neighbour cells.
#include "common.cuh"

#define P3M_WARPS 4
#define P3M_DEG 10

struct P3mParams {
  float rc2;         // r_cut²
  float eps2;        // softening²
  float w_scale;     // u² = r² * w_scale,  w_scale = 1/(4 r_s²)
  float t_scale;     // t = u² * t_scale - 1
  float cs;          // cell side
  float coef[P3M_DEG + 1];   // Q(t) monomial coefficients, low to high
};

__device__ __forceinline__ float p3m_rsqrt(float x) {
  float r;
  asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

__global__ void __launch_bounds__(P3M_WARPS * 32)
p3m_short_kernel(const float4* __restrict__ sorted, const int* __restrict__ cell_start,
                 const int* __restrict__ perm, int ncs, int64_t cell_lo, int64_t cell_hi, P3mParams prm,
                 double scale, double* __restrict__ acc, int accumulate) {
  __shared__ float4 jbuf[P3M_WARPS][32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t wstride = (int64_t)gridDim.x * P3M_WARPS;
  for (int64_t c = cell_lo + (int64_t)blockIdx.x * P3M_WARPS + warp; c < cell_hi; c += wstride) {
    const int is = cell_start[c], ie = cell_start[c + 1];
    if (is == ie) continue;
    const int cz = (int)(c % ncs);
    const int cy = (int)((c / ncs) % ncs);
    const int cx = (int)(c / ((int64_t)ncs * ncs));
    for (int ib = is; ib < ie; ib += 32) {
      const int i = ib + lane;
      const bool ivalid = i < ie;
      const float4 pi = sorted[ivalid ? i : is];
      double ax = 0.0, ay = 0.0, az = 0.0;
      for (int ox = -1; ox <= 1; ++ox) {
        int nx = cx + ox;
        nx = nx < 0 ? nx + ncs : (nx >= ncs ? nx - ncs : nx);
        for (int oy = -1; oy <= 1; ++oy) {
          int ny = cy + oy;
          ny = ny < 0 ? ny + ncs : (ny >= ncs ? ny - ncs : ny);
          for (int oz = -1; oz <= 1; ++oz) {
            int nz = cz + oz;
            nz = nz < 0 ? nz + ncs : (nz >= ncs ? nz - ncs : nz);
            const int64_t n = ((int64_t)nx * ncs + ny) * ncs + nz;
            const int js = cell_start[n], je = cell_start[n + 1];
            // d = x_j + offset - x_i  ->  fold the offset into the (negated) i coordinates
            const float nxi = (float)ox * prm.cs - pi.x;
            const float nyi = (float)oy * prm.cs - pi.y;
            const float nzi = (float)oz * prm.cs - pi.z;
            float fx = 0.f, fy = 0.f, fz = 0.f;
            for (int jb = js; jb < je; jb += 32) {
              const int cnt = min(32, je - jb);
              __syncwarp();
              if (lane < cnt) jbuf[warp][lane] = sorted[jb + lane];
              __syncwarp();
              for (int t = 0; t < cnt; ++t) {
                const float4 pj = jbuf[warp][t];
                const float dx = pj.x + nxi, dy = pj.y + nyi, dz = pj.z + nzi;
                float r2 = dx * dx;
                r2 = fmaf(dy, dy, r2);
                r2 = fmaf(dz, dz, r2);
                const bool in = (r2 < prm.rc2) && (r2 > 0.f);
                if (!__any_sync(0xffffffffu, in)) continue;
                const float rinv = p3m_rsqrt(r2 + prm.eps2);
                const float w = r2 * prm.w_scale;             // u²
                const float u = w * p3m_rsqrt(fmaxf(w, 1e-30f));
                const float tt = fmaf(w, prm.t_scale, -1.f);
                float q = prm.coef[P3M_DEG];
#pragma unroll
                for (int d = P3M_DEG - 1; d >= 0; --d) q = fmaf(q, tt, prm.coef[d]);
                const float f = fmaf(-(w * u), q, 1.f);        // erfc(u) + 2u/sqrt(pi) exp(-u²)
                float s = (pj.w * rinv) * (rinv * rinv) * f;
                s = in ? s : 0.f;
                fx = fmaf(s, dx, fx);
                fy = fmaf(s, dy, fy);
                fz = fmaf(s, dz, fz);
              }
            }
            ax += (double)fx; ay += (double)fy; az += (double)fz;
          }
        }
      }
      if (ivalid) {
        const int64_t o = perm[i];
        if (accumulate) {
          acc[3 * o] += ax * scale; acc[3 * o + 1] += ay * scale; acc[3 * o + 2] += az * scale;
        } else {
          acc[3 * o] = ax * scale; acc[3 * o + 1] = ay * scale; acc[3 * o + 2] = az * scale;
        }
      }
    }
  }
}

// coef_host: P3M_DEG+1 doubles (monomial coefficients of Q in t, low to high), HOST pointer.
extern "C" int nb_p3m_short(nb_ctx* ctx, const void* sorted_xyzw, const int32_t* cell_start, const int32_t* perm,
                            int ncs, int64_t cell_lo, int64_t cell_hi, double cell_side, double r_cut,
                            double r_s, double eps, const double* coef_host, int ncoef, double u_max,
                            double scale, int accumulate, double* acc, void* stream) {
  NB_REQUIRE(ctx && sorted_xyzw && cell_start && perm && coef_host && acc, "NULL argument");
  NB_REQUIRE(ncoef == P3M_DEG + 1, "split polynomial must have 11 coefficients");
  NB_REQUIRE(ncs >= 3, "need at least 3 cells per side (27 distinct neighbours)");
  NB_REQUIRE(cell_side >= r_cut && r_cut > 0 && r_s > 0, "cell side must be >= r_cut > 0");
  NB_REQUIRE(r_cut / (2 * r_s) <= u_max * (1 + 1e-9), "polynomial range does not cover r_cut");
  const int64_t nc = (int64_t)ncs * ncs * ncs;
  NB_REQUIRE(cell_lo >= 0 && cell_hi <= nc && cell_lo <= cell_hi, "bad cell range");
  if (cell_lo == cell_hi) return NB_OK;
  cudaStream_t st = (cudaStream_t)stream;
  NB_CHECK_CUDA(cudaSetDevice(ctx->device));
  P3mParams prm;
  prm.rc2 = (float)(r_cut * r_cut);
  prm.eps2 = (float)(eps * eps);
  prm.w_scale = (float)(1.0 / (4.0 * r_s * r_s));
  prm.t_scale = (float)(2.0 / (u_max * u_max));
  prm.cs = (float)cell_side;
  for (int d = 0; d <= P3M_DEG; ++d) prm.coef[d] = (float)coef_host[d];
  const int64_t ncell = cell_hi - cell_lo;
  int64_t nblk = (ncell + P3M_WARPS - 1) / P3M_WARPS;
  const int64_t cap = (int64_t)ctx->num_sms * 64;
  if (nblk > cap) nblk = cap;
  p3m_short_kernel<<<(int)nblk, P3M_WARPS * 32, 0, st>>>((const float4*)sorted_xyzw, cell_start, perm, ncs, cell_lo,
                                                         cell_hi, prm, scale, acc, accumulate);
  NB_LAUNCHED(ctx);
  return NB_OK;
}

// Pair statistics of the same traversal (not timed; bench.py uses it to convert the kernel
// time into candidate-pair and in-range-pair rates): out[0] = candidate pairs, out[1] = pairs
// with 0 < r < r_cut.
__global__ void __launch_bounds__(P3M_WARPS * 32)
p3m_count_kernel(const float4* __restrict__ sorted, const int* __restrict__ cell_start, int ncs, int64_t cell_lo,
                 int64_t cell_hi, float cs, float rc2, unsigned long long* __restrict__ out) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t wstride = (int64_t)gridDim.x * P3M_WARPS;
  unsigned long long cand = 0, inr = 0;
  for (int64_t c = cell_lo + (int64_t)blockIdx.x * P3M_WARPS + warp; c < cell_hi; c += wstride) {
    const int is = cell_start[c], ie = cell_start[c + 1];
    if (is == ie) continue;
    const int cz = (int)(c % ncs), cy = (int)((c / ncs) % ncs), cx = (int)(c / ((int64_t)ncs * ncs));
    for (int i = is + lane; i < ie; i += 32) {
      const float4 pi = sorted[i];
      for (int ox = -1; ox <= 1; ++ox)
        for (int oy = -1; oy <= 1; ++oy)
          for (int oz = -1; oz <= 1; ++oz) {
            const int nx = (cx + ox + ncs) % ncs, ny = (cy + oy + ncs) % ncs, nz = (cz + oz + ncs) % ncs;
            const int64_t n = ((int64_t)nx * ncs + ny) * ncs + nz;
            const int js = cell_start[n], je = cell_start[n + 1];
            cand += (unsigned long long)(je - js);
            for (int j = js; j < je; ++j) {
              const float4 pj = sorted[j];
              const float dx = pj.x + ox * cs - pi.x, dy = pj.y + oy * cs - pi.y, dz = pj.z + oz * cs - pi.z;
              const float r2 = dx * dx + dy * dy + dz * dz;
              inr += (r2 < rc2 && r2 > 0.f) ? 1ull : 0ull;
            }
          }
    }
  }
  for (int o = 16; o > 0; o >>= 1) {
    cand += __shfl_xor_sync(0xffffffffu, cand, o);
    inr += __shfl_xor_sync(0xffffffffu, inr, o);
  }
  if (lane == 0) {
    atomicAdd(&out[0], cand);
    atomicAdd(&out[1], inr);
  }
}

extern "C" int nb_p3m_count_pairs(nb_ctx* ctx, const void* sorted_xyzw, const int32_t* cell_start, int ncs,
                                  int64_t cell_lo, int64_t cell_hi, double cell_side, double r_cut,
                                  int64_t* out2, void* stream) {
  NB_REQUIRE(ctx && sorted_xyzw && cell_start && out2, "NULL argument");
  NB_REQUIRE(ncs >= 3 && cell_lo >= 0 && cell_lo <= cell_hi && cell_hi <= (int64_t)ncs * ncs * ncs, "bad cells");
  cudaStream_t st = (cudaStream_t)stream;
  NB_CHECK_CUDA(cudaSetDevice(ctx->device));
  NB_CHECK_CUDA(cudaMemsetAsync(out2, 0, 2 * sizeof(int64_t), st));
  if (cell_lo == cell_hi) return NB_OK;
  int64_t nblk = (cell_hi - cell_lo + P3M_WARPS - 1) / P3M_WARPS;
  const int64_t cap = (int64_t)ctx->num_sms * 64;
  if (nblk > cap) nblk = cap;
  p3m_count_kernel<<<(int)nblk, P3M_WARPS * 32, 0, st>>>((const float4*)sorted_xyzw, cell_start, ncs, cell_lo, cell_hi,
                                                         (float)cell_side, (float)(r_cut * r_cut),
                                                         (unsigned long long*)out2);
  NB_LAUNCHED(ctx);
  return NB_OK;
}
