// K4 — PP direct-sum gravity, shared-memory j-tiles, FP32 FMA + rsqrt.
// No reference source exists for this path (CONCEPT's PP is external); the arithmetic is
// our softened Newtonian sum   a_i = G Σ_j m_j d / (|d|² + ε²)^{3/2},  d = x_j - x_i,
// checked against oracle/force.py:pp_accel.
//
// Blackwell notes: this is not a dense contraction, so no tensor cores.  The inner loop is
// 12 FP32-pipe operations + 1 MUFU.RSQ per interaction (20 flop by the usual convention).
// sm_100a has packed FP32 (FFMA2/FADD2/FMUL2 from PTX fma/add/mul.f32x2): two i-particles
// ride in one 64-bit register pair, so one interaction costs 6 issue slots on the FMA pipe
// plus the MUFU.  Both the scalar and the packed loop are kept (template) — bench.py reports
// which one the build uses and tools/kbench.py times both.
//
// Accuracy: FP32 inside a j-tile (512 terms), tile partial sums are folded into FP64
// accumulators (3 DADD per tile per i-particle — free at this arithmetic intensity).
#include "common.cuh"

#define PP_THREADS 256
#define PP_TJ 512

__device__ __forceinline__ float rsqrt_approx(float x) {
  float r;
  asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pack2(float lo, float hi) {
  f32x2 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void unpack2(f32x2 v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) {
  f32x2 d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) {
  f32x2 d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) {
  f32x2 d;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}

// pos (Np,3) f64 -> float4 (x - ox, y - oy, z - oz, w)
__global__ void __launch_bounds__(256)
pp_pack_kernel(const double* __restrict__ pos, int64_t np, double ox, double oy, double oz, float w,
               const double* __restrict__ masses, float4* __restrict__ out) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < np; p += stride) {
    out[p] = make_float4((float)(pos[3 * p] - ox), (float)(pos[3 * p + 1] - oy), (float)(pos[3 * p + 2] - oz),
                         masses ? (float)masses[p] : w);
  }
}

template <int IPT, bool PACKED, bool PERIODIC, bool GUARD>
__global__ void __launch_bounds__(PP_THREADS)
pp_accel_kernel(const float4* __restrict__ ipos, int64_t ni, const float4* __restrict__ jpos, int64_t nj,
                int64_t j_chunk, float eps2, float box, float inv_box, double scale, double* __restrict__ acc) {
  __shared__ float4 tile[PP_TJ];
  const int64_t ibase = (int64_t)blockIdx.x * (PP_THREADS * IPT) + threadIdx.x;
  const int64_t j0 = (int64_t)blockIdx.y * j_chunk;
  const int64_t j1 = (j0 + j_chunk < nj) ? j0 + j_chunk : nj;

  float nxi[IPT], nyi[IPT], nzi[IPT];   // negated i coordinates: d = x_j + (-x_i)
  double ax64[IPT], ay64[IPT], az64[IPT];
#pragma unroll
  for (int q = 0; q < IPT; ++q) {
    int64_t i = ibase + (int64_t)q * PP_THREADS;
    if (i >= ni) i = ni - 1;
    const float4 p = ipos[i];
    nxi[q] = -p.x; nyi[q] = -p.y; nzi[q] = -p.z;
    ax64[q] = ay64[q] = az64[q] = 0.0;
  }
  // GUARD (eps == 0): coincident points (the self term) contribute nothing instead of NaN

  for (int64_t jt = j0; jt < j1; jt += PP_TJ) {
    __syncthreads();
#pragma unroll
    for (int t = threadIdx.x; t < PP_TJ; t += PP_THREADS) {
      const int64_t j = jt + t;
      tile[t] = (j < j1) ? jpos[j] : make_float4(0.f, 0.f, 0.f, 0.f);   // zero mass: no contribution
    }
    __syncthreads();

    if constexpr (!PACKED) {
      float ax[IPT], ay[IPT], az[IPT];
#pragma unroll
      for (int q = 0; q < IPT; ++q) ax[q] = ay[q] = az[q] = 0.f;
#pragma unroll 4
      for (int t = 0; t < PP_TJ; ++t) {
        const float4 pj = tile[t];
#pragma unroll
        for (int q = 0; q < IPT; ++q) {
          float dx = pj.x + nxi[q], dy = pj.y + nyi[q], dz = pj.z + nzi[q];
          if constexpr (PERIODIC) {
            dx = fmaf(-box, rintf(dx * inv_box), dx);
            dy = fmaf(-box, rintf(dy * inv_box), dy);
            dz = fmaf(-box, rintf(dz * inv_box), dz);
          }
          float r2 = fmaf(dx, dx, eps2);
          r2 = fmaf(dy, dy, r2);
          r2 = fmaf(dz, dz, r2);
          const float rinv = rsqrt_approx(r2);
          float s = (pj.w * rinv) * (rinv * rinv);
          if constexpr (GUARD) s = (r2 > 0.f) ? s : 0.f;
          ax[q] = fmaf(s, dx, ax[q]);
          ay[q] = fmaf(s, dy, ay[q]);
          az[q] = fmaf(s, dz, az[q]);
        }
      }
#pragma unroll
      for (int q = 0; q < IPT; ++q) {
        ax64[q] += (double)ax[q]; ay64[q] += (double)ay[q]; az64[q] += (double)az[q];
      }
    } else {
      constexpr int H = IPT / 2;
      f32x2 nx2[H], ny2[H], nz2[H], ax2[H], ay2[H], az2[H];
#pragma unroll
      for (int h = 0; h < H; ++h) {
        nx2[h] = pack2(nxi[2 * h], nxi[2 * h + 1]);
        ny2[h] = pack2(nyi[2 * h], nyi[2 * h + 1]);
        nz2[h] = pack2(nzi[2 * h], nzi[2 * h + 1]);
        ax2[h] = ay2[h] = az2[h] = pack2(0.f, 0.f);
      }
      const f32x2 eps22 = pack2(eps2, eps2);
#pragma unroll 4
      for (int t = 0; t < PP_TJ; ++t) {
        const float4 pj = tile[t];
        const f32x2 xj = pack2(pj.x, pj.x), yj = pack2(pj.y, pj.y), zj = pack2(pj.z, pj.z), wj = pack2(pj.w, pj.w);
#pragma unroll
        for (int h = 0; h < H; ++h) {
          f32x2 dx = add2(xj, nx2[h]), dy = add2(yj, ny2[h]), dz = add2(zj, nz2[h]);
          if constexpr (PERIODIC) {
            float a, b;
            unpack2(dx, a, b); dx = pack2(fmaf(-box, rintf(a * inv_box), a), fmaf(-box, rintf(b * inv_box), b));
            unpack2(dy, a, b); dy = pack2(fmaf(-box, rintf(a * inv_box), a), fmaf(-box, rintf(b * inv_box), b));
            unpack2(dz, a, b); dz = pack2(fmaf(-box, rintf(a * inv_box), a), fmaf(-box, rintf(b * inv_box), b));
          }
          f32x2 r2 = fma2(dx, dx, eps22);
          r2 = fma2(dy, dy, r2);
          r2 = fma2(dz, dz, r2);
          float r2a, r2b;
          unpack2(r2, r2a, r2b);
          const f32x2 rinv = pack2(rsqrt_approx(r2a), rsqrt_approx(r2b));
          f32x2 s = mul2(mul2(wj, rinv), mul2(rinv, rinv));
          if constexpr (GUARD) {
            float sa, sb;
            unpack2(s, sa, sb);
            s = pack2(r2a > 0.f ? sa : 0.f, r2b > 0.f ? sb : 0.f);
          }
          ax2[h] = fma2(s, dx, ax2[h]);
          ay2[h] = fma2(s, dy, ay2[h]);
          az2[h] = fma2(s, dz, az2[h]);
        }
      }
#pragma unroll
      for (int h = 0; h < H; ++h) {
        float a, b;
        unpack2(ax2[h], a, b); ax64[2 * h] += (double)a; ax64[2 * h + 1] += (double)b;
        unpack2(ay2[h], a, b); ay64[2 * h] += (double)a; ay64[2 * h + 1] += (double)b;
        unpack2(az2[h], a, b); az64[2 * h] += (double)a; az64[2 * h + 1] += (double)b;
      }
    }
  }
#pragma unroll
  for (int q = 0; q < IPT; ++q) {
    const int64_t i = ibase + (int64_t)q * PP_THREADS;
    if (i < ni) {
      atomicAdd(&acc[3 * i + 0], ax64[q] * scale);
      atomicAdd(&acc[3 * i + 1], ay64[q] * scale);
      atomicAdd(&acc[3 * i + 2], az64[q] * scale);
    }
  }
}

// ---------------------------------------------------------------------------------------------------
// Near-field fix of the FP32 coordinates.  nb_pp_pack rounds x - origin to float32: 24 bits over the box
// (1.5e-5 Mpc at |x| = 250 Mpc).  For a pair closer than a few softening lengths that is 1e-4 of its
// separation, and ONE such pair puts the particle's acceleration 4e-4 (of the rms) off the float64 sum
// at 128^3 (median 1.6e-7, 99th percentile 1.7e-5: measured, tests/test_gpu_baseline_sizes.py).
// Instead of paying two FP32 words per coordinate for all Np^2 pairs, the few pairs with r < r_c are
// visited once more through a cell list (nb_cell_sort): the float64 pair term minus the term the
// all-pairs kernel added for the same pair — replayed with the same float32 operations, rsqrt.approx
// included — is summed in float64 and added to acc.  ~0.5 pairs per particle at r_c = 20 eps.
template <bool PERIODIC>
__global__ void __launch_bounds__(256)
pp_near_fix_kernel(const double* __restrict__ pos, int64_t i0, int64_t ni, const int* __restrict__ perm,
                   const int* __restrict__ cell_start, int ncs, double L, double inv_cs, double ox, double oy,
                   double oz, float w, double eps2, double rc2, double scale, double* __restrict__ acc) {
  const float eps2f = (float)eps2, boxf = (float)L, inv_boxf = 1.f / (float)L;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = i0 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < i0 + ni; i += stride) {
    const double xi = pos[3 * i], yi = pos[3 * i + 1], zi = pos[3 * i + 2];
    int ci[3];
    {
      const double v[3] = {xi, yi, zi};
#pragma unroll
      for (int a = 0; a < 3; ++a) {
        double xw = v[a];
        if (xw < 0.0 || xw >= L) xw = nb_pymod(xw, L);
        int c = (int)(xw * inv_cs);
        ci[a] = c >= ncs ? ncs - 1 : (c < 0 ? 0 : c);
      }
    }
    const float nxi = -(float)(xi - ox), nyi = -(float)(yi - oy), nzi = -(float)(zi - oz);
    double sx = 0.0, sy = 0.0, sz = 0.0;
    for (int dxc = -1; dxc <= 1; ++dxc)
      for (int dyc = -1; dyc <= 1; ++dyc)
        for (int dzc = -1; dzc <= 1; ++dzc) {
          int cx = ci[0] + dxc, cy = ci[1] + dyc, cz = ci[2] + dzc;
          if (PERIODIC) {
            cx = (cx + ncs) % ncs; cy = (cy + ncs) % ncs; cz = (cz + ncs) % ncs;
          } else if (cx < 0 || cy < 0 || cz < 0 || cx >= ncs || cy >= ncs || cz >= ncs) {
            continue;
          }
          const int c = (cx * ncs + cy) * ncs + cz;
          for (int s = cell_start[c], e = cell_start[c + 1]; s < e; ++s) {
            const int64_t j = perm[s];
            if (j == i) continue;
            const double xj = pos[3 * j], yj = pos[3 * j + 1], zj = pos[3 * j + 2];
            double dx = xj - xi, dy = yj - yi, dz = zj - zi;
            if (PERIODIC) {
              dx -= L * rint(dx / L); dy -= L * rint(dy / L); dz -= L * rint(dz / L);
            }
            const double r2 = dx * dx + dy * dy + dz * dz;
            if (r2 >= rc2) continue;
            // float64 pair term
            const double q2 = r2 + eps2;
            const double inv = q2 > 0.0 ? 1.0 / sqrt(q2) : 0.0;
            const double fp = (double)w * inv * inv * inv;
            // the term pp_accel_kernel added for this pair (same operations, same rounding)
            float fx = (float)(xj - ox) + nxi, fy = (float)(yj - oy) + nyi, fz = (float)(zj - oz) + nzi;
            if (PERIODIC) {
              fx = fmaf(-boxf, rintf(fx * inv_boxf), fx);
              fy = fmaf(-boxf, rintf(fy * inv_boxf), fy);
              fz = fmaf(-boxf, rintf(fz * inv_boxf), fz);
            }
            float r2f = fmaf(fx, fx, eps2f);
            r2f = fmaf(fy, fy, r2f);
            r2f = fmaf(fz, fz, r2f);
            const float rinv = rsqrt_approx(r2f);
            float sf = (w * rinv) * (rinv * rinv);
            if (!(r2f > 0.f)) sf = 0.f;
            sx += fp * dx - (double)sf * (double)fx;
            sy += fp * dy - (double)sf * (double)fy;
            sz += fp * dz - (double)sf * (double)fz;
          }
        }
    acc[3 * (i - i0) + 0] += sx * scale;
    acc[3 * (i - i0) + 1] += sy * scale;
    acc[3 * (i - i0) + 2] += sz * scale;
  }
}

extern "C" int nb_pp_near_fix(nb_ctx* ctx, const double* pos, int64_t np, int64_t i0, int64_t ni, const int32_t* perm,
                              const int32_t* cell_start, int ncs, double L, double ox, double oy, double oz,
                              double mass_w, double eps, double r_c, int periodic, double scale, double* acc,
                              void* stream) {
  NB_REQUIRE(ctx, "ctx is NULL");
  if (np == 0 || ni == 0) return NB_OK;
  NB_REQUIRE(pos && perm && cell_start && acc, "NULL argument");
  NB_REQUIRE(i0 >= 0 && ni >= 0 && i0 + ni <= np, "bad i range");
  NB_REQUIRE(ncs >= 3 && L > 0 && r_c > 0 && L / ncs >= r_c && eps >= 0, "cells must be at least r_c wide, ncs >= 3");
  cudaStream_t st = (cudaStream_t)stream;
  NB_CHECK_CUDA(cudaSetDevice(ctx->device));
  const int nblk = nb_grid_for(ctx, ni, 256, 16);
  if (periodic)
    pp_near_fix_kernel<true><<<nblk, 256, 0, st>>>(pos, i0, ni, perm, cell_start, ncs, L, (double)ncs / L, ox, oy, oz,
                                                   (float)mass_w, eps * eps, r_c * r_c, scale, acc);
  else
    pp_near_fix_kernel<false><<<nblk, 256, 0, st>>>(pos, i0, ni, perm, cell_start, ncs, L, (double)ncs / L, ox, oy, oz,
                                                    (float)mass_w, eps * eps, r_c * r_c, scale, acc);
  NB_LAUNCHED(ctx);
  return NB_OK;
}

extern "C" int nb_pp_pack(nb_ctx* ctx, const double* pos, int64_t np, double ox, double oy, double oz,
                          double mass, const double* masses, void* out_xyzw, void* stream) {
  NB_REQUIRE(ctx, "ctx is NULL");
  if (np == 0) return NB_OK;
  NB_REQUIRE(out_xyzw && pos, "NULL argument");
  cudaStream_t st = (cudaStream_t)stream;
  NB_CHECK_CUDA(cudaSetDevice(ctx->device));
  pp_pack_kernel<<<nb_grid_for(ctx, np, 256, 16), 256, 0, st>>>(pos, np, ox, oy, oz, (float)mass, masses,
                                                               (float4*)out_xyzw);
  NB_LAUNCHED(ctx);
  return NB_OK;
}

template <int IPT, bool PACKED>
static int pp_launch(nb_ctx* ctx, const float4* ipos, int64_t ni, const float4* jpos, int64_t nj, float eps2,
                     float box, double scale, double* acc, cudaStream_t st) {
  const int64_t per_block = (int64_t)PP_THREADS * IPT;
  const int64_t nbi = (ni + per_block - 1) / per_block;
  // split j so that the grid has several waves of CTAs even for small i-shards
  int64_t jsplit = 1;
  const int64_t want = (int64_t)ctx->num_sms * 6;
  if (nbi < want) jsplit = (want + nbi - 1) / nbi;
  int64_t max_split = (nj + 4 * PP_TJ - 1) / (4 * PP_TJ);
  if (max_split < 1) max_split = 1;
  if (jsplit > max_split) jsplit = max_split;
  if (jsplit > 65535) jsplit = 65535;
  int64_t j_chunk = (nj + jsplit - 1) / jsplit;
  j_chunk = ((j_chunk + PP_TJ - 1) / PP_TJ) * PP_TJ;
  jsplit = (nj + j_chunk - 1) / j_chunk;
  dim3 grid((unsigned)nbi, (unsigned)jsplit);
  const float ib = box > 0.f ? 1.f / box : 0.f;
  if (box > 0.f) {
    if (eps2 > 0.f)
      pp_accel_kernel<IPT, PACKED, true, false><<<grid, PP_THREADS, 0, st>>>(ipos, ni, jpos, nj, j_chunk, eps2, box, ib, scale, acc);
    else
      pp_accel_kernel<IPT, PACKED, true, true><<<grid, PP_THREADS, 0, st>>>(ipos, ni, jpos, nj, j_chunk, eps2, box, ib, scale, acc);
  } else {
    if (eps2 > 0.f)
      pp_accel_kernel<IPT, PACKED, false, false><<<grid, PP_THREADS, 0, st>>>(ipos, ni, jpos, nj, j_chunk, eps2, box, ib, scale, acc);
    else
      pp_accel_kernel<IPT, PACKED, false, true><<<grid, PP_THREADS, 0, st>>>(ipos, ni, jpos, nj, j_chunk, eps2, box, ib, scale, acc);
  }
  NB_LAUNCHED(ctx);
  return NB_OK;
}

extern "C" int nb_pp_accel(nb_ctx* ctx, const void* ipos_xyzw, int64_t ni, const void* jpos_xyzw, int64_t nj,
                           double eps, double box, double scale, int variant, int accumulate, double* acc,
                           void* stream) {
  NB_REQUIRE(ctx, "ctx is NULL");
  if (ni == 0) return NB_OK;
  NB_REQUIRE(acc && ipos_xyzw && (nj == 0 || jpos_xyzw), "NULL argument");
  NB_REQUIRE(eps >= 0 && box >= 0, "eps and box must be >= 0");
  NB_REQUIRE(variant >= 0 && variant <= 3, "variant: 0 default, 1 scalar x4, 2 packed x4, 3 packed x8");
  cudaStream_t st = (cudaStream_t)stream;
  NB_CHECK_CUDA(cudaSetDevice(ctx->device));
  if (!accumulate && ni > 0) NB_CHECK_CUDA(cudaMemsetAsync(acc, 0, sizeof(double) * 3 * ni, st));
  if (ni == 0 || nj == 0) return NB_OK;
  const float4* ip = (const float4*)ipos_xyzw;
  const float4* jp = (const float4*)jpos_xyzw;
  const float eps2 = (float)(eps * eps);
  if (variant == 0) variant = 2;
  switch (variant) {
    case 1: return pp_launch<4, false>(ctx, ip, ni, jp, nj, eps2, (float)box, scale, acc, st);
    case 2: return pp_launch<4, true>(ctx, ip, ni, jp, nj, eps2, (float)box, scale, acc, st);
    default: return pp_launch<8, true>(ctx, ip, ni, jp, nj, eps2, (float)box, scale, acc, st);
  }
}
