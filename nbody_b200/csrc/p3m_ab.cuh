// A/B kernels of the P3M short range: the generations v1, v2, v3, v5 and the byte-mask form of the v6 pass.
// Compiled only with -DNB_AB_KERNELS (make AB=1); the product library holds v6 and the v4 overflow pass.
// Included by p3m.cu after the shared helpers (packed FP32 ops, P3m3Packed, p3m4_*, p3m_make_packed).
#pragma once

__global__ void __launch_bounds__(P3M_WARPS * 32)
p3m_short_kernel(const float4* __restrict__ sorted, const int* __restrict__ cell_start,
                 const int* __restrict__ perm, int ncs, int64_t cell_lo, int64_t cell_hi, P3mParams prm,
                 double scale, double* __restrict__ acc, int accumulate, int i_limit) {
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
      // particles with perm >= i_limit are halo copies of another slab's particles: sources only
      const int oi = i < ie ? perm[i] : 0x7fffffff;
      const bool ivalid = oi < i_limit;
      if (!__any_sync(0xffffffffu, ivalid)) continue;
      const float4 pi = sorted[i < ie ? i : is];
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
        const int64_t o = oi;
        if (accumulate) {
          acc[3 * o] += ax * scale; acc[3 * o + 1] += ay * scale; acc[3 * o + 2] += az * scale;
        } else {
          acc[3 * o] = ax * scale; acc[3 * o + 1] = ay * scale; acc[3 * o + 2] = az * scale;
        }
      }
    }
  }
}


// ---------------------------------------------------------------------------------------
// v2 — "uniform i, lanes over j" with in-range compaction (experimental, NB_P3M_KERNEL=2).
//
// The v1 kernel above evaluates the full split polynomial for every candidate pair of the 27
// neighbour cells although only ~16 % of them are inside r_cut (27 cs³ against 4.19 r_cut³):
// measured 61 issue slots per candidate on B200.  v2 separates the cheap test from the
// expensive evaluation:
//   tile build  the j-particles of all 27 neighbour cells are copied once per (cell, i-block)
//               into a per-warp SoA shared-memory tile, already translated into the frame of
//               the i-cell (periodic wrap costs nothing later);
//   phase 1     for one i-particle (warp-uniform, broadcast by shuffle) the 32 lanes test 64
//               candidates per iteration with packed FP32 (LDS.64 + FADD2/FFMA2), and the
//               in-range separations (dx,dy,dz) are compacted with ballot/popc into a
//               per-warp list — conflict-free, no atomics;
//   phase 2     the list is drained 32 pairs at a time with every lane busy: softened
//               rsqrt, split polynomial (two interleaved Horner chains), FMA accumulation;
//   reduce      one xor-shuffle reduction per i-particle, added to the float64 accumulator
//               held by the lane that owns that i-particle.
// sorted_xyzw.w (the mass) must be uniform: nb_cell_sort writes the scalar `mass` there.
#define P3M2_WARPS 4
#define P3M2_CAP 896     // candidate tile capacity per warp (multiple of 64)
#define P3M2_LCAP 256    // in-range list capacity per warp
#define P3M2_FLOATS_PER_WARP (3 * P3M2_CAP + 3 * P3M2_LCAP)
// phase 2: evaluate the `cnt` compacted separations, 32 per iteration
__device__ __forceinline__ void p3m2_drain(const float* __restrict__ lx, const float* __restrict__ ly,
                                           const float* __restrict__ lz, int cnt, int lane, const P3mParams& prm,
                                           float wj, float& fx, float& fy, float& fz) {
  __syncwarp();
  for (int q0 = 0; q0 < cnt; q0 += 32) {
    const int q = q0 + lane;
    if (q < cnt) {
      const float dx = lx[q], dy = ly[q], dz = lz[q];
      float r2 = dx * dx;
      r2 = fmaf(dy, dy, r2);
      r2 = fmaf(dz, dz, r2);
      // coincident points (the self pair): d = 0, the clamp keeps rinv³ finite -> contributes 0
      const float rinv = p3m_rsqrt(fmaxf(r2 + prm.eps2, 1e-20f));
      const float w = r2 * prm.w_scale;                 // u²
      const float u = w * p3m_rsqrt(fmaxf(w, 1e-30f));
      const float tt = fmaf(w, prm.t_scale, -1.f);
      const float t2 = tt * tt;
      float qe = prm.coef[10], qo = prm.coef[9];
      qe = fmaf(qe, t2, prm.coef[8]); qo = fmaf(qo, t2, prm.coef[7]);
      qe = fmaf(qe, t2, prm.coef[6]); qo = fmaf(qo, t2, prm.coef[5]);
      qe = fmaf(qe, t2, prm.coef[4]); qo = fmaf(qo, t2, prm.coef[3]);
      qe = fmaf(qe, t2, prm.coef[2]); qo = fmaf(qo, t2, prm.coef[1]);
      qe = fmaf(qe, t2, prm.coef[0]);
      const float qq = fmaf(qo, tt, qe);
      const float f = fmaf(-(w * u), qq, 1.f);          // erfc(u) + 2u/sqrt(pi) exp(-u²)
      const float s = (wj * rinv) * (rinv * rinv) * f;
      fx = fmaf(s, dx, fx);
      fy = fmaf(s, dy, fy);
      fz = fmaf(s, dz, fz);
    }
  }
  __syncwarp();
}

// all i-particles of the block against the `nt` candidates in the tile
__device__ __forceinline__ void p3m2_process(float* __restrict__ tx, float* __restrict__ ty, float* __restrict__ tz,
                                             float* __restrict__ lx, float* __restrict__ ly, float* __restrict__ lz,
                                             int nt, int nib, const float4& pi, int lane, const P3mParams& prm,
                                             float wj, double& ax, double& ay, double& az) {
  const int ntpad = (nt + 63) & ~63;
  for (int t = nt + lane; t < ntpad; t += 32) {   // sentinels: far outside any cutoff
    tx[t] = 1e18f;
    ty[t] = 0.f;
    tz[t] = 0.f;
  }
  __syncwarp();
  const unsigned lt = (1u << lane) - 1u;
  for (int ii = 0; ii < nib; ++ii) {
    const float xi = __shfl_sync(0xffffffffu, pi.x, ii);
    const float yi = __shfl_sync(0xffffffffu, pi.y, ii);
    const float zi = __shfl_sync(0xffffffffu, pi.z, ii);
    const p3m_f32x2 nx2 = p3m_pack2(-xi, -xi), ny2 = p3m_pack2(-yi, -yi), nz2 = p3m_pack2(-zi, -zi);
    int cnt = 0;
    float fx = 0.f, fy = 0.f, fz = 0.f;
#pragma unroll 2
    for (int v = 0; v < ntpad; v += 64) {
      const p3m_f32x2 X = *reinterpret_cast<const p3m_f32x2*>(tx + v + 2 * lane);
      const p3m_f32x2 Y = *reinterpret_cast<const p3m_f32x2*>(ty + v + 2 * lane);
      const p3m_f32x2 Z = *reinterpret_cast<const p3m_f32x2*>(tz + v + 2 * lane);
      const p3m_f32x2 dx = p3m_add2(X, nx2), dy = p3m_add2(Y, ny2), dz = p3m_add2(Z, nz2);
      p3m_f32x2 r2 = p3m_mul2(dx, dx);
      r2 = p3m_fma2(dy, dy, r2);
      r2 = p3m_fma2(dz, dz, r2);
      float r2a, r2b;
      p3m_unpack2(r2, r2a, r2b);
      const bool in0 = r2a < prm.rc2, in1 = r2b < prm.rc2;
      const unsigned m0 = __ballot_sync(0xffffffffu, in0);
      const unsigned m1 = __ballot_sync(0xffffffffu, in1);
      const int c0 = __popc(m0);
      if (in0) {
        const int p = cnt + __popc(m0 & lt);
        float a, b;
        p3m_unpack2(dx, a, b); lx[p] = a;
        p3m_unpack2(dy, a, b); ly[p] = a;
        p3m_unpack2(dz, a, b); lz[p] = a;
      }
      if (in1) {
        const int p = cnt + c0 + __popc(m1 & lt);
        float a, b;
        p3m_unpack2(dx, a, b); lx[p] = b;
        p3m_unpack2(dy, a, b); ly[p] = b;
        p3m_unpack2(dz, a, b); lz[p] = b;
      }
      cnt += c0 + __popc(m1);
      if (cnt > P3M2_LCAP - 64) {
        p3m2_drain(lx, ly, lz, cnt, lane, prm, wj, fx, fy, fz);
        cnt = 0;
      }
    }
    p3m2_drain(lx, ly, lz, cnt, lane, prm, wj, fx, fy, fz);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      fx += __shfl_xor_sync(0xffffffffu, fx, o);
      fy += __shfl_xor_sync(0xffffffffu, fy, o);
      fz += __shfl_xor_sync(0xffffffffu, fz, o);
    }
    if (lane == ii) {
      ax += (double)fx;
      ay += (double)fy;
      az += (double)fz;
    }
  }
  __syncwarp();
}

__global__ void __launch_bounds__(P3M2_WARPS * 32)
p3m_short_v2_kernel(const float4* __restrict__ sorted, const int* __restrict__ cell_start,
                    const int* __restrict__ perm, int ncs, int64_t cell_lo, int64_t cell_hi, P3mParams prm,
                    double scale, double* __restrict__ acc, int accumulate, int i_limit) {
  extern __shared__ __align__(16) float p3m2_smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float* tx = p3m2_smem + warp * P3M2_FLOATS_PER_WARP;
  float* ty = tx + P3M2_CAP;
  float* tz = ty + P3M2_CAP;
  float* lx = tz + P3M2_CAP;
  float* ly = lx + P3M2_LCAP;
  float* lz = ly + P3M2_LCAP;
  const float wj = __ldg(&sorted[0].w);   // uniform mass (nb_cell_sort)
  const int64_t wstride = (int64_t)gridDim.x * P3M2_WARPS;
  for (int64_t c = cell_lo + (int64_t)blockIdx.x * P3M2_WARPS + warp; c < cell_hi; c += wstride) {
    const int is = cell_start[c], ie = cell_start[c + 1];
    if (is == ie) continue;
    const int cz = (int)(c % ncs);
    const int cy = (int)((c / ncs) % ncs);
    const int cx = (int)(c / ((int64_t)ncs * ncs));
    // lanes 0..26 look up the j-ranges of the 27 neighbour cells
    int my_js = 0, my_je = 0;
    if (lane < 27) {
      int nx = cx + lane / 9 - 1, ny = cy + (lane / 3) % 3 - 1, nz = cz + lane % 3 - 1;
      nx = nx < 0 ? nx + ncs : (nx >= ncs ? nx - ncs : nx);
      ny = ny < 0 ? ny + ncs : (ny >= ncs ? ny - ncs : ny);
      nz = nz < 0 ? nz + ncs : (nz >= ncs ? nz - ncs : nz);
      const int64_t n = ((int64_t)nx * ncs + ny) * ncs + nz;
      my_js = cell_start[n];
      my_je = cell_start[n + 1];
    }
    for (int ib = is; ib < ie; ib += 32) {
      const int i = ib + lane;
      // particles with perm >= i_limit are halo copies of another slab's particles: sources only
      const int oi = i < ie ? perm[i] : 0x7fffffff;
      const bool ivalid = oi < i_limit;
      if (!__any_sync(0xffffffffu, ivalid)) continue;
      const float4 pi = sorted[i < ie ? i : is];
      const int nib = min(32, ie - ib);
      double ax = 0.0, ay = 0.0, az = 0.0;
      int fill = 0;
      for (int n = 0; n < 27; ++n) {
        int js = __shfl_sync(0xffffffffu, my_js, n);
        const int je = __shfl_sync(0xffffffffu, my_je, n);
        const float offx = (float)(n / 9 - 1) * prm.cs;
        const float offy = (float)((n / 3) % 3 - 1) * prm.cs;
        const float offz = (float)(n % 3 - 1) * prm.cs;
        while (js < je) {
          const int take = min(je - js, P3M2_CAP - fill);
          for (int t = lane; t < take; t += 32) {
            const float4 p = sorted[js + t];
            tx[fill + t] = p.x + offx;
            ty[fill + t] = p.y + offy;
            tz[fill + t] = p.z + offz;
          }
          fill += take;
          js += take;
          if (fill == P3M2_CAP) {
            p3m2_process(tx, ty, tz, lx, ly, lz, fill, nib, pi, lane, prm, wj, ax, ay, az);
            fill = 0;
          }
        }
      }
      if (fill > 0) p3m2_process(tx, ty, tz, lx, ly, lz, fill, nib, pi, lane, prm, wj, ax, ay, az);
      if (ivalid) {
        const int64_t o = oi;
        if (accumulate) {
          acc[3 * o] += ax * scale; acc[3 * o + 1] += ay * scale; acc[3 * o + 2] += az * scale;
        } else {
          acc[3 * o] = ax * scale; acc[3 * o + 1] = ay * scale; acc[3 * o + 2] = az * scale;
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------------------
// v3 — "8 i x 4 j-streams, packed pairs" (default).
//
// Two things held v1 back (profiles/r1a, r1b): (a) a cell holds ~22 particles at BASELINE
// config 5, so a warp with one i-particle per lane runs at ~65 % lane occupancy; (b) every
// candidate costs ~40 scalar issue slots.  v3 changes the lane mapping and the arithmetic:
//   lanes     lane = 8·s + i: 8 i-particles of the cell (i-blocks of 8: 22 particles -> 3
//             passes at 92 % occupancy) times 4 j-streams s; the three xor-shuffle levels a
//             per-i reduction needs shrink to two (over s) once per i-block;
//   tile      the j-particles of the 27 neighbour cells are copied once per CELL (not per
//             i-block) into a per-warp SoA shared-memory tile, translated into the frame of
//             the i-cell, padded with far-away sentinels to a multiple of 8;
//   pairs     every lane takes TWO consecutive candidates per iteration as one LDS.64 per
//             coordinate and evaluates them with packed FP32 (add/mul/fma.f32x2 -> FADD2 /
//             FMUL2 / FFMA2): separation, r², softened r^-3, u = w·rsqrt(w), the degree-10
//             split polynomial, accumulation — ~21 issue slots per candidate instead of 40;
//             only the 4 MUFU.RSQ, 2 FSETP and 2 FSEL per pair stay scalar.
// Coincident points (the self pair) have d = 0 and a finite r^-3 (ε² is clamped to >= 1e-20),
// so they contribute exactly 0 like the oracle's r > 0 condition.
// sorted_xyzw.w (the mass) must be uniform: nb_cell_sort writes the scalar `mass` there.
#define P3M3_WARPS 4
#define P3M3_CAP 768     // candidate tile capacity per warp (multiple of 8); typical 27-cell load ~620
// the 8 i-particles of this i-block (lane & 7) against the `npad` (multiple of 8) tile slots
__device__ __forceinline__ void p3m3_process(const float* __restrict__ tx, const float* __restrict__ ty,
                                             const float* __restrict__ tz, int npad, int strm, p3m_f32x2 nxi,
                                             p3m_f32x2 nyi, p3m_f32x2 nzi, const P3m3Packed& k, double& ax,
                                             double& ay, double& az) {
  p3m_f32x2 FX = 0ull, FY = 0ull, FZ = 0ull;   // (+0.f, +0.f)
#pragma unroll 2
  for (int v = 2 * strm; v < npad; v += 8) {
    const p3m_f32x2 X = *reinterpret_cast<const p3m_f32x2*>(tx + v);
    const p3m_f32x2 Y = *reinterpret_cast<const p3m_f32x2*>(ty + v);
    const p3m_f32x2 Z = *reinterpret_cast<const p3m_f32x2*>(tz + v);
    const p3m_f32x2 dx = p3m_add2(X, nxi), dy = p3m_add2(Y, nyi), dz = p3m_add2(Z, nzi);
    p3m_f32x2 r2 = p3m_mul2(dx, dx);
    r2 = p3m_fma2(dy, dy, r2);
    r2 = p3m_fma2(dz, dz, r2);
    float r2a, r2b;
    p3m_unpack2(r2, r2a, r2b);
    const bool in0 = r2a < k.rc2, in1 = r2b < k.rc2;
    const p3m_f32x2 rinv = p3m_rsqrt2(p3m_add2(r2, k.eps2));
    const p3m_f32x2 w = p3m_mul2(r2, k.ws);                          // u²
    const p3m_f32x2 u = p3m_mul2(w, p3m_rsqrt2(p3m_add2(w, k.tiny)));
    const p3m_f32x2 tt = p3m_fma2(w, k.ts, k.m1);
    p3m_f32x2 q = k.nq[P3M_DEG];
#pragma unroll
    for (int d = P3M_DEG - 1; d >= 0; --d) q = p3m_fma2(q, tt, k.nq[d]);
    const p3m_f32x2 f = p3m_fma2(p3m_mul2(w, u), q, k.one);          // erfc(u) + 2u/sqrt(pi) exp(-u²)
    const p3m_f32x2 s = p3m_mul2(p3m_mul2(p3m_mul2(rinv, rinv), p3m_mul2(rinv, k.wj)), f);
    float s0, s1;
    p3m_unpack2(s, s0, s1);
    const p3m_f32x2 sm = p3m_pack2(in0 ? s0 : 0.f, in1 ? s1 : 0.f);
    FX = p3m_fma2(sm, dx, FX);
    FY = p3m_fma2(sm, dy, FY);
    FZ = p3m_fma2(sm, dz, FZ);
  }
  float a, b;
  p3m_unpack2(FX, a, b); ax += (double)(a + b);
  p3m_unpack2(FY, a, b); ay += (double)(a + b);
  p3m_unpack2(FZ, a, b); az += (double)(a + b);
}

__global__ void __launch_bounds__(P3M3_WARPS * 32, 4)
p3m_short_v3_kernel(const float4* __restrict__ sorted, const int* __restrict__ cell_start,
                    const int* __restrict__ perm, int ncs, int64_t cell_lo, int64_t cell_hi, P3mParams prm,
                    double scale, double* __restrict__ acc, int accumulate, int i_limit, int cap) {
  extern __shared__ __align__(16) float p3m3_smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int li = lane & 7, strm = lane >> 3;
  float* tx = p3m3_smem + warp * (3 * cap);   // cap: tile capacity per warp, a multiple of 8
  float* ty = tx + cap;
  float* tz = ty + cap;
  P3m3Packed k;
  {
    const float e2 = fmaxf(prm.eps2, 1e-20f);
    const float wj = __ldg(&sorted[0].w);   // uniform mass (nb_cell_sort)
    k.eps2 = p3m_pack2(e2, e2);
    k.ws = p3m_pack2(prm.w_scale, prm.w_scale);
    k.tiny = p3m_pack2(1e-30f, 1e-30f);
    k.ts = p3m_pack2(prm.t_scale, prm.t_scale);
    k.m1 = p3m_pack2(-1.f, -1.f);
    k.one = p3m_pack2(1.f, 1.f);
    k.wj = p3m_pack2(wj, wj);
#pragma unroll
    for (int d = 0; d <= P3M_DEG; ++d) k.nq[d] = p3m_pack2(-prm.coef[d], -prm.coef[d]);
    k.rc2 = prm.rc2;
  }
  const int64_t wstride = (int64_t)gridDim.x * P3M3_WARPS;
  for (int64_t c = cell_lo + (int64_t)blockIdx.x * P3M3_WARPS + warp; c < cell_hi; c += wstride) {
    const int is = cell_start[c], ie = cell_start[c + 1];
    if (is == ie) continue;
    const int cz = (int)(c % ncs);
    const int cy = (int)((c / ncs) % ncs);
    const int cx = (int)(c / ((int64_t)ncs * ncs));
    // lanes 0..26 look up the j-ranges of the 27 neighbour cells
    int my_js = 0, my_je = 0;
    if (lane < 27) {
      int nx = cx + lane / 9 - 1, ny = cy + (lane / 3) % 3 - 1, nz = cz + lane % 3 - 1;
      nx = nx < 0 ? nx + ncs : (nx >= ncs ? nx - ncs : nx);
      ny = ny < 0 ? ny + ncs : (ny >= ncs ? ny - ncs : ny);
      nz = nz < 0 ? nz + ncs : (nz >= ncs ? nz - ncs : nz);
      const int64_t n = ((int64_t)nx * ncs + ny) * ncs + nz;
      my_js = cell_start[n];
      my_je = cell_start[n + 1];
    }
    int ncand = my_je - my_js;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) ncand += __shfl_xor_sync(0xffffffffu, ncand, o);
    const bool single = ncand <= cap;   // whole neighbourhood fits: fill once per cell
    int resident = 0;                        // padded slot count of the resident tile (0 = not yet filled)
    for (int ib = is; ib < ie; ib += 8) {
      const int i = ib + li;
      // particles with perm >= i_limit are halo copies of another slab's particles: sources only
      const int oi = i < ie ? perm[i] : 0x7fffffff;
      const bool ivalid = oi < i_limit;
      if (!__any_sync(0xffffffffu, ivalid)) continue;
      const float4 pi = sorted[i < ie ? i : is];
      const p3m_f32x2 nxi = p3m_pack2(-pi.x, -pi.x), nyi = p3m_pack2(-pi.y, -pi.y), nzi = p3m_pack2(-pi.z, -pi.z);
      double ax = 0.0, ay = 0.0, az = 0.0;
      if (resident) {
        p3m3_process(tx, ty, tz, resident, strm, nxi, nyi, nzi, k, ax, ay, az);
      } else {
        int fill = 0;
        for (int n = 0; n < 27; ++n) {
          int js = __shfl_sync(0xffffffffu, my_js, n);
          const int je = __shfl_sync(0xffffffffu, my_je, n);
          const float offx = (float)(n / 9 - 1) * prm.cs;
          const float offy = (float)((n / 3) % 3 - 1) * prm.cs;
          const float offz = (float)(n % 3 - 1) * prm.cs;
          while (js < je) {
            const int take = min(je - js, cap - fill);
            for (int t = lane; t < take; t += 32) {
              const float4 p = sorted[js + t];
              tx[fill + t] = p.x + offx;
              ty[fill + t] = p.y + offy;
              tz[fill + t] = p.z + offz;
            }
            fill += take;
            js += take;
            if (fill == cap) {   // only when !single
              __syncwarp();
              p3m3_process(tx, ty, tz, fill, strm, nxi, nyi, nzi, k, ax, ay, az);
              __syncwarp();
              fill = 0;
            }
          }
        }
        const int npad = (fill + 7) & ~7;
        if (lane < npad - fill) {   // sentinels: far outside any cutoff
          tx[fill + lane] = 1e18f;
          ty[fill + lane] = 0.f;
          tz[fill + lane] = 0.f;
        }
        __syncwarp();
        if (npad > 0) p3m3_process(tx, ty, tz, npad, strm, nxi, nyi, nzi, k, ax, ay, az);
        __syncwarp();
        if (single) resident = npad;
      }
      // fold the 4 j-streams of each i-particle
      ax += __shfl_xor_sync(0xffffffffu, ax, 8);  ay += __shfl_xor_sync(0xffffffffu, ay, 8);
      az += __shfl_xor_sync(0xffffffffu, az, 8);
      ax += __shfl_xor_sync(0xffffffffu, ax, 16); ay += __shfl_xor_sync(0xffffffffu, ay, 16);
      az += __shfl_xor_sync(0xffffffffu, az, 16);
      if (ivalid && strm == 0) {
        const int64_t o = oi;
        if (accumulate) {
          acc[3 * o] += ax * scale; acc[3 * o + 1] += ay * scale; acc[3 * o + 2] += az * scale;
        } else {
          acc[3 * o] = ax * scale; acc[3 * o + 1] = ay * scale; acc[3 * o + 2] = az * scale;
        }
      }
    }
    __syncwarp();   // the next cell overwrites the tile
  }
}

// ---------------------------------------------------------------------------------------
// v5 — v4's pair loop on a SLIDING z-window tile.
//
// ncu on v4 (profiles/r1d_ncu_full_p3m_short_v4_256.txt): the tile fill is 11 % of the
// instructions but 24 % of the stall samples (dependent global loads at 14 warps/SM), and every
// cell re-loads all 27 neighbour cells although the next cell along z shares 18 of them.
// v5 gives a warp a z-run of P3M5_SEG cells of one (cx,cy) column.  The tile is three regions,
// one per z-slab (the 3x3 cells at one z), each padded to whole 32-slot groups:
//   first cell of the run : fill the three slabs cz-1, cz, cz+1;
//   every further cell    : subtract cs from the z of the two slabs that stay (about 13 LDS /
//                           FADD / STS per warp; the error stays one rounding of a value < 2cs,
//                           what neighbour-cell candidates always carried), and fill ONLY the
//                           new slab cz+2 into the region of the evicted one;
// i.e. 1.1 slab fills per cell instead of 3.  Phase 1 / phase 2 are v4's (p3m4_eval2, lane-
// private queues), run region by region with one queue.  A cell whose slab does not fit a
// region (strong clustering) is handled by v4's chunked loop over the 27 runs on the same
// tile memory, after which the window is rebuilt.
#define P3M5_WARPS 4
#define P3M5_SEG 16

// fill of one slab: the 9 cells (cx+ox, cy+oy, zs) as one flattened list -> tile slots
// [0, nt) of a region, translated by (ox*cs, oy*cs, oz*cs) (offs = offtab + 3*(oz+1), run r
// at offs[9*r ..]); pads to a multiple of 32 with far-away sentinels.
__device__ __forceinline__ void p3m5_fill_slab(const float4* __restrict__ sorted, const int* __restrict__ roff,
                                               const int* __restrict__ rjs, const float* __restrict__ offs,
                                               float* __restrict__ tx, float* __restrict__ ty,
                                               float* __restrict__ tz, int nt, int lane) {
  for (int t0 = lane; t0 < nt + lane; t0 += 128) {   // warp-uniform trip count
    int n[4];
    float4 p[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int g = min(t0 + 32 * j, nt - 1);
      int m = 0;   // largest m with roff[m] <= g  (roff[9..31] = INT_MAX)
#pragma unroll
      for (int s = 8; s > 0; s >>= 1) m = (roff[m + s] <= g) ? m + s : m;
      n[j] = m;
      p[j] = sorted[rjs[m] + (g - roff[m])];
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int t = t0 + 32 * j;
      if (t < nt) {
        tx[t] = p[j].x + offs[9 * n[j]];
        ty[t] = p[j].y + offs[9 * n[j] + 1];
        tz[t] = p[j].z + offs[9 * n[j] + 2];
      }
    }
  }
  const int npad = (nt + 31) & ~31;
  if (lane < npad - nt) {
    tx[nt + lane] = 1e18f;
    ty[nt + lane] = 0.f;
    tz[nt + lane] = 0.f;
  }
}

// phase 1 over the three regions (p0, p1, p2 padded slots, multiples of 32) as ONE loop over
// their 32-slot groups — a single copy of the test loop and of the drain keeps the kernel in
// the instruction cache (three inlined copies measured 141 ms against v4's 90 ms).
template <int QD>
__device__ __forceinline__ void p3m5_pass_all(const float* __restrict__ tx, const float* __restrict__ ty,
                                              const float* __restrict__ tz, unsigned short* __restrict__ qcol,
                                              int caps, int p0, int p1, int p2, int strm, p3m_f32x2 nxi,
                                              p3m_f32x2 nyi, p3m_f32x2 nzi, const P3m3Packed& k, unsigned qbase,
                                              double& ax, double& ay, double& az) {
  unsigned qa = qbase;
  const unsigned qfull = qbase + (QD - 8) * 64;   // a group pushes at most 8 entries per lane
#pragma unroll 1
  for (int r = 0; r < 3; ++r) {
    const int vend = r * caps + (r == 0 ? p0 : (r == 1 ? p1 : p2));
    for (int vb = r * caps + 4 * strm; vb < vend; vb += 32) {
      p3m_f32x2 r2[4];
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const int v = vb + 16 * j;
        const float4 X = *reinterpret_cast<const float4*>(tx + v);
        const float4 Y = *reinterpret_cast<const float4*>(ty + v);
        const float4 Z = *reinterpret_cast<const float4*>(tz + v);
        const p3m_f32x2 dx0 = p3m_add2(p3m_pack2(X.x, X.y), nxi), dx1 = p3m_add2(p3m_pack2(X.z, X.w), nxi);
        const p3m_f32x2 dy0 = p3m_add2(p3m_pack2(Y.x, Y.y), nyi), dy1 = p3m_add2(p3m_pack2(Y.z, Y.w), nyi);
        const p3m_f32x2 dz0 = p3m_add2(p3m_pack2(Z.x, Z.y), nzi), dz1 = p3m_add2(p3m_pack2(Z.z, Z.w), nzi);
        r2[2 * j] = p3m_fma2(dz0, dz0, p3m_fma2(dy0, dy0, p3m_mul2(dx0, dx0)));
        r2[2 * j + 1] = p3m_fma2(dz1, dz1, p3m_fma2(dy1, dy1, p3m_mul2(dx1, dx1)));
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int v = vb + 16 * (j >> 1) + 2 * (j & 1);
        float r2a, r2b;
        p3m_unpack2(r2[j], r2a, r2b);
        if (r2a < k.rc2) { p3m_sts_u16(qa, (unsigned short)v); qa += 64; }
        if (r2b < k.rc2) { p3m_sts_u16(qa, (unsigned short)(v + 1)); qa += 64; }
      }
      if (__any_sync(0xffffffffu, qa > qfull)) {
        p3m4_drain(tx, ty, tz, qcol, (int)((qa - qbase) >> 6), nxi, nyi, nzi, k, ax, ay, az);
        qa = qbase;
      }
    }
  }
  if (__any_sync(0xffffffffu, qa != qbase))
    p3m4_drain(tx, ty, tz, qcol, (int)((qa - qbase) >> 6), nxi, nyi, nzi, k, ax, ay, az);
}

// One cell by v4's chunked loop over the flattened 27-run list (tile = the whole 3*caps slots).
// Out of line: only cells whose slab overflows a region come here.
template <int QD>
__device__ __noinline__ void p3m5_cell_chunked(const float4* __restrict__ sorted, const int* __restrict__ cell_start,
                                               const int* __restrict__ perm, int ncs, int cx, int cy, int cz, int is,
                                               int ie, const float* __restrict__ offtab,
                                               const P3mParams* __restrict__ sprm, float* __restrict__ tx,
                                               float* __restrict__ ty, float* __restrict__ tz, int* __restrict__ roff,
                                               int* __restrict__ rjs, unsigned short* __restrict__ qcol, int cap,
                                               double scale, double* __restrict__ acc, int accumulate, int i_limit) {
  const int lane = threadIdx.x & 31, li = lane & 7, strm = lane >> 3;
  const P3m3Packed k = p3m_make_packed(*sprm, __ldg(&sorted[0].w));
  int my_js = 0, my_len = 0;
  if (lane < 27) {
    int nx = cx + lane / 9 - 1, ny = cy + (lane / 3) % 3 - 1, nz = cz + lane % 3 - 1;
    nx = nx < 0 ? nx + ncs : (nx >= ncs ? nx - ncs : nx);
    ny = ny < 0 ? ny + ncs : (ny >= ncs ? ny - ncs : ny);
    nz = nz < 0 ? nz + ncs : (nz >= ncs ? nz - ncs : nz);
    const int64_t n = ((int64_t)nx * ncs + ny) * ncs + nz;
    my_js = cell_start[n];
    my_len = cell_start[n + 1] - my_js;
  }
  int incl = my_len;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int up = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += up;
  }
  const int ncand = __shfl_sync(0xffffffffu, incl, 31);
  __syncwarp();
  roff[lane] = lane < 27 ? incl - my_len : 0x7fffffff;
  rjs[lane] = my_js;
  __syncwarp();
  bool resident = false;
  for (int ib = is; ib < ie; ib += 8) {
    const int i = ib + li;
    const int oi = i < ie ? perm[i] : 0x7fffffff;
    const bool ivalid = oi < i_limit;
    if (!__any_sync(0xffffffffu, ivalid)) continue;
    float4 pi = sorted[i < ie ? i : is];
    if (!ivalid) pi.x = 3e18f;
    const p3m_f32x2 nxi = p3m_pack2(-pi.x, -pi.x), nyi = p3m_pack2(-pi.y, -pi.y), nzi = p3m_pack2(-pi.z, -pi.z);
    double ax = 0.0, ay = 0.0, az = 0.0;
    for (int first = 0; first < ncand; first += cap) {
      const int nt = min(cap, ncand - first);
      if (!resident) {
        __syncwarp();   // every lane is done reading the previous chunk
        p3m4_fill(sorted, roff, rjs, offtab, tx, ty, tz, first, nt, lane);
      }
      p3m4_pass<QD>(tx, ty, tz, qcol, (nt + 31) & ~31, strm, nxi, nyi, nzi, k, ax, ay, az);
    }
    resident = ncand <= cap;   // whole neighbourhood fits: filled once per cell
    ax += __shfl_xor_sync(0xffffffffu, ax, 8);  ay += __shfl_xor_sync(0xffffffffu, ay, 8);
    az += __shfl_xor_sync(0xffffffffu, az, 8);
    ax += __shfl_xor_sync(0xffffffffu, ax, 16); ay += __shfl_xor_sync(0xffffffffu, ay, 16);
    az += __shfl_xor_sync(0xffffffffu, az, 16);
    if (ivalid && strm == 0) p3m5_store_acc(acc, oi, ax, ay, az, scale, accumulate);
  }
  __syncwarp();
}

// per-warp shared memory of v5: tile (3 regions x caps slots x 3 coordinates), run table, queue
__host__ __device__ inline size_t p3m5_warp_bytes(int caps, int qd) { return p3m4_warp_bytes(3 * caps, qd); }

template <int QD>
__global__ void __launch_bounds__(P3M5_WARPS * 32, 4)
p3m_short_v5_kernel(const float4* __restrict__ sorted, const int* __restrict__ cell_start,
                    const int* __restrict__ perm, int ncs, int64_t cell_lo, int64_t cell_hi, P3mParams prm,
                    double scale, double* __restrict__ acc, int accumulate, int i_limit, int caps) {
  extern __shared__ __align__(16) unsigned char p3m5_smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int li = lane & 7, strm = lane >> 3;
  const int cap = 3 * caps;
  float* offtab = reinterpret_cast<float*>(p3m5_smem);                         // 27*3 floats (96 reserved)
  P3mParams* sprm = reinterpret_cast<P3mParams*>(offtab + 96);                 // copy for the out-of-line path
  unsigned char* wbase = p3m5_smem + 128 * sizeof(float) + (size_t)warp * p3m5_warp_bytes(caps, QD);
  float* tx = reinterpret_cast<float*>(wbase);
  float* ty = tx + cap;
  float* tz = ty + cap;
  int* roff = reinterpret_cast<int*>(tz + cap);
  int* rjs = roff + 32;
  unsigned short* qcol = reinterpret_cast<unsigned short*>(rjs + 32) + lane;
  if (threadIdx.x < 27) {
    const int n = threadIdx.x;
    offtab[3 * n] = (float)(n / 9 - 1) * prm.cs;
    offtab[3 * n + 1] = (float)((n / 3) % 3 - 1) * prm.cs;
    offtab[3 * n + 2] = (float)(n % 3 - 1) * prm.cs;
  }
  if (threadIdx.x == 32) *sprm = prm;
  for (int r = 0; r < QD + 4; ++r) qcol[r * 32] = 0;   // stale reads must be valid tile indices
  __syncthreads();
  const P3m3Packed k = p3m_make_packed(prm, __ldg(&sorted[0].w));
  const unsigned qbase = (unsigned)__cvta_generic_to_shared(qcol);
  const float cs = prm.cs;

  const int nseg = (ncs + P3M5_SEG - 1) / P3M5_SEG;
  const int64_t col_lo = cell_lo / ncs, col_hi = (cell_hi + ncs - 1) / ncs;
  const int64_t nunits = (col_hi - col_lo) * nseg;
  for (int64_t u = (int64_t)blockIdx.x * P3M5_WARPS + warp; u < nunits; u += (int64_t)gridDim.x * P3M5_WARPS) {
    const int64_t col = col_lo + u / nseg;
    const int seg = (int)(u % nseg);
    const int cx = (int)(col / ncs), cy = (int)(col % ncs);
    const int zb = seg * P3M5_SEG, ze = min(ncs, zb + P3M5_SEG);
    bool win = false;            // regions hold the slabs cz-1, cz, cz+1 of the PREVIOUS cell
    int n0 = 0, n1 = 0, n2 = 0;  // candidates per region (region of slab zs = (zs + 3) % 3)
    int c0 = 0, c1 = 0, c2 = 0;  // slot of the centre column's cell inside each region
    for (int cz = zb; cz < ze; ++cz) {
      const int64_t c = col * ncs + cz;
      if (c < cell_lo || c >= cell_hi) { win = false; continue; }
      const int is = cell_start[c], ie = cell_start[c + 1];
      if (is == ie) { win = false; continue; }
      __syncwarp();   // every lane is done with the previous cell's tile
      if (win) {      // slide: the slabs of cz-1 and cz move one cell down in the new frame
#pragma unroll
        for (int d = -1; d <= 0; ++d) {
          const int reg = (cz + d + 3) % 3;
          const int nr = reg == 0 ? n0 : (reg == 1 ? n1 : n2);
          float* z = tz + reg * caps;
          for (int t = lane; t < nr; t += 32) z[t] -= cs;
        }
      }
      bool fits = true;
      for (int oz = win ? 1 : -1; oz <= 1; ++oz) {
        const int zs = cz + oz;
        const int reg = (zs + 3) % 3;
        int my_js = 0, my_len = 0;
        if (lane < 9) {
          int nx = cx + lane / 3 - 1, ny = cy + lane % 3 - 1, nz = zs;
          nx = nx < 0 ? nx + ncs : (nx >= ncs ? nx - ncs : nx);
          ny = ny < 0 ? ny + ncs : (ny >= ncs ? ny - ncs : ny);
          nz = nz < 0 ? nz + ncs : (nz >= ncs ? nz - ncs : nz);
          const int64_t n = ((int64_t)nx * ncs + ny) * ncs + nz;
          my_js = cell_start[n];
          my_len = cell_start[n + 1] - my_js;
        }
        int incl = my_len;
#pragma unroll
        for (int o = 1; o < 16; o <<= 1) {
          const int up = __shfl_up_sync(0xffffffffu, incl, o);
          if (lane >= o) incl += up;
        }
        const int nt = __shfl_sync(0xffffffffu, incl, 15);
        if (nt > caps) { fits = false; break; }
        const int cen = __shfl_sync(0xffffffffu, incl - my_len, 4);   // first slot of the centre column's cell
        if (reg == 0) c0 = cen; else if (reg == 1) c1 = cen; else c2 = cen;
        __syncwarp();   // the previous slab's fill is done with the run table
        roff[lane] = lane < 9 ? incl - my_len : 0x7fffffff;
        rjs[lane] = my_js;
        __syncwarp();
        if (nt > 0)
          p3m5_fill_slab(sorted, roff, rjs, offtab + 3 * (oz + 1), tx + reg * caps, ty + reg * caps, tz + reg * caps,
                         nt, lane);
        if (reg == 0) n0 = nt; else if (reg == 1) n1 = nt; else n2 = nt;
      }
      if (!fits) {
        p3m5_cell_chunked<QD>(sorted, cell_start, perm, ncs, cx, cy, cz, is, ie, offtab, sprm, tx, ty, tz, roff, rjs,
                              qcol, cap, scale, acc, accumulate, i_limit);
        win = false;
        continue;
      }
      win = true;
      __syncwarp();   // tile complete
      const int p0 = (n0 + 31) & ~31, p1 = (n1 + 31) & ~31, p2 = (n2 + 31) & ~31;
      const int rown = cz % 3;
      const float* own_z = tz + rown * caps + (rown == 0 ? c0 : (rown == 1 ? c1 : c2));
      for (int ib = is; ib < ie; ib += 8) {
        const int i = ib + li;
        // particles with perm >= i_limit are halo copies of another slab's particles: sources only
        const int oi = i < ie ? perm[i] : 0x7fffffff;
        const bool ivalid = oi < i_limit;
        if (!__any_sync(0xffffffffu, ivalid)) continue;
        float4 pi = sorted[i < ie ? i : is];
        // the i-particle takes its z from its OWN tile slot: a slab that slid twice carries
        // (z + cs) - cs, which can differ from z by a rounding of ~1e-7 cs; reading it back keeps
        // the self pair (and coincident points) at exactly zero separation.  x and y of the
        // centre column are stored untranslated.
        pi.z = own_z[min(i, ie - 1) - is];
        if (!ivalid) pi.x = 3e18f;   // idle lanes never pass the range test (no queue traffic)
        const p3m_f32x2 nxi = p3m_pack2(-pi.x, -pi.x), nyi = p3m_pack2(-pi.y, -pi.y), nzi = p3m_pack2(-pi.z, -pi.z);
        double ax = 0.0, ay = 0.0, az = 0.0;
        p3m5_pass_all<QD>(tx, ty, tz, qcol, caps, p0, p1, p2, strm, nxi, nyi, nzi, k, qbase, ax, ay, az);
        // fold the 4 j-streams of each i-particle
        ax += __shfl_xor_sync(0xffffffffu, ax, 8);  ay += __shfl_xor_sync(0xffffffffu, ay, 8);
        az += __shfl_xor_sync(0xffffffffu, az, 8);
        ax += __shfl_xor_sync(0xffffffffu, ax, 16); ay += __shfl_xor_sync(0xffffffffu, ay, 16);
        az += __shfl_xor_sync(0xffffffffu, az, 16);
        if (ivalid && strm == 0) p3m5_store_acc(acc, oi, ax, ay, az, scale, accumulate);
      }
    }
  }
}

// Mask form of the v6 pass (experimental, NB_P3M_MASK=1): phase 1 does not push slot indices.  The 8 range
// tests of a lane and group are the SIGN BITS of r² - r_cut² (4 packed adds), shifted together
// into one byte with 8 funnel shifts and stored at qmask[group][lane] — 14 instructions instead of
// 8 compares + 16 predicated push instructions + 8 index materialisations.  Phase 2 walks the
// lane's bytes (find-first-set, clear, next byte when empty) and evaluates four hits per round
// exactly like p3m4_drain.  One byte per group also means the whole window fits (52 groups x 32 B
// per warp instead of a 3.8 KB index queue), so there is a single drain per i-block.
__device__ __forceinline__ int p3m7_next(const unsigned char* __restrict__ qm, int& g, int ng, unsigned& bits) {
  while (bits == 0u) {
    if (g >= ng) return -1;
    bits = qm[g * 32];
    ++g;
  }
  const int b = __ffs((int)bits) - 1;
  bits &= bits - 1u;
  return 32 * (g - 1) + (b & 3) + 16 * (b >> 2);   // slot offset inside the lane's quads of group g-1
}

__device__ __forceinline__ void p3m7_pass(const float* __restrict__ tx, const float* __restrict__ ty,
                                          const float* __restrict__ tz, unsigned txs, unsigned cap4,
                                          unsigned char* __restrict__ qm, int vbeg, int vend, int strm,
                                          p3m_f32x2 nxi, p3m_f32x2 nyi, p3m_f32x2 nzi, const P3m3Packed& k,
                                          double& ax, double& ay, double& az) {
  const int vfirst = vbeg + 4 * strm;
  unsigned sx = txs + 4u * (unsigned)vfirst;
  const unsigned send = txs + 4u * (unsigned)vend;
  const p3m_f32x2 nrc2 = p3m_pack2(-k.rc2, -k.rc2);
  unsigned qs = (unsigned)__cvta_generic_to_shared(qm);
  for (; sx < send; sx += 128u, qs += 32u) {
    const unsigned sy = sx + cap4, sz = sy + cap4;
    p3m_f32x2 d[4];
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const float4 X = p3m_lds128(sx + 64u * j);
      const float4 Y = p3m_lds128(sy + 64u * j);
      const float4 Z = p3m_lds128(sz + 64u * j);
      const p3m_f32x2 dx0 = p3m_add2(p3m_pack2(X.x, X.y), nxi), dx1 = p3m_add2(p3m_pack2(X.z, X.w), nxi);
      const p3m_f32x2 dy0 = p3m_add2(p3m_pack2(Y.x, Y.y), nyi), dy1 = p3m_add2(p3m_pack2(Y.z, Y.w), nyi);
      const p3m_f32x2 dz0 = p3m_add2(p3m_pack2(Z.x, Z.y), nzi), dz1 = p3m_add2(p3m_pack2(Z.z, Z.w), nzi);
      d[2 * j] = p3m_add2(p3m_fma2(dz0, dz0, p3m_fma2(dy0, dy0, p3m_mul2(dx0, dx0))), nrc2);
      d[2 * j + 1] = p3m_add2(p3m_fma2(dz1, dz1, p3m_fma2(dy1, dy1, p3m_mul2(dx1, dx1))), nrc2);
    }
    unsigned m = 0u;   // bit b <-> slot (b & 3) + 16 * (b >> 2): d[3].hi first, d[0].lo last
#pragma unroll
    for (int j = 3; j >= 0; --j) {
      float lo, hi;
      p3m_unpack2(d[j], lo, hi);
      m = __funnelshift_l(__float_as_uint(hi), m, 1);
      m = __funnelshift_l(__float_as_uint(lo), m, 1);
    }
    asm volatile("st.shared.u8 [%0], %1;" ::"r"(qs), "r"(m) : "memory");
  }
  // phase 2: four hits per lane and round (two independent packed chains)
  const int ng = (vend - vbeg) >> 5;
  int g = 0;
  unsigned bits = 0u;
  p3m_f32x2 FX = 0ull, FY = 0ull, FZ = 0ull, GX = 0ull, GY = 0ull, GZ = 0ull;
  __syncwarp();
  while (true) {
    const int o0 = p3m7_next(qm, g, ng, bits), o1 = p3m7_next(qm, g, ng, bits);
    const int o2 = p3m7_next(qm, g, ng, bits), o3 = p3m7_next(qm, g, ng, bits);
    if (!__any_sync(0xffffffffu, o0 >= 0)) break;
    p3m4_eval2(tx, ty, tz, vfirst + max(o0, 0), vfirst + max(o1, 0), o0 >= 0, o1 >= 0, nxi, nyi, nzi, k, FX, FY, FZ);
    p3m4_eval2(tx, ty, tz, vfirst + max(o2, 0), vfirst + max(o3, 0), o2 >= 0, o3 >= 0, nxi, nyi, nzi, k, GX, GY, GZ);
  }
  float a, b, c, e;
  p3m_unpack2(FX, a, b); p3m_unpack2(GX, c, e); ax += (double)((a + b) + (c + e));
  p3m_unpack2(FY, a, b); p3m_unpack2(GY, c, e); ay += (double)((a + b) + (c + e));
  p3m_unpack2(FZ, a, b); p3m_unpack2(GZ, c, e); az += (double)((a + b) + (c + e));
}

