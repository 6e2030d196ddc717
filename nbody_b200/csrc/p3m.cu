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
#include <stdlib.h>

#include <type_traits>

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


typedef unsigned long long p3m_f32x2;
__device__ __forceinline__ p3m_f32x2 p3m_pack2(float lo, float hi) {
  p3m_f32x2 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void p3m_unpack2(p3m_f32x2 v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ p3m_f32x2 p3m_add2(p3m_f32x2 a, p3m_f32x2 b) {
  p3m_f32x2 d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ p3m_f32x2 p3m_mul2(p3m_f32x2 a, p3m_f32x2 b) {
  p3m_f32x2 d;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ p3m_f32x2 p3m_fma2(p3m_f32x2 a, p3m_f32x2 b, p3m_f32x2 c) {
  p3m_f32x2 d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}


struct P3m3Packed {
  p3m_f32x2 eps2, ws, tiny, ts, m1, one, wj;
  p3m_f32x2 nq[P3M_DEG + 1];   // NEGATED split coefficients: f = 1 + w·u·nQ(t)
  float rc2;
};

__device__ __forceinline__ p3m_f32x2 p3m_rsqrt2(p3m_f32x2 v) {
  float a, b;
  p3m_unpack2(v, a, b);
  return p3m_pack2(p3m_rsqrt(a), p3m_rsqrt(b));
}

#ifdef NB_AB_KERNELS
// v3's evaluate-all loop (p3m_ab.cuh), used by the A/B form of the v4 kernel (NB_P3M_KERNEL=5)
__device__ __forceinline__ void p3m3_process(const float* __restrict__ tx, const float* __restrict__ ty,
                                             const float* __restrict__ tz, int npad, int strm, p3m_f32x2 nxi,
                                             p3m_f32x2 nyi, p3m_f32x2 nzi, const P3m3Packed& k, double& ax,
                                             double& ay, double& az);
#endif

// ---------------------------------------------------------------------------------------
// v4 — v3's lane mapping with a two-phase pair loop: test all, evaluate in-range only.
//
// ncu on v3 (profiles/r1d_*): the FMA pipe is the limiter (63 % active) and 29 of its lane-ops
// per candidate are spent on pairs of which only 16 % are inside r_cut.  v4 keeps the tile and
// the 8 i x 4 j-stream mapping and splits the loop:
//   fill      the candidates of the 27 neighbour cells form ONE flattened list (exclusive
//             prefix of the 27 cell populations by warp shuffle, kept in shared memory);
//             lane t loads candidate t, t+32, ... after a 5-step binary search for its cell —
//             independent loads (the v3 fill ran 27 dependent load->store rounds per cell);
//   phase 1   per lane and iteration TWO candidates: 3 LDS.64, 6 packed FP32 ops for r², two
//             compares; an in-range candidate's tile index (u16) is pushed on the lane's
//             PRIVATE queue column in shared memory — no ballot, no prefix, no atomics;
//   phase 2   when any lane's queue is nearly full, and at the end of the i-block: every lane
//             pops two entries per round, re-reads the two j-particles from the tile and runs
//             the packed evaluation of v3 (rounds = max over lanes of its hit count / 2);
// so the expensive part runs ~6x less often.  Queue columns are lane-private (bank = lane/2,
// u16 entries), stale entries beyond a lane's count are valid tile indices and are masked.
#define P3M4_WARPS 4
#define P3M4_QD 64       // queue depth per lane (entries); drained when a lane exceeds QD - 16
#define P3M4_CAP 768     // tile capacity per warp (candidates; multiple of 32)

// per-warp shared memory: tile (3·cap floats), run table (2·32 ints), queue (QD·32 u16)
__host__ __device__ inline size_t p3m4_warp_bytes(int cap, int qd) {
  return sizeof(float) * 3 * (size_t)cap + sizeof(int) * 64 + sizeof(unsigned short) * (size_t)(qd + 4) * 32;
}

// Load the flattened candidates [first, first + nt) into tile slots [0, nt), translated into
// the frame of the i-cell, and pad with far-away sentinels to a multiple of 32 (one group of
// phase 1).  Four rounds of 32 candidates are handled together so that their binary searches
// and global loads overlap (no early exit between them: out-of-range lanes re-load the last
// candidate and skip the store).
__device__ __forceinline__ void p3m4_fill(const float4* __restrict__ sorted, const int* __restrict__ roff,
                                          const int* __restrict__ rjs, const float* __restrict__ offtab,
                                          float* __restrict__ tx, float* __restrict__ ty, float* __restrict__ tz,
                                          int first, int nt, int lane) {
  for (int t0 = lane; t0 < nt + lane; t0 += 128) {   // warp-uniform trip count
    int n[4];
    float4 p[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int g = first + min(t0 + 32 * j, nt - 1);
      int m = 0;   // largest m with roff[m] <= g  (roff[27..31] = INT_MAX)
#pragma unroll
      for (int s = 16; s > 0; s >>= 1) m = (roff[m + s] <= g) ? m + s : m;
      n[j] = m;
      p[j] = sorted[rjs[m] + (g - roff[m])];
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int t = t0 + 32 * j;
      if (t < nt) {
        tx[t] = p[j].x + offtab[3 * n[j]];
        ty[t] = p[j].y + offtab[3 * n[j] + 1];
        tz[t] = p[j].z + offtab[3 * n[j] + 2];
      }
    }
  }
  const int npad = (nt + 31) & ~31;
  if (lane < npad - nt) {
    tx[nt + lane] = 1e18f;
    ty[nt + lane] = 0.f;
    tz[nt + lane] = 0.f;
  }
  __syncwarp();
}

__device__ __forceinline__ void p3m_sts_u16(unsigned saddr, unsigned short v) {
  asm volatile("st.shared.u16 [%0], %1;" ::"r"(saddr), "h"(v) : "memory");
}

// ---------------------------------------------------------------------------------------
// Near-pair list (nb_p3m_short_exact).  Tile coordinates are float32 in a frame of up to 4 cells, so a separation
// carries a quantum of ~2.4e-7 x 4 r_cut; a pair at r of a few softening lengths turns that into 1e-4..1e-3 of the
// rms force.  With a list, phase 2 does not evaluate an in-range pair with 0 < r^2 < rn2: it appends
// (sorted slot of i, sorted slot of j) to a global list and a float64 kernel evaluates those pairs afterwards from
// the original positions.  The decision is made once, here, so no pair is counted twice or lost.  When the list is
// full the pair is evaluated in float32 as before (count keeps counting: the caller can read the overflow).
struct P3mNearList {
  int2* list;
  int* count;
  int cap;
  float rn2;
  float rb2;    // pairs with r^2 >= rb2 (the last float32 quanta below the widened cut-off) are listed too: the float64
                // kernel applies the cut-off to them, so a pair is inside r_cut for both partners or for neither
};
struct P3mNoNear {};   // tag: kernel generation without a list

// the packed pair term of two candidates: s = w_j (r^2 + eps^2)^(-3/2) (erfc(u) + 2u/sqrt(pi) exp(-u^2)) and the
// separations it multiplies
__device__ __forceinline__ void p3m4_term2(const float* __restrict__ tx, const float* __restrict__ ty,
                                           const float* __restrict__ tz, int v0, int v1, p3m_f32x2 nxi, p3m_f32x2 nyi,
                                           p3m_f32x2 nzi, const P3m3Packed& k, p3m_f32x2& s, p3m_f32x2& dx,
                                           p3m_f32x2& dy, p3m_f32x2& dz, p3m_f32x2& r2) {
  dx = p3m_add2(p3m_pack2(tx[v0], tx[v1]), nxi);
  dy = p3m_add2(p3m_pack2(ty[v0], ty[v1]), nyi);
  dz = p3m_add2(p3m_pack2(tz[v0], tz[v1]), nzi);
  r2 = p3m_mul2(dx, dx);
  r2 = p3m_fma2(dy, dy, r2);
  r2 = p3m_fma2(dz, dz, r2);
  const p3m_f32x2 rinv = p3m_rsqrt2(p3m_add2(r2, k.eps2));
  const p3m_f32x2 w = p3m_mul2(r2, k.ws);                          // u²
  const p3m_f32x2 u = p3m_mul2(w, p3m_rsqrt2(p3m_add2(w, k.tiny)));
  const p3m_f32x2 tt = p3m_fma2(w, k.ts, k.m1);
  p3m_f32x2 q = k.nq[P3M_DEG];
#pragma unroll
  for (int d = P3M_DEG - 1; d >= 0; --d) q = p3m_fma2(q, tt, k.nq[d]);
  const p3m_f32x2 f = p3m_fma2(p3m_mul2(w, u), q, k.one);
  s = p3m_mul2(p3m_mul2(p3m_mul2(rinv, rinv), p3m_mul2(rinv, k.wj)), f);
}

// Rare path of the near-pair drain: bit b of nm = queue entry e_base + b of this lane was left out of the float32
// sum.  One entry per lane and trip goes to the list (warp-aggregated append); the self pair (zero term) is
// dropped; with a full list the pair is evaluated in float32 after all, like the plain kernel does.
template <typename SLOT>
__device__ __noinline__ void p3m_near_flush(const P3mNearList& nl, int i_slot, unsigned nm, int e_base,
                                            const unsigned short* __restrict__ qcol, const float* __restrict__ tx,
                                            const float* __restrict__ ty, const float* __restrict__ tz,
                                            p3m_f32x2 nxi, p3m_f32x2 nyi, p3m_f32x2 nzi, const P3m3Packed& k,
                                            double& ax, double& ay, double& az, const SLOT& slot_of) {
  const unsigned lane = threadIdx.x & 31u, lt = (1u << lane) - 1u;
  while (__any_sync(0xffffffffu, nm != 0u)) {
    const bool have = nm != 0u;
    const int b = have ? __ffs((int)nm) - 1 : 0;
    nm &= nm - 1u;                                   // clears the lowest set bit; 0 stays 0
    const int v = qcol[(e_base + b) * 32];
    const int j = slot_of(v);
    const bool want = have && j != i_slot;
    const unsigned m = __ballot_sync(0xffffffffu, want);
    int base = 0;
    if (lane == 0 && m) base = atomicAdd(nl.count, __popc(m));
    base = __shfl_sync(0xffffffffu, base, 0);
    if (want) {
      const int t = base + __popc(m & lt);
      if (t < nl.cap) {
        nl.list[t] = make_int2(i_slot, j);
      } else {
        p3m_f32x2 s2, dx, dy, dz, r2;
        p3m4_term2(tx, ty, tz, v, v, nxi, nyi, nzi, k, s2, dx, dy, dz, r2);
        float s0, s1, d0, d1;
        p3m_unpack2(s2, s0, s1);
        p3m_unpack2(dx, d0, d1); ax += (double)(s0 * d0);
        p3m_unpack2(dy, d0, d1); ay += (double)(s0 * d0);
        p3m_unpack2(dz, d0, d1); az += (double)(s0 * d0);
      }
    }
  }
}

// one packed evaluation of two queued candidates (tile slots v0, v1; ok = entry is live)
template <typename SLOT = P3mNoNear>
__device__ __forceinline__ void p3m4_eval2(const float* __restrict__ tx, const float* __restrict__ ty,
                                           const float* __restrict__ tz, int v0, int v1, bool ok0, bool ok1,
                                           p3m_f32x2 nxi, p3m_f32x2 nyi, p3m_f32x2 nzi, const P3m3Packed& k,
                                           p3m_f32x2& FX, p3m_f32x2& FY, p3m_f32x2& FZ,
                                           const P3mNearList& nl = P3mNearList{}, int i_slot = 0,
                                           const SLOT& slot_of = SLOT{}) {
  const p3m_f32x2 dx = p3m_add2(p3m_pack2(tx[v0], tx[v1]), nxi);
  const p3m_f32x2 dy = p3m_add2(p3m_pack2(ty[v0], ty[v1]), nyi);
  const p3m_f32x2 dz = p3m_add2(p3m_pack2(tz[v0], tz[v1]), nzi);
  p3m_f32x2 r2 = p3m_mul2(dx, dx);
  r2 = p3m_fma2(dy, dy, r2);
  r2 = p3m_fma2(dz, dz, r2);
  const p3m_f32x2 rinv = p3m_rsqrt2(p3m_add2(r2, k.eps2));
  const p3m_f32x2 w = p3m_mul2(r2, k.ws);                          // u²
  const p3m_f32x2 u = p3m_mul2(w, p3m_rsqrt2(p3m_add2(w, k.tiny)));
  const p3m_f32x2 tt = p3m_fma2(w, k.ts, k.m1);
  p3m_f32x2 q = k.nq[P3M_DEG];   // Horner (an Estrin form measured no faster on B200: 92.4 vs 90.5 ms)
#pragma unroll
  for (int d = P3M_DEG - 1; d >= 0; --d) q = p3m_fma2(q, tt, k.nq[d]);
  const p3m_f32x2 f = p3m_fma2(p3m_mul2(w, u), q, k.one);          // erfc(u) + 2u/sqrt(pi) exp(-u²)
  const p3m_f32x2 s = p3m_mul2(p3m_mul2(p3m_mul2(rinv, rinv), p3m_mul2(rinv, k.wj)), f);
  float s0, s1;
  p3m_unpack2(s, s0, s1);
  const p3m_f32x2 sm = p3m_pack2(ok0 ? s0 : 0.f, ok1 ? s1 : 0.f);
  FX = p3m_fma2(sm, dx, FX);
  FY = p3m_fma2(sm, dy, FY);
  FZ = p3m_fma2(sm, dz, FZ);
}

// phase 2: evaluate the queued candidates of every lane, four per round as two independent
// packed chains (the queue has 4 rows of slack beyond QD, so reading ahead is always in range)
template <typename SLOT = P3mNoNear>
__device__ __forceinline__ void p3m4_drain(const float* __restrict__ tx, const float* __restrict__ ty,
                                           const float* __restrict__ tz, const unsigned short* __restrict__ qcol,
                                           int cnt, p3m_f32x2 nxi, p3m_f32x2 nyi, p3m_f32x2 nzi,
                                           const P3m3Packed& k, double& ax, double& ay, double& az,
                                           const P3mNearList& nl = P3mNearList{}, int i_slot = 0,
                                           const SLOT& slot_of = SLOT{}) {
  const int rounds = (__reduce_max_sync(0xffffffffu, cnt) + 3) >> 2;
  p3m_f32x2 FX = 0ull, FY = 0ull, FZ = 0ull, GX = 0ull, GY = 0ull, GZ = 0ull;
  if constexpr (!std::is_same<SLOT, P3mNoNear>::value) {
    // Near-pair list: four rounds (16 entries per lane) per trip, branch-free like the plain drain so that ptxas
    // schedules the eight packed chains together; an entry that stays out of the float32 sum (closer than r_near, in
    // the cut-off shell, or the self pair) sets a bit of nm, and ONE vote per trip sends the warp to the rare path.
    // Non-negative floats order like their bit patterns: rn2 <= r^2 < rb2 is one unsigned compare of r^2 - rn2.
    const unsigned lo = __float_as_uint(nl.rn2), span = __float_as_uint(nl.rb2) - lo;
    auto round = [&](int rr, int q, unsigned& nm) {
      const int e = 4 * rr;
      const int v0 = qcol[e * 32], v1 = qcol[(e + 1) * 32], v2 = qcol[(e + 2) * 32], v3 = qcol[(e + 3) * 32];
      p3m_f32x2 sa, sb, dxa, dya, dza, dxb, dyb, dzb, r2a, r2b;
      p3m4_term2(tx, ty, tz, v0, v1, nxi, nyi, nzi, k, sa, dxa, dya, dza, r2a);
      p3m4_term2(tx, ty, tz, v2, v3, nxi, nyi, nzi, k, sb, dxb, dyb, dzb, r2b);
      float q0, q1, q2, q3, s0, s1, s2, s3;
      p3m_unpack2(r2a, q0, q1); p3m_unpack2(r2b, q2, q3);
      p3m_unpack2(sa, s0, s1); p3m_unpack2(sb, s2, s3);
      const bool ok0 = e < cnt, ok1 = e + 1 < cnt, ok2 = e + 2 < cnt, ok3 = e + 3 < cnt;
      const bool f0 = ok0 && (__float_as_uint(q0) - lo) < span, f1 = ok1 && (__float_as_uint(q1) - lo) < span;
      const bool f2 = ok2 && (__float_as_uint(q2) - lo) < span, f3 = ok3 && (__float_as_uint(q3) - lo) < span;
      if (ok0 && !f0) nm |= 1u << (4 * q);
      if (ok1 && !f1) nm |= 2u << (4 * q);
      if (ok2 && !f2) nm |= 4u << (4 * q);
      if (ok3 && !f3) nm |= 8u << (4 * q);
      const p3m_f32x2 ma = p3m_pack2(f0 ? s0 : 0.f, f1 ? s1 : 0.f);
      const p3m_f32x2 mb = p3m_pack2(f2 ? s2 : 0.f, f3 ? s3 : 0.f);
      FX = p3m_fma2(ma, dxa, FX); FY = p3m_fma2(ma, dya, FY); FZ = p3m_fma2(ma, dza, FZ);
      GX = p3m_fma2(mb, dxb, GX); GY = p3m_fma2(mb, dyb, GY); GZ = p3m_fma2(mb, dzb, GZ);
    };
    int r = 0;
    for (; r + 4 <= rounds; r += 4) {
      unsigned nm = 0u;
#pragma unroll
      for (int q = 0; q < 4; ++q) round(r + q, q, nm);
      if (__any_sync(0xffffffffu, nm != 0u))
        p3m_near_flush(nl, i_slot, nm, 4 * r, qcol, tx, ty, tz, nxi, nyi, nzi, k, ax, ay, az, slot_of);
    }
    for (; r < rounds; ++r) {
      unsigned nm = 0u;
      round(r, 0, nm);
      if (__any_sync(0xffffffffu, nm != 0u))
        p3m_near_flush(nl, i_slot, nm, 4 * r, qcol, tx, ty, tz, nxi, nyi, nzi, k, ax, ay, az, slot_of);
    }
  } else
  for (int r = 0; r < rounds; ++r) {
    const int e = 4 * r;
    const int v0 = qcol[e * 32], v1 = qcol[(e + 1) * 32], v2 = qcol[(e + 2) * 32], v3 = qcol[(e + 3) * 32];
    p3m4_eval2<SLOT>(tx, ty, tz, v0, v1, e < cnt, e + 1 < cnt, nxi, nyi, nzi, k, FX, FY, FZ, nl, i_slot, slot_of);
    p3m4_eval2<SLOT>(tx, ty, tz, v2, v3, e + 2 < cnt, e + 3 < cnt, nxi, nyi, nzi, k, GX, GY, GZ, nl, i_slot, slot_of);
  }
  float a, b, c, d;
  p3m_unpack2(FX, a, b); p3m_unpack2(GX, c, d); ax += (double)((a + b) + (c + d));
  p3m_unpack2(FY, a, b); p3m_unpack2(GY, c, d); ay += (double)((a + b) + (c + d));
  p3m_unpack2(FZ, a, b); p3m_unpack2(GZ, c, d); az += (double)((a + b) + (c + d));
}

// phase 1 over the resident tile slots [0, npad), npad a multiple of 32; drains when a queue
// runs full and at the end.  One group = 32 slots: every lane takes slots 4s..4s+3 of both
// 16-slot rows (s = its j-stream) as one LDS.128 per coordinate, i.e. 8 candidates as four
// independent packed r² chains; the pushes go through a running pointer into the lane's column.
template <int QD, typename SLOT = P3mNoNear>
__device__ __forceinline__ void p3m4_pass(const float* __restrict__ tx, const float* __restrict__ ty,
                                          const float* __restrict__ tz, unsigned short* __restrict__ qcol,
                                          int npad, int strm, p3m_f32x2 nxi, p3m_f32x2 nyi, p3m_f32x2 nzi,
                                          const P3m3Packed& k, double& ax, double& ay, double& az,
                                          const P3mNearList& nl = P3mNearList{}, int i_slot = 0,
                                          const SLOT& slot_of = SLOT{}) {
  // 32-bit shared-memory address of the next free entry of this lane's column (64 B per row)
  const unsigned qbase = (unsigned)__cvta_generic_to_shared(qcol);
  unsigned qa = qbase;
  const unsigned qfull = qbase + (QD - 8) * 64;   // a group pushes at most 8 entries per lane
  for (int vb = 4 * strm; vb < npad; vb += 32) {
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
      p3m4_drain<SLOT>(tx, ty, tz, qcol, (int)((qa - qbase) >> 6), nxi, nyi, nzi, k, ax, ay, az, nl, i_slot, slot_of);
      qa = qbase;
    }
  }
  if (__any_sync(0xffffffffu, qa != qbase))
    p3m4_drain<SLOT>(tx, ty, tz, qcol, (int)((qa - qbase) >> 6), nxi, nyi, nzi, k, ax, ay, az, nl, i_slot, slot_of);
}

template <bool QUEUE, int QD, bool NEAR = false>
__global__ void __launch_bounds__(P3M4_WARPS * 32, QD >= 64 ? 4 : 5)
p3m_short_v4_kernel(const float4* __restrict__ sorted, const int* __restrict__ cell_start,
                    const int* __restrict__ perm, int ncs, int64_t cell_lo, int64_t cell_hi, P3mParams prm,
                    double scale, double* __restrict__ acc, int accumulate, int i_limit, int cap,
                    const int* __restrict__ cell_list, const int* __restrict__ list_count,
                    P3mNearList nl = P3mNearList{}) {
  extern __shared__ __align__(16) unsigned char p3m4_smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int li = lane & 7, strm = lane >> 3;
  // CTA-wide: translation of neighbour n = (ox+1)*9 + (oy+1)*3 + (oz+1) into the i-cell frame
  float* offtab = reinterpret_cast<float*>(p3m4_smem);                       // 27*3 floats (96 reserved)
  unsigned char* wbase = p3m4_smem + 96 * sizeof(float) + (size_t)warp * p3m4_warp_bytes(cap, QD);
  float* tx = reinterpret_cast<float*>(wbase);
  float* ty = tx + cap;
  float* tz = ty + cap;
  int* roff = reinterpret_cast<int*>(tz + cap);                               // 32: exclusive prefix
  int* rjs = roff + 32;                                                       // 32: first slot of the cell
  unsigned short* qcol = reinterpret_cast<unsigned short*>(rjs + 32) + lane;  // this lane's queue column
  if (threadIdx.x < 27) {
    const int n = threadIdx.x;
    offtab[3 * n] = (float)(n / 9 - 1) * prm.cs;
    offtab[3 * n + 1] = (float)((n / 3) % 3 - 1) * prm.cs;
    offtab[3 * n + 2] = (float)(n % 3 - 1) * prm.cs;
  }
  for (int r = 0; r < QD + 4; ++r) qcol[r * 32] = 0;   // stale reads must be valid tile indices
  __syncthreads();
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
  // work items: the cells [cell_lo, cell_hi), or (v6's overflow pass) the cells of a device-side list
  const int64_t wstride = (int64_t)gridDim.x * P3M4_WARPS;
  const int64_t nwork = cell_list ? (int64_t)__ldg(list_count) : cell_hi - cell_lo;
  for (int64_t wi = (int64_t)blockIdx.x * P3M4_WARPS + warp; wi < nwork; wi += wstride) {
    const int64_t c = cell_list ? (int64_t)cell_list[wi] : cell_lo + wi;
    const int is = cell_start[c], ie = cell_start[c + 1];
    if (is == ie) continue;
    const int cz = (int)(c % ncs);
    const int cy = (int)((c / ncs) % ncs);
    const int cx = (int)(c / ((int64_t)ncs * ncs));
    // lanes 0..26: population of the 27 neighbour cells, then their exclusive prefix
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
    __syncwarp();   // the previous cell's fills are done with the run table
    roff[lane] = lane < 27 ? incl - my_len : 0x7fffffff;
    rjs[lane] = my_js;
    __syncwarp();
    const bool single = ncand <= cap;   // whole neighbourhood fits: fill once per cell
    bool resident = false;
    for (int ib = is; ib < ie; ib += 8) {
      const int i = ib + li;
      // particles with perm >= i_limit are halo copies of another slab's particles: sources only
      const int oi = i < ie ? perm[i] : 0x7fffffff;
      const bool ivalid = oi < i_limit;
      if (!__any_sync(0xffffffffu, ivalid)) continue;
      float4 pi = sorted[i < ie ? i : is];
      if (!ivalid) pi.x = 3e18f;   // idle lanes never pass the range test (no queue traffic)
      const p3m_f32x2 nxi = p3m_pack2(-pi.x, -pi.x), nyi = p3m_pack2(-pi.y, -pi.y), nzi = p3m_pack2(-pi.z, -pi.z);
      double ax = 0.0, ay = 0.0, az = 0.0;
      for (int first = 0; first < ncand; first += cap) {
        const int nt = min(cap, ncand - first);
        if (!resident) {
          __syncwarp();   // every lane is done reading the previous chunk
          p3m4_fill(sorted, roff, rjs, offtab, tx, ty, tz, first, nt, lane);
        }
        if constexpr (QUEUE && NEAR) {
          // tile slot v of this chunk is candidate first + v of the flattened 27-cell list
          auto slot_of = [&](int v) {
            const int g = first + v;
            int m = 0;
#pragma unroll
            for (int sb = 16; sb > 0; sb >>= 1) m = (roff[m + sb] <= g) ? m + sb : m;
            return rjs[m] + (g - roff[m]);
          };
          p3m4_pass<QD, decltype(slot_of)>(tx, ty, tz, qcol, (nt + 31) & ~31, strm, nxi, nyi, nzi, k, ax, ay, az, nl, i,
                                           slot_of);
        } else if constexpr (QUEUE)
          p3m4_pass<QD>(tx, ty, tz, qcol, (nt + 31) & ~31, strm, nxi, nyi, nzi, k, ax, ay, az);
#ifdef NB_AB_KERNELS
        else   // v3's evaluate-all loop on the v4 tile (A/B: NB_P3M_KERNEL=5)
          p3m3_process(tx, ty, tz, (nt + 31) & ~31, strm, nxi, nyi, nzi, k, ax, ay, az);
#endif
      }
      resident = single;
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
  }
}

__device__ __forceinline__ P3m3Packed p3m_make_packed(const P3mParams& prm, float wj) {
  P3m3Packed k;
  const float e2 = fmaxf(prm.eps2, 1e-20f);
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
  return k;
}

__device__ __forceinline__ void p3m5_store_acc(double* __restrict__ acc, int64_t o, double ax, double ay, double az,
                                               double scale, int accumulate) {
  if (accumulate) {
    acc[3 * o] += ax * scale; acc[3 * o + 1] += ay * scale; acc[3 * o + 2] += az * scale;
  } else {
    acc[3 * o] = ax * scale; acc[3 * o + 1] = ay * scale; acc[3 * o + 2] = az * scale;
  }
}

// v6's form of p3m4_pass: tile slots [vbeg, vend) (multiples of 32) of the CTA tile, ABSOLUTE slot
// indices in the queue — a stale queue entry then always names a slot that some earlier fill of
// this kernel wrote (finite coordinates); window-relative indices could point at never-written
// shared memory, and 0 * NaN in the masked accumulate is NaN.
// The loads go through explicit 32-bit shared addresses (one running address, ty/tz at +cap4 /
// +2*cap4 bytes): with generic pointers ptxas re-derived the three tile bases and the queue limit
// from S2R/S2UR in every iteration (84 instructions per group, 17 of them address arithmetic).
__device__ __forceinline__ float4 p3m_lds128(unsigned saddr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(saddr));
  return v;
}

template <int QD, int CAPC, typename SLOT = P3mNoNear>
__device__ __forceinline__ void p3m6_pass(const float* __restrict__ tx, const float* __restrict__ ty,
                                          const float* __restrict__ tz, unsigned txs, unsigned cap4_rt,
                                          unsigned short* __restrict__ qcol, unsigned qbase, int vbeg, int vend,
                                          int strm, p3m_f32x2 nxi, p3m_f32x2 nyi, p3m_f32x2 nzi,
                                          const P3m3Packed& k, double& ax, double& ay, double& az,
                                          const P3mNearList& nl = P3mNearList{}, int i_slot = 0,
                                          const SLOT& slot_of = SLOT{}) {
  // CAPC > 0: the tile capacity is a compile-time constant, so the y / z planes are immediate offsets of
  // the running x address (the runtime form cost an LDCU + MOV + 2 LEA per group)
  const unsigned cap4 = CAPC > 0 ? 4u * (unsigned)CAPC : cap4_rt;
  unsigned qa = qbase;
  unsigned qfull = qbase + (QD - 8) * 64;   // a group pushes at most 8 entries per lane
  // opaque to the optimiser: ptxas otherwise re-derives this limit from %tid, the CTA's shared window and the
  // tile size in EVERY group (S2R, SHF, LOP3, 2 IMAD, 2 IADD3, 2 MOV: 9 of the 92 instructions of the loop)
  asm volatile("" : "+r"(qfull));
  unsigned sx = txs + 4u * (unsigned)(vbeg + 4 * strm);
  const unsigned send = txs + 4u * (unsigned)vend;
  int v0 = vbeg + 4 * strm;
  for (; sx < send; sx += 128u, v0 += 32) {
    const unsigned sy = sx + cap4, sz = sy + cap4;
    p3m_f32x2 r2[4];
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const float4 X = p3m_lds128(sx + 64u * j);
      const float4 Y = p3m_lds128(sy + 64u * j);
      const float4 Z = p3m_lds128(sz + 64u * j);
      const p3m_f32x2 dx0 = p3m_add2(p3m_pack2(X.x, X.y), nxi), dx1 = p3m_add2(p3m_pack2(X.z, X.w), nxi);
      const p3m_f32x2 dy0 = p3m_add2(p3m_pack2(Y.x, Y.y), nyi), dy1 = p3m_add2(p3m_pack2(Y.z, Y.w), nyi);
      const p3m_f32x2 dz0 = p3m_add2(p3m_pack2(Z.x, Z.y), nzi), dz1 = p3m_add2(p3m_pack2(Z.z, Z.w), nzi);
      r2[2 * j] = p3m_fma2(dz0, dz0, p3m_fma2(dy0, dy0, p3m_mul2(dx0, dx0)));
      r2[2 * j + 1] = p3m_fma2(dz1, dz1, p3m_fma2(dy1, dy1, p3m_mul2(dx1, dx1)));
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int v = v0 + 16 * (j >> 1) + 2 * (j & 1);
      float r2a, r2b;
      p3m_unpack2(r2[j], r2a, r2b);
      if (r2a < k.rc2) { p3m_sts_u16(qa, (unsigned short)v); qa += 64; }
      if (r2b < k.rc2) { p3m_sts_u16(qa, (unsigned short)(v + 1)); qa += 64; }
    }
    if (__any_sync(0xffffffffu, qa > qfull)) {
      p3m4_drain<SLOT>(tx, ty, tz, qcol, (int)((qa - qbase) >> 6), nxi, nyi, nzi, k, ax, ay, az, nl, i_slot, slot_of);
      qa = qbase;
    }
  }
  if (__any_sync(0xffffffffu, qa != qbase))
    p3m4_drain<SLOT>(tx, ty, tz, qcol, (int)((qa - qbase) >> 6), nxi, nyi, nzi, k, ax, ay, az, nl, i_slot, slot_of);
}

#ifdef NB_AB_KERNELS
#include "p3m_ab.cuh"   // superseded kernel generations v1, v2, v3, v5 and the byte-mask pass (A/B builds only)
#endif

// ---------------------------------------------------------------------------------------
// v6 — v4's pair loop with ONE tile per CTA shared by four z-adjacent cells.
//
// ncu on v4 at the C5 size (profiles/r1e_ncu_p3m_short_v4_v5_c5.txt): the per-warp tile fill is
// 14 % of the instructions but 28 % of the stall samples, and the 13.8 KB of shared memory per
// warp hold the kernel at 16 warps/SM.  The four warps of a v6 CTA take the cells z0..z0+3 of one
// (cx,cy) column; the union of their neighbourhoods is 6 z-slabs of 9 cells (54 cells instead of
// 4 x 27), loaded ONCE by all 128 threads into a tile ordered slab by slab.  The slabs w..w+2 that
// warp w needs are then one contiguous slot range; rounding it out to whole 32-slot groups only
// adds candidates of the slabs w-1 / w+3, which are a full cell (>= r_cut) away and fail the
// range test like any other candidate.  Shared memory per warp drops to ~10 KB (5 CTAs/SM) and the
// fill work per cell from 3 slabs to 1.5.  Tile coordinates are in the frame of cell z0+1
// (|z| < 4 cs); the i-particle reads its own z back from its tile slot, so the self pair (and
// coincident points) stay at exactly zero separation.  A run whose 54 cells do not fit the tile
// is appended to a device-side list and handled afterwards by the v4 kernel (no host sync).
#define P3M6_WARPS 4

__host__ __device__ inline size_t p3m6_cta_bytes(int cap, int qd) {
  return sizeof(float) * (3 * (size_t)cap + 192) + sizeof(int) * 160 +
         P3M6_WARPS * sizeof(unsigned short) * (size_t)(qd + 4) * 32;
}

template <int QD, int CTAS, bool MASK, int CAPC = 0, bool NEAR = false>
__global__ void __launch_bounds__(P3M6_WARPS * 32, CTAS)
p3m_short_v6_kernel(const float4* __restrict__ sorted, const int* __restrict__ cell_start,
                    const int* __restrict__ perm, int ncs, int64_t cell_lo, int64_t cell_hi, P3mParams prm,
                    double scale, double* __restrict__ acc, int accumulate, int i_limit, int cap,
                    int* __restrict__ overflow_list, int* __restrict__ overflow_count,
                    int* __restrict__ unit_counter = nullptr, int sm_limit = 0,
                    const int* __restrict__ stop_flag = nullptr, P3mNearList nl = P3mNearList{}) {
  // SM partition for the overlapped P3M step (nb_p3m_set_sm_limit): CTAs that land on an SM >= sm_limit leave
  // at once, the others pull their work units from an atomic counter — the pair kernel then occupies the first
  // sm_limit SMs completely (6 CTAs fill their shared memory) and the memory-bound mesh chain of the same step,
  // launched on another stream, gets the remaining SMs to itself.  When the mesh chain is done it raises
  // *stop_flag (nb_p3m_mesh_done): the confined CTAs stop fetching and a second, unconfined launch on the same
  // counter finishes the units on the whole GPU.
  if (sm_limit > 0) {
    unsigned smid;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    if ((int)smid >= sm_limit) return;
  }
  extern __shared__ __align__(16) unsigned char p3m6_smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int li = lane & 7, strm = lane >> 3;
  if (CAPC > 0) cap = CAPC;        // compile-time tile capacity: the y / z planes become immediate offsets
  float* tx = reinterpret_cast<float*>(p3m6_smem);
  float* ty = tx + cap;
  float* tz = ty + cap;
  float* offx = tz + cap;          // 64: translation of run m = 9*slab + 3*(ox+1) + (oy+1)
  float* offy = offx + 64;
  float* offz = offy + 64;
  int* roff = reinterpret_cast<int*>(offz + 64);   // 64 (+1): exclusive prefix of the run lengths
  int* rjs = roff + 80;                            // 64: first sorted slot of each run
  int* s_tot = rjs + 64;                           // [0] total candidates
  unsigned short* qcol = reinterpret_cast<unsigned short*>(s_tot + 16) + (size_t)warp * (QD + 4) * 32 + lane;
  if (threadIdx.x < 64) {
    const int m = threadIdx.x, sl = m / 9, r = m % 9;
    offx[m] = (float)(r / 3 - 1) * prm.cs;
    offy[m] = (float)(r % 3 - 1) * prm.cs;
    offz[m] = (float)(sl - 2) * prm.cs;            // frame of cell z0+1: slab sl holds z0-1+sl
  }
  for (int r = 0; r < QD + 4; ++r) qcol[r * 32] = 0;   // stale reads must be valid tile indices
  __syncthreads();
  const P3m3Packed k = p3m_make_packed(prm, __ldg(&sorted[0].w));
  const unsigned txs = (unsigned)__cvta_generic_to_shared(tx), cap4 = 4u * (unsigned)cap;
  const unsigned qbase = (unsigned)__cvta_generic_to_shared(qcol);

  const int nrun = (ncs + P3M6_WARPS - 1) / P3M6_WARPS;
  const int64_t col_lo = cell_lo / ncs, col_hi = (cell_hi + ncs - 1) / ncs;
  const int64_t nunits = (col_hi - col_lo) * nrun;
  for (int64_t u = blockIdx.x;; u += gridDim.x) {
    if (unit_counter) {      // dynamic: one unit per fetch (s_tot[2] is not used by the tile code)
      __syncthreads();
      if (threadIdx.x == 0)
        s_tot[2] = (stop_flag && *(const volatile int*)stop_flag) ? 0x7fffffff : atomicAdd(unit_counter, 1);
      __syncthreads();
      u = s_tot[2];
    }
    if (u >= nunits) break;
    const int64_t col = col_lo + u / nrun;
    const int z0 = (int)(u % nrun) * P3M6_WARPS;
    const int cx = (int)(col / ncs), cy = (int)(col % ncs);
    const int nact = min(P3M6_WARPS, ncs - z0);     // cells of this run
    __syncthreads();   // every warp is done with the previous tile and run table
    // run table: 9 cells of each of the nact+2 slabs
    int my_len = 0;
    if (threadIdx.x < 64) {
      const int m = threadIdx.x, sl = m / 9, r = m % 9;
      int js = 0;
      if (sl < nact + 2) {
        int nx = cx + r / 3 - 1, ny = cy + r % 3 - 1, nz = z0 - 1 + sl;
        nx = nx < 0 ? nx + ncs : (nx >= ncs ? nx - ncs : nx);
        ny = ny < 0 ? ny + ncs : (ny >= ncs ? ny - ncs : ny);
        nz = nz < 0 ? nz + ncs : (nz >= ncs ? nz - ncs : nz);
        const int64_t n = ((int64_t)nx * ncs + ny) * ncs + nz;
        js = cell_start[n];
        my_len = cell_start[n + 1] - js;
      }
      rjs[m] = js;
      // exclusive prefix over the 64 entries: warp scan + the first warp's total
      int incl = my_len;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int up = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += up;
      }
      if (m == 31) s_tot[1] = incl;
      roff[m] = incl - my_len;    // warp-local for now
    }
    __syncthreads();
    if (threadIdx.x >= 32 && threadIdx.x < 64) roff[threadIdx.x] += s_tot[1];
    if (threadIdx.x == 63) { const int t = roff[63] + my_len; s_tot[0] = t; roff[64] = t; }
    __syncthreads();
    const int total = s_tot[0];
    if (total > cap) {   // does not fit: leave the cells of this run to the v4 pass
      if (threadIdx.x < nact) {
        const int64_t c = col * ncs + z0 + threadIdx.x;
        if (c >= cell_lo && c < cell_hi && cell_start[c + 1] > cell_start[c])
          overflow_list[atomicAdd(overflow_count, 1)] = (int)c;
      }
      continue;
    }
    // cooperative fill (4 independent loads per thread and round); runs of length 0 are skipped by
    // the search because the next run starts at the same offset
    for (int t0 = threadIdx.x; t0 < total; t0 += 4 * P3M6_WARPS * 32) {
      int mm[4];
      float4 p[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int g = min(t0 + j * P3M6_WARPS * 32, total - 1);
        int m = 0;   // largest m with roff[m] <= g; entries past the last slab equal `total`
#pragma unroll
        for (int sstep = 32; sstep > 0; sstep >>= 1) m = (roff[m + sstep] <= g) ? m + sstep : m;
        mm[j] = m;
        p[j] = sorted[rjs[m] + (g - roff[m])];
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int t = t0 + j * P3M6_WARPS * 32;
        if (t < total) {
          tx[t] = p[j].x + offx[mm[j]];
          ty[t] = p[j].y + offy[mm[j]];
          tz[t] = p[j].z + offz[mm[j]];
        }
      }
    }
    {
      const int npad = (total + 31) & ~31;
      if ((int)threadIdx.x < npad - total) {
        tx[total + threadIdx.x] = 1e18f;
        ty[total + threadIdx.x] = 0.f;
        tz[total + threadIdx.x] = 0.f;
      }
    }
    __syncthreads();
    // The i-blocks (8 particles) of the run's cells form one work list dealt round-robin to the
    // four warps: a warp per cell would leave the CTA waiting for its most populated cell at the
    // next barrier.  Cell w is run 4 of slab w+1, so its first slot and size come from the table.
    int e[P3M6_WARPS + 1];
    e[0] = 0;
#pragma unroll
    for (int w = 0; w < P3M6_WARPS; ++w) {
      const int64_t cw = col * ncs + z0 + w;
      const int m = 9 * (w + 1) + 4;
      const int cnt = (w < nact && cw >= cell_lo && cw < cell_hi) ? roff[m + 1] - roff[m] : 0;
      e[w + 1] = e[w] + ((cnt + 7) >> 3);
    }
    for (int blk = warp; blk < e[P3M6_WARPS]; blk += P3M6_WARPS) {
      const int w = (blk >= e[1]) + (blk >= e[2]) + (blk >= e[3]);
      const int m = 9 * (w + 1) + 4;
      const int is = rjs[m], ie = is + (roff[m + 1] - roff[m]);
      const int ib = is + 8 * (blk - (w == 0 ? e[0] : (w == 1 ? e[1] : (w == 2 ? e[2] : e[3]))));
      // this cell's candidates: slabs w .. w+2, rounded out to whole groups
      const int a_w = roff[9 * w] & ~31;
      const int b_w = (roff[9 * (w + 3)] + 31) & ~31;
      const float* own_z = tz + roff[m];
      const int i = ib + li;
      // particles with perm >= i_limit are halo copies of another slab's particles: sources only
      const int oi = i < ie ? perm[i] : 0x7fffffff;
      const bool ivalid = oi < i_limit;
      if (!__any_sync(0xffffffffu, ivalid)) continue;
      float4 pi = sorted[i < ie ? i : is];
      pi.z = own_z[min(i, ie - 1) - is];
      if (!ivalid) pi.x = 3e18f;   // idle lanes never pass the range test (no queue traffic)
      const p3m_f32x2 nxi = p3m_pack2(-pi.x, -pi.x), nyi = p3m_pack2(-pi.y, -pi.y), nzi = p3m_pack2(-pi.z, -pi.z);
      double ax = 0.0, ay = 0.0, az = 0.0;
#ifdef NB_AB_KERNELS
      if constexpr (MASK)
        p3m7_pass(tx, ty, tz, txs, cap4, reinterpret_cast<unsigned char*>(qcol - lane) + lane, a_w, b_w, strm, nxi,
                  nyi, nzi, k, ax, ay, az);
      else
#endif
      if constexpr (NEAR) {
        // tile slot -> run (largest m with roff[m] <= v; entries past the last slab equal `total`) -> sorted slot
        auto slot_of = [&](int v) {
          int m = 0;
#pragma unroll
          for (int sb = 32; sb > 0; sb >>= 1) m = (roff[m + sb] <= v) ? m + sb : m;
          return rjs[m] + (v - roff[m]);
        };
        p3m6_pass<QD, CAPC, decltype(slot_of)>(tx, ty, tz, txs, cap4, qcol, qbase, a_w, b_w, strm, nxi, nyi, nzi, k, ax,
                                               ay, az, nl, i, slot_of);
      } else
        p3m6_pass<QD, CAPC>(tx, ty, tz, txs, cap4, qcol, qbase, a_w, b_w, strm, nxi, nyi, nzi, k, ax, ay, az);
      ax += __shfl_xor_sync(0xffffffffu, ax, 8);  ay += __shfl_xor_sync(0xffffffffu, ay, 8);
      az += __shfl_xor_sync(0xffffffffu, az, 8);
      ax += __shfl_xor_sync(0xffffffffu, ax, 16); ay += __shfl_xor_sync(0xffffffffu, ay, 16);
      az += __shfl_xor_sync(0xffffffffu, az, 16);
      if (ivalid && strm == 0) p3m5_store_acc(acc, oi, ax, ay, az, scale, accumulate);
    }
  }
}

extern "C" int nb_p3m_mesh_done(nb_ctx* ctx, void* stream) {
  NB_REQUIRE(ctx, "bad arguments");
  NB_CHECK_CUDA(cudaSetDevice(ctx->device));
  void* fl = nullptr;
  int rc = nb_scratch(ctx, NB_SCR_FLAGS, 64, &fl);
  if (rc) return rc;
  NB_CHECK_CUDA(cudaMemsetAsync((int*)fl + 1, 1, sizeof(int), (cudaStream_t)stream));
  return NB_OK;
}

extern "C" int nb_p3m_set_sm_limit(nb_ctx* ctx, int sm_limit) {
  NB_REQUIRE(ctx && sm_limit >= 0, "bad arguments");
  ctx->p3m_sm_limit = sm_limit >= ctx->num_sms ? 0 : sm_limit;
  return NB_OK;
}

extern "C" int nb_p3m_set_ctas_per_sm(nb_ctx* ctx, int ctas_per_sm) {
  NB_REQUIRE(ctx && ctas_per_sm >= 0 && ctas_per_sm <= 64, "ctas_per_sm must be in [0, 64]");
  ctx->p3m_ctas_per_sm = ctas_per_sm;
  return NB_OK;
}

// float64 evaluation of the listed near pairs from the original positions (see P3mNearList): for entry (i, j)
// acc[perm[i]] += scale * w_j * d / (r^2 + eps^2)^(3/2) * (erfc(u) + 2u/sqrt(pi) exp(-u^2)), d = x_j - x_i (nearest
// image), u = r / (2 r_s) — the pair term of the short-range sum without the float32 tile coordinates.
__global__ void __launch_bounds__(256)
p3m_near_kernel(const int2* __restrict__ list, const int* __restrict__ count, int cap,
                const float4* __restrict__ sorted, const int* __restrict__ perm, const double* __restrict__ pos,
                double box, double inv_box, double inv_2rs, double eps2, double rc2, double scale,
                double* __restrict__ acc) {
  const int n = min(*count, cap);
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (int64_t)gridDim.x * blockDim.x) {
    const int2 e = list[t];
    const int64_t i = perm[e.x], j = perm[e.y];
    double dx = pos[3 * j] - pos[3 * i], dy = pos[3 * j + 1] - pos[3 * i + 1], dz = pos[3 * j + 2] - pos[3 * i + 2];
    dx -= box * rint(dx * inv_box);
    dy -= box * rint(dy * inv_box);
    dz -= box * rint(dz * inv_box);
    const double r2 = dx * dx + dy * dy + dz * dz;
    if (!(r2 > 0.0) || r2 >= rc2) continue;      // the cut-off, decided in float64 for the pairs of the boundary shell
    const double u = sqrt(r2) * inv_2rs;
    const double f = erfc(u) + 1.1283791670955126 * u * exp(-u * u);
    const double rinv = 1.0 / sqrt(r2 + eps2);
    const double sfac = scale * (double)sorted[e.y].w * rinv * rinv * rinv * f;
    atomicAdd(&acc[3 * i], sfac * dx);
    atomicAdd(&acc[3 * i + 1], sfac * dy);
    atomicAdd(&acc[3 * i + 2], sfac * dz);
  }
}

// coef_host: P3M_DEG+1 doubles (monomial coefficients of Q in t, low to high), HOST pointer.
// pos != NULL and r_near > 0: near-pair list (nb_p3m_short_exact).
static int p3m_short_impl(nb_ctx* ctx, const void* sorted_xyzw, const int32_t* cell_start, const int32_t* perm,
                          int ncs, int64_t cell_lo, int64_t cell_hi, double cell_side, double r_cut,
                          double r_s, double eps, const double* coef_host, int ncoef, double u_max,
                          double scale, int accumulate, int64_t i_limit, double* acc, const double* pos, int64_t np,
                          double box, double r_near, int32_t* near_count_out, void* stream) {
  NB_REQUIRE(ctx && sorted_xyzw && cell_start && perm && coef_host && acc, "NULL argument");
  NB_REQUIRE(ncoef == P3M_DEG + 1, "split polynomial must have 11 coefficients");
  NB_REQUIRE(ncs >= 3, "need at least 3 cells per side (27 distinct neighbours)");
  NB_REQUIRE(cell_side >= r_cut && r_cut > 0 && r_s > 0, "cell side must be >= r_cut > 0");
  NB_REQUIRE(r_cut / (2 * r_s) <= u_max * (1 + 1e-9), "polynomial range does not cover r_cut");
  const int64_t nc = (int64_t)ncs * ncs * ncs;
  NB_REQUIRE(cell_lo >= 0 && cell_hi <= nc && cell_lo <= cell_hi, "bad cell range");
  cudaStream_t st = (cudaStream_t)stream;
  NB_CHECK_CUDA(cudaSetDevice(ctx->device));
  if (cell_lo == cell_hi) {
    if (near_count_out) NB_CHECK_CUDA(cudaMemsetAsync(near_count_out, 0, sizeof(int32_t), st));
    return NB_OK;
  }
  P3mParams prm;
  prm.rc2 = (float)(r_cut * r_cut);
  prm.eps2 = (float)(eps * eps);
  prm.w_scale = (float)(1.0 / (4.0 * r_s * r_s));
  prm.t_scale = (float)(2.0 / (u_max * u_max));
  prm.cs = (float)cell_side;
  for (int d = 0; d <= P3M_DEG; ++d) prm.coef[d] = (float)coef_host[d];
  const int ilim = (i_limit <= 0 || i_limit > 0x7fffffff) ? 0x7fffffff : (int)i_limit;
  const int64_t ncell = cell_hi - cell_lo;
  const int64_t cap = (int64_t)ctx->num_sms * (ctx->p3m_ctas_per_sm > 0 ? ctx->p3m_ctas_per_sm : 64);
#ifndef NB_AB_KERNELS
  // Product path: v6 (one tile per CTA shared by four z-adjacent cells: queue depth 56, 1664-slot tile, 6 CTAs
  // per SM at 80 registers) and, for the runs whose 54 cells do not fit the tile, the v4 kernel over the
  // device-side overflow list.  The superseded generations and their getenv switches are in p3m_ab.cuh
  // (make AB=1).
  {
    constexpr int QD6 = 56, CAP6 = 1664;
    const size_t smem6 = p3m6_cta_bytes(CAP6, QD6);
    void* ovf = nullptr;
    int rc = nb_scratch(ctx, NB_SCR_SORT3, sizeof(int) * (size_t)(ncell + 4), &ovf);
    if (rc) return rc;
    int* ovf_count = (int*)ovf;
    int* ovf_list = ovf_count + 4;
    NB_CHECK_CUDA(cudaMemsetAsync(ovf_count, 0, sizeof(int), st));
    // near-pair list: 2 entries per particle (a uniform set has 0.5 partners per particle inside 20 softening
    // lengths); pairs beyond the capacity stay with the float32 evaluation
    const bool near = pos != nullptr && r_near > 0.0;
    P3mNearList nl{};
    if (near) {
      const int64_t want = 2 * np > ((int64_t)1 << 20) ? 2 * np : ((int64_t)1 << 20);
      const int64_t ncap = want < 0x7ffffff0 ? want : 0x7ffffff0;
      void* nb = nullptr;
      rc = nb_scratch(ctx, NB_SCR_NEAR, sizeof(int2) * (size_t)ncap + 256, &nb);
      if (rc) return rc;
      nl.count = (int*)nb;
      nl.list = (int2*)((char*)nb + 256);
      nl.cap = (int)ncap;
      nl.rn2 = (float)(r_near * r_near);
      // boundary shell: the float32 range test runs against r_cut^2 (1 + 8e-6) and everything from r_cut^2 (1 - 8e-6)
      // up goes to the list (tile coordinates carry 2^-21 of 4 cell sides: 2e-6 of r_cut^2 in a squared separation)
      nl.rb2 = (float)(r_cut * r_cut * (1.0 - 8e-6));
      prm.rc2 = (float)(r_cut * r_cut * (1.0 + 8e-6));
      NB_CHECK_CUDA(cudaMemsetAsync(nl.count, 0, sizeof(int), st));
    }
    const int nrun = (ncs + P3M6_WARPS - 1) / P3M6_WARPS;
    const int64_t nunits = ((cell_hi + ncs - 1) / ncs - cell_lo / ncs) * nrun;
    int64_t nblk = nunits < cap ? nunits : cap;
    int* unit_counter = nullptr;
    if (ctx->p3m_sm_limit > 0) {      // SM-partitioned persistent form: every resident slot gets a CTA
      void* fl = nullptr;
      rc = nb_scratch(ctx, NB_SCR_FLAGS, 64, &fl);
      if (rc) return rc;
      unit_counter = (int*)fl;        // [0] unit counter, [1] stop flag (nb_p3m_mesh_done)
      NB_CHECK_CUDA(cudaMemsetAsync(unit_counter, 0, 2 * sizeof(int), st));
      nblk = (int64_t)ctx->num_sms * 6;
    }
    auto k6 = near ? p3m_short_v6_kernel<QD6, 6, false, CAP6, true> : p3m_short_v6_kernel<QD6, 6, false, CAP6, false>;
    NB_CHECK_CUDA(cudaFuncSetAttribute(k6, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem6));
    k6<<<(int)nblk, P3M6_WARPS * 32, smem6, st>>>(
        (const float4*)sorted_xyzw, cell_start, perm, ncs, cell_lo, cell_hi, prm, scale, acc, accumulate, ilim, CAP6,
        ovf_list, ovf_count, unit_counter, ctx->p3m_sm_limit, unit_counter ? unit_counter + 1 : nullptr, nl);
    NB_LAUNCHED(ctx);
    if (unit_counter) {
      // second, unconfined launch on the same counter: the units left when the mesh chain raised the stop flag —
      // and every unit, should the confined launch have found the SMs busy and lost its CTAs (a CTA that exits
      // on a foreign SM is not replaced).  Empty when the confined CTAs got through the list.
      k6<<<(int)nblk, P3M6_WARPS * 32, smem6, st>>>(
          (const float4*)sorted_xyzw, cell_start, perm, ncs, cell_lo, cell_hi, prm, scale, acc, accumulate, ilim,
          CAP6, ovf_list, ovf_count, unit_counter, 0, nullptr, nl);
      NB_LAUNCHED(ctx);
    }
    const int tile_cap = P3M4_CAP;
    const size_t smem4 = 96 * sizeof(float) + P3M4_WARPS * p3m4_warp_bytes(tile_cap, 64);
    auto k4 = near ? p3m_short_v4_kernel<true, 64, true> : p3m_short_v4_kernel<true, 64, false>;
    NB_CHECK_CUDA(cudaFuncSetAttribute(k4, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem4));
    k4<<<ctx->num_sms * 4, P3M4_WARPS * 32, smem4, st>>>(
        (const float4*)sorted_xyzw, cell_start, perm, ncs, cell_lo, cell_hi, prm, scale, acc, accumulate, ilim,
        tile_cap, ovf_list, ovf_count, nl);
    if (near) {
      NB_LAUNCHED(ctx);
      p3m_near_kernel<<<ctx->num_sms * 8, 256, 0, st>>>(nl.list, nl.count, nl.cap, (const float4*)sorted_xyzw, perm, pos,
                                                        box, 1.0 / box, 1.0 / (2.0 * r_s), eps * eps, r_cut * r_cut,
                                                        scale, acc);
      if (near_count_out)
        NB_CHECK_CUDA(cudaMemcpyAsync(near_count_out, nl.count, sizeof(int32_t), cudaMemcpyDeviceToDevice, st));
    }
  }
#else
  (void)pos; (void)np; (void)box; (void)r_near; (void)near_count_out;   // A/B builds: no near-pair list
#include "p3m_ab_dispatch.inc"
#endif
  NB_LAUNCHED(ctx);
  return NB_OK;
}

extern "C" int nb_p3m_short(nb_ctx* ctx, const void* sorted_xyzw, const int32_t* cell_start, const int32_t* perm,
                            int ncs, int64_t cell_lo, int64_t cell_hi, double cell_side, double r_cut,
                            double r_s, double eps, const double* coef_host, int ncoef, double u_max,
                            double scale, int accumulate, int64_t i_limit, double* acc, void* stream) {
  return p3m_short_impl(ctx, sorted_xyzw, cell_start, perm, ncs, cell_lo, cell_hi, cell_side, r_cut, r_s, eps, coef_host,
                        ncoef, u_max, scale, accumulate, i_limit, acc, nullptr, 0, 0.0, 0.0, nullptr, stream);
}

// nb_p3m_short with the pairs closer than r_near evaluated in float64 from the original positions `pos` ((np,3)
// float64, the array nb_cell_sort sorted; box = side of the periodic box).  near_count_out (device int32, may be
// NULL) receives the number of near pairs found; pairs beyond the list capacity (2 np entries, at least 2^20) were
// evaluated in float32 like nb_p3m_short does.
extern "C" int nb_p3m_short_exact(nb_ctx* ctx, const void* sorted_xyzw, const int32_t* cell_start,
                                  const int32_t* perm, int ncs, int64_t cell_lo, int64_t cell_hi, double cell_side,
                                  double r_cut, double r_s, double eps, const double* coef_host, int ncoef,
                                  double u_max, double scale, int accumulate, int64_t i_limit, double* acc,
                                  const double* pos, int64_t np, double box, double r_near, int32_t* near_count_out,
                                  void* stream) {
  NB_REQUIRE(pos && np > 0 && box > 0 && r_near > 0 && r_near < r_cut, "bad near-pair arguments");
  return p3m_short_impl(ctx, sorted_xyzw, cell_start, perm, ncs, cell_lo, cell_hi, cell_side, r_cut, r_s, eps, coef_host,
                        ncoef, u_max, scale, accumulate, i_limit, acc, pos, np, box, r_near, near_count_out, stream);
}

// Pair statistics of the same traversal (not timed; bench.py uses it to convert the kernel
// time into candidate-pair and in-range-pair rates): out[0] = candidate pairs, out[1] = pairs
// with 0 < r < r_cut.
__global__ void __launch_bounds__(P3M_WARPS * 32)
p3m_count_kernel(const float4* __restrict__ sorted, const int* __restrict__ cell_start,
                 const int* __restrict__ perm, int i_limit, int ncs, int64_t cell_lo, int64_t cell_hi, float cs,
                 float rc2, unsigned long long* __restrict__ out) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t wstride = (int64_t)gridDim.x * P3M_WARPS;
  unsigned long long cand = 0, inr = 0;
  for (int64_t c = cell_lo + (int64_t)blockIdx.x * P3M_WARPS + warp; c < cell_hi; c += wstride) {
    const int is = cell_start[c], ie = cell_start[c + 1];
    if (is == ie) continue;
    const int cz = (int)(c % ncs), cy = (int)((c / ncs) % ncs), cx = (int)(c / ((int64_t)ncs * ncs));
    for (int i = is + lane; i < ie; i += 32) {
      if (perm && perm[i] >= i_limit) continue;   // halo copy: a source, not an i-particle
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

extern "C" int nb_p3m_count_pairs(nb_ctx* ctx, const void* sorted_xyzw, const int32_t* cell_start,
                                  const int32_t* perm, int64_t i_limit, int ncs, int64_t cell_lo, int64_t cell_hi,
                                  double cell_side, double r_cut, int64_t* out2, void* stream) {
  NB_REQUIRE(ctx && sorted_xyzw && cell_start && out2, "NULL argument");
  NB_REQUIRE(ncs >= 3 && cell_lo >= 0 && cell_lo <= cell_hi && cell_hi <= (int64_t)ncs * ncs * ncs, "bad cells");
  cudaStream_t st = (cudaStream_t)stream;
  NB_CHECK_CUDA(cudaSetDevice(ctx->device));
  NB_CHECK_CUDA(cudaMemsetAsync(out2, 0, 2 * sizeof(int64_t), st));
  if (cell_lo == cell_hi) return NB_OK;
  int64_t nblk = (cell_hi - cell_lo + P3M_WARPS - 1) / P3M_WARPS;
  const int64_t cap = (int64_t)ctx->num_sms * 64;
  if (nblk > cap) nblk = cap;
  const int ilim = (i_limit <= 0 || i_limit > 0x7fffffff) ? 0x7fffffff : (int)i_limit;
  p3m_count_kernel<<<(int)nblk, P3M_WARPS * 32, 0, st>>>((const float4*)sorted_xyzw, cell_start,
                                                         i_limit > 0 ? perm : nullptr, ilim, ncs, cell_lo, cell_hi,
                                                         (float)cell_side, (float)(r_cut * r_cut),
                                                         (unsigned long long*)out2);
  NB_LAUNCHED(ctx);
  return NB_OK;
}
