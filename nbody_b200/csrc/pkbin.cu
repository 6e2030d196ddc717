// K6 — |delta_k|^2 shell binning with warp-segmented reductions.
// Reference: Power_compare.py:166-200 (np.histogram), :442-465 (np.digitize loop).
//
// Exactness of the bin assignment (north_star: bin counts bit-exact):
//   * k² = (kx²+ky²)+kz² is formed in float64 from the host table k2_tab = k_vals**2 with
//     __dadd_rn in numpy's association order (no FMA);
//   * "sqrt(k²) >= edge" is decided against host-computed thresholds in k² space
//     (thresh[e] = smallest double x with sqrt(x) >= edges[e]); sqrt is monotone and
//     correctly rounded on both sides, so the decision equals numpy's;
//   * a float32 log2 only GUESSES the bin, the thresholds correct the guess.
// Half spectrum (R2C): modes with 0 < kz < N/2 stand for themselves and their conjugate
// (weight 2); kz = 0 and kz = N/2 (N even) have weight 1 — full-grid counts are reproduced.
//
// One warp owns one (ix,iy) row at a time: along kz in [0, N/2] |k| is non-decreasing, so
// equal bins form contiguous lane segments; a segmented shuffle reduction leaves one
// (count,sum) per segment, written by the segment head into a warp-private shared-memory
// histogram (no atomics).  CTAs flush their histograms to global memory once at the end.
#include <stdlib.h>

#include "common.cuh"

#define PK_WARPS 8

template <typename C, typename T>
__global__ void __launch_bounds__(PK_WARPS * 32)
pk_shell_bin_kernel(const C* __restrict__ spec, int N, int y0, int nyl, const double* __restrict__ k2_tab,
                    const double* __restrict__ sinc_tab, const double* __restrict__ thresh,
                    double thresh_hi, int n_edges, int window_power, int clip_mode, double clip_val,
                    double norm, int closed_last, float log2_k2min, float inv_dlog2,
                    unsigned long long* __restrict__ counts, double* __restrict__ sums) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int nb = n_edges - 1;
  double* s_thresh = reinterpret_cast<double*>(smem_raw);           // n_edges
  double* s_k2 = s_thresh + n_edges;                                // N
  double* s_sinc = s_k2 + N;                                        // N
  double* s_sums = s_sinc + N;                                      // PK_WARPS * nb
  unsigned* s_cnt = reinterpret_cast<unsigned*>(s_sums + PK_WARPS * nb);  // PK_WARPS * nb

  for (int t = threadIdx.x; t < n_edges; t += blockDim.x) s_thresh[t] = thresh[t];
  for (int t = threadIdx.x; t < N; t += blockDim.x) {
    s_k2[t] = k2_tab[t];
    s_sinc[t] = sinc_tab[t];
  }
  for (int t = threadIdx.x; t < PK_WARPS * nb; t += blockDim.x) {
    s_sums[t] = 0.0;
    s_cnt[t] = 0u;
  }
  __syncthreads();

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nzh = N / 2 + 1;
  const int kz_nyq = (N % 2 == 0) ? N / 2 : -1;
  double* my_sums = s_sums + warp * nb;
  unsigned* my_cnt = s_cnt + warp * nb;
  const int64_t nrows = (int64_t)N * nyl;   // spec is [N (kx)][nyl (ky = y0..)][N/2+1]
  const int64_t wstride = (int64_t)gridDim.x * PK_WARPS;

  for (int64_t row = (int64_t)blockIdx.x * PK_WARPS + warp; row < nrows; row += wstride) {
    const int j = y0 + (int)(row % nyl), i = (int)(row / nyl);
    const double kxy2 = __dadd_rn(s_k2[i], s_k2[j]);
    const double sxy = __dmul_rn(s_sinc[i], s_sinc[j]);
    const C* rowp = spec + row * nzh;
    for (int kz0 = 0; kz0 < nzh; kz0 += 32) {
      const int kz = kz0 + lane;
      int bin = nb;  // "invalid / beyond the last bin"
      T val = 0;
      unsigned cnt = 0;
      if (kz < nzh) {
        const double x = __dadd_rn(kxy2, s_k2[kz]);
        if (x > 0.0) {
          int g = (int)floorf((__log2f((float)x) - log2_k2min) * inv_dlog2);
          g = max(-1, min(g, n_edges - 1));
          while (g + 1 < n_edges && s_thresh[g + 1] <= x) ++g;
          while (g >= 0 && s_thresh[g] > x) --g;
          // g = (#edges <= k) - 1
          if (g == nb && closed_last && x < thresh_hi) g = nb - 1;  // k == edges[-1]
          if (g >= 0 && g < nb) {
            bin = g;
            const C v = rowp[kz];
            T p = (T)(v.x * v.x + v.y * v.y) * (T)norm;
            if (window_power) {
              double w = __dmul_rn(sxy, s_sinc[kz]);
              w = __dmul_rn(w, w);
              if (clip_mode == 1) {
                if (w == 0.0) w = 1.0;
              } else if (clip_mode == 2) {
                w = fmax(w, clip_val);
              }
              if (window_power == 2) w = __dmul_rn(w, w);
              p = p / (T)w;
            }
            const unsigned wt = (kz == 0 || kz == kz_nyq) ? 1u : 2u;
            val = p * (T)wt;
            cnt = wt;
          } else if (g < 0) {
            bin = -1;
          }
        } else {
          bin = -1;  // k == 0 is masked out (Power_compare.py:183)
        }
      }
      // segmented (by equal, monotone bin) suffix reduction
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const int ob = __shfl_down_sync(0xffffffffu, bin, d);
        const T ov = __shfl_down_sync(0xffffffffu, val, d);
        const unsigned oc = __shfl_down_sync(0xffffffffu, cnt, d);
        if (lane + d < 32 && ob == bin) {
          val += ov;
          cnt += oc;
        }
      }
      const int pb = __shfl_up_sync(0xffffffffu, bin, 1);
      const bool head = (lane == 0) || (pb != bin);
      if (head && bin >= 0 && bin < nb) {
        my_sums[bin] += (double)val;
        my_cnt[bin] += cnt;
      }
      __syncwarp();
    }
  }
  __syncthreads();
  for (int b = threadIdx.x; b < nb; b += blockDim.x) {
    double s = 0.0;
    unsigned long long c = 0;
#pragma unroll
    for (int w = 0; w < PK_WARPS; ++w) {
      s += s_sums[w * nb + b];
      c += s_cnt[w * nb + b];
    }
    if (c) {
      atomicAdd(&sums[b], s);
      atomicAdd(&counts[b], c);
    }
  }
}

// ---------------------------------------------------------------------------------------
// pk_shell_bin_rows_kernel — the complex64 production path (HBM-bound form).
//
// The kernel above spends ~100 warp instructions per 32 modes on the bin search and the
// segmented shuffle reduction and keeps one 256-byte load in flight per warp.  This one turns
// the mapping by 90 degrees: a LANE owns one (kx,ky) row and walks it along kz, so the
// monotone bin index becomes a running register (g, next two thresholds) and a partial sum
// stays in registers until the row crosses a shell boundary (about every 27 modes at the C5
// size).  The 32 rows of a warp are contiguous in memory; PKF_CHUNK-mode pieces of them (8 by
// default) are staged through shared memory with 8-byte cp.async in a 2- or 3-deep pipeline
// (coalesced 64-byte row segments in, conflict-free LDS.128 per lane out; row pitch 80 B): 4 KB
// per warp in flight, 20 warps per SM.
//   * kz = 1 .. N/2-1 (weight 2) is the streamed range; the two weight-1 columns kz = 0 and
//     kz = N/2 are binned by a generic search at the start of each row;
//   * a piece with at most one shell boundary is summed branch-free into a "current"
//     and a "next" float accumulator (x < th1 selects); pieces with more boundaries (low |k|
//     only) take a per-mode loop;
//   * partial sums are flushed into ONE shared histogram per CTA (native u32 atomics for the
//     counts, CAS-loop f64 adds for the sums — a flush happens once per ~27 modes), which goes
//     to global memory once, scaled by `norm`, at the end.
// Bin decisions use the same float64 k² tables and thresholds as above: counts stay bit-exact.
#ifndef PKF_WARPS
#define PKF_WARPS 4
#endif
#ifndef PKF_CHUNK
#define PKF_CHUNK 8            // modes per staged row piece (r1h A/B at 1024^3: 16 / 4 CTAs 1.81 ms, 8 / 5 CTAs 1.39 ms, 8 / 6 CTAs 2.14 ms)
#endif
#define PKF_PITCH (PKF_CHUNK * 8 + 16)          // bytes per staged row piece
#define PKF_STAGE_BYTES (32 * PKF_PITCH)
// L2 sector promotion of the staged row pieces: a 64-byte piece is half of a 128-byte line whose other half
// is the next piece of the same row
#ifndef PKF_L2HINT
#define PKF_L2HINT ""            // ".L2::128B" measured: no difference (1.3250 vs 1.3246 ms)
#endif
#ifndef PKF_MIN_CTAS
#define PKF_MIN_CTAS 5         // 96 registers: 20 warps per SM
#endif

__device__ __forceinline__ int pkf_find_bin(double x, const double* __restrict__ s_th, int n_edges,
                                            float log2_k2min, float inv_dlog2) {
  int g = (int)floorf((__log2f((float)x) - log2_k2min) * inv_dlog2);
  g = max(-1, min(g, n_edges - 1));
  while (g + 1 < n_edges && s_th[g + 1] <= x) ++g;
  while (g >= 0 && s_th[g] > x) --g;
  return g;   // (#thresholds <= x) - 1
}

__device__ __forceinline__ void pkf_flush(double* s_sums, unsigned* s_cnt, int nb, int g, double acc,
                                          unsigned cnt, unsigned wt) {
  if (g >= 0 && g < nb && cnt) {
    atomicAdd(&s_cnt[g], cnt * wt);
    atomicAdd(&s_sums[g], acc * (double)wt);
  }
}

// Warp-collective flush (all 32 lanes call it, converged).  The 32 rows of a warp are
// neighbours in k-space, so lanes that cross a shell boundary in the same piece mostly flush
// into the SAME bin; sent one by one they serialise in the CAS loop of the shared f64 add.
// Lanes are grouped by bin (leader loop), each group is summed with shuffles and its leader
// issues one add.
__device__ __forceinline__ void pkf_flush_warp(double* s_sums, unsigned* s_cnt, int nb, bool want, int g,
                                               double acc, unsigned cnt, unsigned wt, int lane) {
  bool pend = want && g >= 0 && g < nb && cnt != 0;
  unsigned m = __ballot_sync(0xffffffffu, pend);
  while (m) {
    const int leader = __ffs(m) - 1;
    const int gl = __shfl_sync(0xffffffffu, g, leader);
    const bool mine = pend && g == gl;
    const unsigned mm = __ballot_sync(0xffffffffu, mine);
    double a = mine ? acc : 0.0;
    unsigned c = mine ? cnt : 0u;
    if (mm & (mm - 1)) {   // more than one lane in this group
      a = nb_warp_sum(a);
      c = __reduce_add_sync(0xffffffffu, c);
    }
    if (lane == leader) {
      atomicAdd(&s_cnt[gl], c * wt);
      atomicAdd(&s_sums[gl], a * (double)wt);
    }
    m &= ~mm;
    pend = pend && !mine;
  }
}

template <bool WINDOW>
__device__ __forceinline__ float pkf_power(float2 v, double sxy, double sz, int window_power, int clip_mode,
                                           double clip_val) {
  float p = v.x * v.x + v.y * v.y;
  if constexpr (WINDOW) {
    double w = __dmul_rn(sxy, sz);
    w = __dmul_rn(w, w);
    if (clip_mode == 1) {
      if (w == 0.0) w = 1.0;
    } else if (clip_mode == 2) {
      w = fmax(w, clip_val);
    }
    if (window_power == 2) w = __dmul_rn(w, w);
    p = p / (float)w;
  }
  return p;
}

template <bool WINDOW, int STAGES>
__global__ void __launch_bounds__(PKF_WARPS * 32, PKF_MIN_CTAS)
pk_shell_bin_rows_kernel(const float2* __restrict__ spec, int N, int y0, int nyl, const double* __restrict__ k2_tab,
                         const double* __restrict__ sinc_tab, const double* __restrict__ thresh, double thresh_hi,
                         int n_edges, int window_power, int clip_mode, double clip_val, double norm,
                         int closed_last, float log2_k2min, float inv_dlog2,
                         unsigned long long* __restrict__ counts, double* __restrict__ sums) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int nb = n_edges - 1;
  unsigned char* s_stage = smem_raw;                                              // 16-byte aligned
  double* s_th = reinterpret_cast<double*>(smem_raw + PKF_WARPS * STAGES * PKF_STAGE_BYTES);  // n_edges + 3
  // k² and the sinc factor are even in the FFT index (fftfreq: k[N-m] = -k[m] exactly), so only the
  // entries 0 .. N/2 are kept (4 KB instead of 8 KB at N = 1024) and kx, ky fold with min(m, N-m)
  const int nfold = N / 2 + 1;
  double* s_k2 = s_th + n_edges + 3;                                              // nfold
  double* s_sinc = s_k2 + nfold;                                                  // nfold (WINDOW only)
  double* s_sums = s_sinc + (WINDOW ? nfold : 0);                                 // nb
  unsigned* s_cnt = reinterpret_cast<unsigned*>(s_sums + nb);                     // nb

  for (int t = threadIdx.x; t < n_edges + 3; t += blockDim.x) {
    double v = __longlong_as_double(0x7ff0000000000000ll);                        // +inf sentinels
    if (t < n_edges) v = (t == n_edges - 1 && closed_last) ? thresh_hi : thresh[t];
    s_th[t] = v;
  }
  for (int t = threadIdx.x; t < nfold; t += blockDim.x) {
    s_k2[t] = k2_tab[t];
    if (WINDOW) s_sinc[t] = sinc_tab[t];
  }
  for (int t = threadIdx.x; t < nb; t += blockDim.x) {
    s_sums[t] = 0.0;
    s_cnt[t] = 0u;
  }
  __syncthreads();

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nzh = N / 2 + 1;
  const int kz_nyq = (N % 2 == 0) ? N / 2 : -1;
  const int kz_hi = (N % 2 == 0) ? N / 2 : nzh;          // streamed range is kz = 1 .. kz_hi-1
  const int nmain = kz_hi - 1;
  const int nchunks = (nmain + PKF_CHUNK - 1) / PKF_CHUNK;
  const int64_t nrows = (int64_t)N * nyl;
  const int64_t ngroups = (nrows + 31) / 32;
  unsigned char* my_stage = s_stage + warp * STAGES * PKF_STAGE_BYTES;
  const unsigned my_stage_s = (unsigned)__cvta_generic_to_shared(my_stage);
  constexpr int RPI = 32 / PKF_CHUNK;          // rows covered by one cp.async step of the warp
  static_assert(PKF_CHUNK == 8 || PKF_CHUNK == 16 || PKF_CHUNK == 32, "PKF_CHUNK: 8, 16 or 32");
  const int ld_r = lane / PKF_CHUNK, ld_m = lane % PKF_CHUNK;

  // Work units.  Every warp takes `rounds` whole groups (32 rows, all of kz); the ngroups % nwarps groups
  // left over would cost a full extra round for a few warps (1024^3 on 2960 warps: 11.07 -> 12 rounds), so
  // each of them is cut along kz into `pieces` chunk ranges that are spread over all warps instead.
  const int64_t nwarps = (int64_t)gridDim.x * PKF_WARPS;
  const int64_t wid = (int64_t)blockIdx.x * PKF_WARPS + warp;
  const int64_t rounds = ngroups / nwarps;
  const int64_t left = ngroups - rounds * nwarps;
  int pieces = 1;
  if (left > 0) pieces = (int)min((int64_t)max(nchunks, 1), nwarps / left);
  const int64_t tail_units = left * pieces;
  for (int64_t it = 0; it <= rounds; ++it) {
    int64_t grp;
    int c_lo = 0, c_hi = nchunks;
    bool first = true;
    if (it < rounds) {
      grp = wid + it * nwarps;
    } else {
      if (wid >= tail_units) break;
      grp = rounds * nwarps + wid / pieces;
      const int pc = (int)(wid % pieces);
      c_lo = (int)((int64_t)pc * nchunks / pieces);
      c_hi = (int)((int64_t)(pc + 1) * nchunks / pieces);
      first = pc == 0;
    }
    const int64_t rowbase = grp * 32;
    const int64_t row = rowbase + lane;
    const bool active = row < nrows;
    const int64_t rowc = active ? row : nrows - 1;
    const int jj = y0 + (int)(rowc % nyl), ii = (int)(rowc / nyl);
    const int j = min(jj, N - jj), i = min(ii, N - ii);
    const double kxy2 = __dadd_rn(s_k2[i], s_k2[j]);
    double sxy = 1.0;
    if (WINDOW) sxy = __dmul_rn(s_sinc[i], s_sinc[j]);

    // stage issue: piece c of the 32 rows -> buffer c % STAGES (lane = PKF_CHUNK*(row % RPI) + mode)
    const bool full = rowbase + 32 <= nrows;
    const float2* src0 = spec + (rowbase + ld_r) * nzh;
    const int64_t stride2 = RPI * (int64_t)nzh;
    auto issue = [&](int c) {
      if (c < c_hi) {
        const int kz = 1 + c * PKF_CHUNK + ld_m;
        if (kz < kz_hi) {
          const unsigned dst0 = my_stage_s + (unsigned)((c % STAGES) * PKF_STAGE_BYTES + ld_m * 8 + ld_r * PKF_PITCH);
          if (full) {   // all 32 rows exist: one pointer, a constant stride of RPI rows
            const float2* src = src0 + kz;
#pragma unroll
            for (int t = 0; t < PKF_CHUNK; ++t)
              asm volatile("cp.async.ca.shared.global" PKF_L2HINT " [%0], [%1], 8;" ::"r"(dst0 + (unsigned)(RPI * t * PKF_PITCH)),
                           "l"(src + (int64_t)t * stride2)
                           : "memory");
          } else {      // last group: clamp the row
#pragma unroll
            for (int t = 0; t < PKF_CHUNK; ++t) {
              int64_t gr = rowbase + RPI * t + ld_r;
              gr = gr < nrows ? gr : nrows - 1;
              asm volatile("cp.async.ca.shared.global" PKF_L2HINT " [%0], [%1], 8;" ::"r"(dst0 + (unsigned)(RPI * t * PKF_PITCH)),
                           "l"(spec + gr * nzh + kz)
                           : "memory");
            }
          }
        }
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
    };
#pragma unroll
    for (int c = 0; c < STAGES - 1; ++c) issue(c_lo + c);

    // the two weight-1 columns (generic search; every lane does this once per row)
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const int kz = e == 0 ? 0 : kz_nyq;
      if (kz < 0 || !first) continue;  // warp-uniform
      const double x = __dadd_rn(kxy2, s_k2[kz]);
      int ge = -1;
      float pe = 0.f;
      if (active && x > 0.0) {
        ge = pkf_find_bin(x, s_th, n_edges, log2_k2min, inv_dlog2);
        if (ge >= 0 && ge < nb)
          pe = pkf_power<WINDOW>(__ldg(spec + row * nzh + kz), sxy, WINDOW ? s_sinc[kz] : 1.0, window_power,
                                 clip_mode, clip_val);
      }
      __syncwarp();
      pkf_flush_warp(s_sums, s_cnt, nb, active && x > 0.0, ge, (double)pe, 1u, 1u, lane);
    }

    // running state of the streamed range
    int g = c_lo < c_hi ? pkf_find_bin(__dadd_rn(kxy2, s_k2[1 + c_lo * PKF_CHUNK]), s_th, n_edges, log2_k2min, inv_dlog2) : nb;
    double th1 = s_th[g + 1], th2 = s_th[g + 2], th3 = s_th[g + 3];
    double acc = 0.0;
    unsigned cnt = 0;

    for (int c = c_lo; c < c_hi; ++c) {
      issue(c + STAGES - 1);
      asm volatile("cp.async.wait_group %0;" ::"n"(STAGES - 1) : "memory");
      __syncwarp();
      const int kz0 = 1 + c * PKF_CHUNK;
      const int nv = min(PKF_CHUNK, nmain - c * PKF_CHUNK);
      const unsigned char* piece = my_stage + (c % STAGES) * PKF_STAGE_BYTES + lane * PKF_PITCH;
      const double* k2c = s_k2 + kz0;
      const double x_last = __dadd_rn(kxy2, k2c[nv - 1]);
      // up to two closed (bin, sum, count) triples of the branch-free path, flushed collectively below
      bool flA = false, flB = false;
      int flA_g = -1, flB_g = -1;
      double flA_acc = 0.0, flB_acc = 0.0;
      unsigned flA_cnt = 0, flB_cnt = 0;
      if (x_last < th3) {
        // at most two shell boundaries (th1 <= th2) inside this piece.  x is non-decreasing in m, so
        // the numbers n1 / n2 of modes below th1 / th2 come from a 5-step counting search (same
        // float64 add and compare as the per-mode test), and the 16 powers are then split by index.
        auto count_below = [&](double th) {
          int cb = 0;
#pragma unroll
          for (int sstep = PKF_CHUNK; sstep > 0; sstep >>= 1) {
            const int t = cb + sstep;
            if (t <= nv) {
              if (__dadd_rn(kxy2, k2c[t - 1]) < th) cb = t;
            }
          }
          return cb;
        };
        const int n1 = count_below(th1);
        const int n2 = x_last >= th2 ? count_below(th2) : nv;   // (no warp collective here: lanes diverge)
        float f1 = 0.f, f2 = 0.f, f3 = 0.f;
#pragma unroll
        for (int mm = 0; mm < PKF_CHUNK / 2; ++mm) {
          const float4 q = *reinterpret_cast<const float4*>(piece + mm * 16);
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            const int m = 2 * mm + h;
            if (m < nv) {
              const float2 v = h == 0 ? make_float2(q.x, q.y) : make_float2(q.z, q.w);
              const float p = pkf_power<WINDOW>(v, sxy, WINDOW ? s_sinc[kz0 + m] : 1.0, window_power, clip_mode,
                                                clip_val);
              const bool b1 = m < n1, b2 = m < n2;
              f1 += b1 ? p : 0.f;
              f2 += (!b1 && b2) ? p : 0.f;
              f3 += b2 ? 0.f : p;
            }
          }
        }
        acc += (double)f1;
        cnt += (unsigned)n1;
        if (n1 < nv) {            // crossed th1: the running bin g is closed
          flA = active;
          flA_g = g;
          flA_acc = acc;
          flA_cnt = cnt;
          if (n2 < nv) {          // crossed th2 as well: bin g+1 lies entirely inside this piece
            flB = active;
            flB_g = g + 1;
            flB_acc = (double)f2;
            flB_cnt = (unsigned)(n2 - n1);
            g += 2;
            acc = (double)f3;
            cnt = (unsigned)(nv - n2);
            th1 = th3;
          } else {
            g += 1;
            acc = (double)f2;
            cnt = (unsigned)(nv - n1);
            th1 = th2;
          }
          th2 = s_th[min(g + 2, n_edges + 2)];
          th3 = s_th[min(g + 3, n_edges + 2)];
        }
      } else {
        for (int m = 0; m < nv; ++m) {
          const double x = __dadd_rn(kxy2, s_k2[kz0 + m]);
          const float2 v = *reinterpret_cast<const float2*>(piece + m * 8);
          const float p = pkf_power<WINDOW>(v, sxy, WINDOW ? s_sinc[kz0 + m] : 1.0, window_power, clip_mode,
                                            clip_val);
          if (x >= th1) {
            if (active) pkf_flush(s_sums, s_cnt, nb, g, acc, cnt, 2u);
            acc = 0.0;
            cnt = 0;
            do {
              ++g;
              th1 = th2;
              th2 = th3;
              th3 = s_th[min(g + 3, n_edges + 2)];
            } while (x >= th1);
          }
          acc += (double)p;
          cnt += 1;
        }
      }
      __syncwarp();   // everyone is done with this buffer before it is refilled
      pkf_flush_warp(s_sums, s_cnt, nb, flA, flA_g, flA_acc, flA_cnt, 2u, lane);
      if (__any_sync(0xffffffffu, flB)) pkf_flush_warp(s_sums, s_cnt, nb, flB, flB_g, flB_acc, flB_cnt, 2u, lane);
    }
    __syncwarp();
    pkf_flush_warp(s_sums, s_cnt, nb, active, g, acc, cnt, 2u, lane);
    asm volatile("cp.async.wait_group 0;" ::: "memory");
  }
  __syncthreads();
  for (int b = threadIdx.x; b < nb; b += blockDim.x) {
    const unsigned c = s_cnt[b];
    if (c) {
      atomicAdd(&sums[b], s_sums[b] * norm);
      atomicAdd(&counts[b], (unsigned long long)c);
    }
  }
}

// Bins the half spectrum `spec` ([N][N][N/2+1]) — shared by nb_power_spectrum and the
// slab-decomposed path.
// NB_PK_KERNEL=1 selects the warp-per-row kernel for complex64 too (development / A-B runs)
static bool pk_rows_enabled() {
  static const bool on = [] { const char* e = getenv("NB_PK_KERNEL"); return !(e && atoi(e) == 1); }();
  return on;
}

int nb_pk_bin_spectrum(nb_ctx* ctx, const void* spec, int N, int y0, int nyl, int prec, const double* k2_tab,
                       const double* sinc_tab, const double* thresh, double thresh_hi, int n_edges,
                       int window_power, int clip_mode, double clip_val, double norm, int closed_last,
                       double k2min_host, double k2max_host, int64_t* counts, double* sums,
                       cudaStream_t st) {
  const int nb = n_edges - 1;
  const size_t smem = sizeof(double) * ((size_t)n_edges + 2 * (size_t)N + (size_t)PK_WARPS * nb) +
                      sizeof(unsigned) * (size_t)PK_WARPS * nb;
  NB_REQUIRE(smem <= 200 * 1024, "too many bins / mesh too large for the shared-memory tables");
  // bin guess: edges are uniform in log k, so in log2(k²) the spacing is dlog2 per edge
  const float log2_k2min = (float)log2(k2min_host);
  const float inv_dlog2 = (float)((double)(n_edges - 1) / (log2(k2max_host) - log2(k2min_host)));
  int nblk = ctx->num_sms * 2;
  const int64_t need = ((int64_t)N * nyl + PK_WARPS - 1) / PK_WARPS;
  if (nblk > need) nblk = (int)need;
  if (prec == NB_F64) {
    auto kern = pk_shell_bin_kernel<double2, double>;
    NB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<nblk, PK_WARPS * 32, smem, st>>>((const double2*)spec, N, y0, nyl, k2_tab, sinc_tab, thresh, thresh_hi,
                                            n_edges, window_power, clip_mode, clip_val, norm, closed_last,
                                            log2_k2min, inv_dlog2, (unsigned long long*)counts, sums);
  } else if (pk_rows_enabled() && N >= 4) {
    // production path for complex64 spectra: lane-per-row streaming kernel
    static const int stages = [] { const char* e = getenv("NB_PK_STAGES"); return e && atoi(e) == 3 ? 3 : 2; }();
    const size_t smem2 = (size_t)PKF_WARPS * stages * PKF_STAGE_BYTES +
                         sizeof(double) * ((size_t)n_edges + 3 + (size_t)(N / 2 + 1) * (window_power ? 2 : 1) + nb) +
                         sizeof(unsigned) * (size_t)nb;
    NB_REQUIRE(smem2 <= 200 * 1024, "too many bins / mesh too large for the shared-memory tables");
    const int64_t ngroups = ((int64_t)N * nyl + 31) / 32;
    int per_sm = (int)((size_t)220 * 1024 / (smem2 + 1024));
    per_sm = per_sm < 1 ? 1 : (per_sm > PKF_MIN_CTAS ? PKF_MIN_CTAS : per_sm);
    int nblk2 = ctx->num_sms * per_sm;
    const int64_t need2 = (ngroups + PKF_WARPS - 1) / PKF_WARPS;
    if (nblk2 > need2) nblk2 = (int)need2;
    // NB_PK_GRID: force the CTA count (tests use it to reach the split-tail path at small N)
    if (const char* e = getenv("NB_PK_GRID")) { const int v = atoi(e); if (v >= 1 && v < nblk2) nblk2 = v; }
    auto launch = [&](auto kern) -> int {
      NB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2));
      kern<<<nblk2, PKF_WARPS * 32, smem2, st>>>((const float2*)spec, N, y0, nyl, k2_tab, sinc_tab, thresh, thresh_hi,
                                                 n_edges, window_power, clip_mode, clip_val, norm, closed_last,
                                                 log2_k2min, inv_dlog2, (unsigned long long*)counts, sums);
      return NB_OK;
    };
    int rc2;
    if (stages == 3) rc2 = window_power ? launch(pk_shell_bin_rows_kernel<true, 3>) : launch(pk_shell_bin_rows_kernel<false, 3>);
    else rc2 = window_power ? launch(pk_shell_bin_rows_kernel<true, 2>) : launch(pk_shell_bin_rows_kernel<false, 2>);
    if (rc2) return rc2;
  } else {
    auto kern = pk_shell_bin_kernel<float2, float>;
    NB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<nblk, PK_WARPS * 32, smem, st>>>((const float2*)spec, N, y0, nyl, k2_tab, sinc_tab, thresh, thresh_hi,
                                            n_edges, window_power, clip_mode, clip_val, norm, closed_last,
                                            log2_k2min, inv_dlog2, (unsigned long long*)counts, sums);
  }
  NB_LAUNCHED(ctx);
  return NB_OK;
}

// Upstream interlacing for the estimator (CONCEPT `powerspec_options['interlace']`,
// SnapVsPK/Uncertain.param:43-45): s1 <- 0.5 (s1 + s2 · e^{+i k·s}), s = half a cell along x, y and z,
// s2 the spectrum of the deposit on the half-cell-shifted mesh.  Same phase convention and Nyquist
// treatment (Hermitian part of e^{±iπ/2} = 0) as the force path's K2 (kspace.cu).
template <typename C2, typename T>
__global__ void __launch_bounds__(256)
pk_interlace_combine_kernel(C2* __restrict__ s1, const C2* __restrict__ s2, int N) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  T* s_pc = reinterpret_cast<T*>(smem_raw);
  T* s_ps = s_pc + N;
  const int nyq = (N % 2 == 0) ? N / 2 : -1;
  for (int m = threadIdx.x; m < N; m += blockDim.x) {
    const int ms = (m <= N / 2 && m != nyq) ? m : m - N;
    double sn, cs;
    sincospi((double)ms / (double)N, &sn, &cs);
    s_pc[m] = (m == nyq) ? (T)0 : (T)cs;
    s_ps[m] = (m == nyq) ? (T)0 : (T)sn;
  }
  __syncthreads();
  const int nzh = N / 2 + 1;
  const int64_t nrows = (int64_t)N * N;
  const int warps = blockDim.x >> 5, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int64_t row = (int64_t)blockIdx.x * warps + warp; row < nrows; row += (int64_t)gridDim.x * warps) {
    const int j = (int)(row % N), i = (int)(row / N);
    const T pcxy = s_pc[i] * s_pc[j] - s_ps[i] * s_ps[j];
    const T psxy = s_pc[i] * s_ps[j] + s_ps[i] * s_pc[j];
    const int64_t base = row * nzh;
    for (int kz = lane; kz < nzh; kz += 32) {
      const T pc = pcxy * s_pc[kz] - psxy * s_ps[kz];
      const T ps = pcxy * s_ps[kz] + psxy * s_pc[kz];
      C2 r = s1[base + kz];
      const C2 q = s2[base + kz];
      r.x = (T)0.5 * (r.x + q.x * pc - q.y * ps);
      r.y = (T)0.5 * (r.y + q.x * ps + q.y * pc);
      s1[base + kz] = r;
    }
  }
}

static int pk_forward(nb_ctx* ctx, const void* delta, int N, int prec, void* spec, cudaStream_t st) {
  cufftHandle fwd;
  int rc;
  NbPhase phf(ctx, NB_PH_FFT_FWD, st);
  if (prec == NB_F64) {
    if ((rc = nb_plan3d(ctx, N, CUFFT_D2Z, st, &fwd))) return rc;
    NB_CHECK_CUFFT(cufftExecD2Z(fwd, (cufftDoubleReal*)const_cast<void*>(delta), (cufftDoubleComplex*)spec));
  } else {
    if ((rc = nb_plan3d(ctx, N, CUFFT_R2C, st, &fwd))) return rc;
    NB_CHECK_CUFFT(cufftExecR2C(fwd, (cufftReal*)const_cast<void*>(delta), (cufftComplex*)spec));
  }
  return NB_OK;
}

static int pk_spectrum_impl(nb_ctx* ctx, const void* delta, const void* delta2, int N, int prec, const double* k2_tab,
                            const double* sinc_tab, const double* thresh, double thresh_first, double thresh_last,
                            double thresh_hi, int n_edges, int window_power, int clip_mode, double clip_val,
                            double norm, int closed_last, int64_t* counts, double* sums, void* stream) {
  NB_REQUIRE(ctx && delta && k2_tab && sinc_tab && thresh && counts && sums, "NULL argument");
  NB_REQUIRE(N >= 2 && n_edges >= 2, "bad N / n_edges");
  NB_REQUIRE(prec == NB_F32 || prec == NB_F64, "bad prec");
  NB_REQUIRE(window_power >= 0 && window_power <= 2 && clip_mode >= 0 && clip_mode <= 2, "bad window options");
  cudaStream_t st = (cudaStream_t)stream;
  NB_CHECK_CUDA(cudaSetDevice(ctx->device));
  const int64_t nspec = (int64_t)N * N * (N / 2 + 1);
  const size_t cbytes = (size_t)nspec * (prec == NB_F64 ? 16 : 8);
  void *spec = nullptr, *spec2 = nullptr;
  int rc = nb_scratch(ctx, NB_SCR_SPEC0, cbytes, &spec);
  if (rc) return rc;
  if ((rc = pk_forward(ctx, delta, N, prec, spec, st))) return rc;
  if (delta2) {
    if ((rc = nb_scratch(ctx, NB_SCR_SPEC1, cbytes, &spec2))) return rc;
    if ((rc = pk_forward(ctx, delta2, N, prec, spec2, st))) return rc;
    int nblk = ctx->num_sms * 4;
    const int64_t need = ((int64_t)N * N + 7) / 8;
    if (nblk > need) nblk = (int)need;
    if (prec == NB_F64)
      pk_interlace_combine_kernel<double2, double><<<nblk, 256, 2 * sizeof(double) * N, st>>>((double2*)spec, (const double2*)spec2, N);
    else
      pk_interlace_combine_kernel<float2, float><<<nblk, 256, 2 * sizeof(float) * N, st>>>((float2*)spec, (const float2*)spec2, N);
    NB_LAUNCHED(ctx);
  }
  NB_CHECK_CUDA(cudaMemsetAsync(counts, 0, sizeof(int64_t) * (n_edges - 1), st));
  NB_CHECK_CUDA(cudaMemsetAsync(sums, 0, sizeof(double) * (n_edges - 1), st));
  NB_REQUIRE(thresh_first > 0 && thresh_last > thresh_first, "bin thresholds must be positive and increasing");
  NbPhase ph(ctx, NB_PH_PK_BIN, st);
  return nb_pk_bin_spectrum(ctx, spec, N, 0, N, prec, k2_tab, sinc_tab, thresh, thresh_hi, n_edges, window_power,
                            clip_mode, clip_val, norm, closed_last, thresh_first, thresh_last, counts, sums, st);
}

extern "C" int nb_power_spectrum(nb_ctx* ctx, const void* delta, int N, int prec, const double* k2_tab,
                                 const double* sinc_tab, const double* thresh, double thresh_first,
                                 double thresh_last, double thresh_hi, int n_edges, int window_power,
                                 int clip_mode, double clip_val, double norm, int closed_last,
                                 int64_t* counts, double* sums, void* stream) {
  return pk_spectrum_impl(ctx, delta, nullptr, N, prec, k2_tab, sinc_tab, thresh, thresh_first, thresh_last,
                          thresh_hi, n_edges, window_power, clip_mode, clip_val, norm, closed_last, counts, sums,
                          stream);
}

extern "C" int nb_power_spectrum_interlaced(nb_ctx* ctx, const void* delta, const void* delta_shifted, int N,
                                            int prec, const double* k2_tab, const double* sinc_tab,
                                            const double* thresh, double thresh_first, double thresh_last,
                                            double thresh_hi, int n_edges, int window_power, int clip_mode,
                                            double clip_val, double norm, int closed_last, int64_t* counts,
                                            double* sums, void* stream) {
  NB_REQUIRE(delta_shifted, "NULL argument");
  return pk_spectrum_impl(ctx, delta, delta_shifted, N, prec, k2_tab, sinc_tab, thresh, thresh_first, thresh_last,
                          thresh_hi, n_edges, window_power, clip_mode, clip_val, norm, closed_last, counts, sums,
                          stream);
}

// Slab form: bins the y-slab [N][ny_local][N/2+1] (complex64) a rank holds after the slab-FFT
// transpose and ADDS into counts / sums (the caller zeroes them and all-reduces afterwards).
extern "C" int nb_pk_bin_slab(nb_ctx* ctx, const void* spec, int N, int y0, int ny_local, const double* k2_tab,
                              const double* sinc_tab, const double* thresh, double thresh_first,
                              double thresh_last, double thresh_hi, int n_edges, int window_power, int clip_mode,
                              double clip_val, double norm, int closed_last, int64_t* counts, double* sums,
                              void* stream) {
  NB_REQUIRE(ctx && spec && k2_tab && sinc_tab && thresh && counts && sums, "NULL argument");
  NB_REQUIRE(N >= 2 && n_edges >= 2 && y0 >= 0 && ny_local >= 1 && y0 + ny_local <= N, "bad N / n_edges / slab");
  NB_REQUIRE(window_power >= 0 && window_power <= 2 && clip_mode >= 0 && clip_mode <= 2, "bad window options");
  NB_REQUIRE(thresh_first > 0 && thresh_last > thresh_first, "bin thresholds must be positive and increasing");
  NB_CHECK_CUDA(cudaSetDevice(ctx->device));
  return nb_pk_bin_spectrum(ctx, spec, N, y0, ny_local, NB_F32, k2_tab, sinc_tab, thresh, thresh_hi, n_edges,
                            window_power, clip_mode, clip_val, norm, closed_last, thresh_first, thresh_last,
                            counts, sums, (cudaStream_t)stream);
}
