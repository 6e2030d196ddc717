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

// Bins the half spectrum `spec` ([N][N][N/2+1]) — shared by nb_power_spectrum and the
// slab-decomposed path.
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

extern "C" int nb_power_spectrum(nb_ctx* ctx, const void* delta, int N, int prec, const double* k2_tab,
                                 const double* sinc_tab, const double* thresh, double thresh_first,
                                 double thresh_last, double thresh_hi, int n_edges, int window_power,
                                 int clip_mode, double clip_val, double norm, int closed_last,
                                 int64_t* counts, double* sums, void* stream) {
  NB_REQUIRE(ctx && delta && k2_tab && sinc_tab && thresh && counts && sums, "NULL argument");
  NB_REQUIRE(N >= 2 && n_edges >= 2, "bad N / n_edges");
  NB_REQUIRE(prec == NB_F32 || prec == NB_F64, "bad prec");
  NB_REQUIRE(window_power >= 0 && window_power <= 2 && clip_mode >= 0 && clip_mode <= 2, "bad window options");
  cudaStream_t st = (cudaStream_t)stream;
  NB_CHECK_CUDA(cudaSetDevice(ctx->device));
  const int64_t nspec = (int64_t)N * N * (N / 2 + 1);
  void* spec = nullptr;
  int rc = nb_scratch(ctx, NB_SCR_SPEC0, (size_t)nspec * (prec == NB_F64 ? 16 : 8), &spec);
  if (rc) return rc;
  cufftHandle fwd;
  {
    NbPhase phf(ctx, NB_PH_FFT_FWD, st);
    if (prec == NB_F64) {
      if ((rc = nb_plan3d(ctx, N, CUFFT_D2Z, st, &fwd))) return rc;
      NB_CHECK_CUFFT(cufftExecD2Z(fwd, (cufftDoubleReal*)const_cast<void*>(delta), (cufftDoubleComplex*)spec));
    } else {
      if ((rc = nb_plan3d(ctx, N, CUFFT_R2C, st, &fwd))) return rc;
      NB_CHECK_CUFFT(cufftExecR2C(fwd, (cufftReal*)const_cast<void*>(delta), (cufftComplex*)spec));
    }
  }
  NB_CHECK_CUDA(cudaMemsetAsync(counts, 0, sizeof(int64_t) * (n_edges - 1), st));
  NB_CHECK_CUDA(cudaMemsetAsync(sums, 0, sizeof(double) * (n_edges - 1), st));
  NB_REQUIRE(thresh_first > 0 && thresh_last > thresh_first, "bin thresholds must be positive and increasing");
  NbPhase ph(ctx, NB_PH_PK_BIN, st);
  return nb_pk_bin_spectrum(ctx, spec, N, 0, N, prec, k2_tab, sinc_tab, thresh, thresh_hi, n_edges, window_power,
                            clip_mode, clip_val, norm, closed_last, thresh_first, thresh_last, counts, sums, st);
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
