/* TEST INFRASTRUCTURE ONLY — plain-C float64 restatement of the loops that are too slow in
 * numpy at bench sizes.  Built by oracle/Makefile into oracle/liboracle_c.so and loaded by
 * oracle/cport.py; used by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs only.  The product package never links or loads it.
 *
 * Each function cites what it follows (reference paths relative to the reference root):
 *   oc_cic_deposit   FraP_kTilDelta.py:93-123 (= Power_compare.py:100-132): serial loop over
 *                    particles, i = int(p/cell) % N, d = (p % cell)/cell, 8 weighted adds in
 *                    (di,dj,dk) order — bit-identical to the reference's numba loop
 *   oc_cic_gather    adjoint of the deposit, same index/weight expressions
 *   oc_pp_accel      oracle/force.py:pp_accel (our spec; CONCEPT PP has no source here)
 *   oc_p3m_short     oracle/force.py:p3m_short_accel with a cell list instead of all pairs
 *   oc_shell_bin     Power_compare.py:179-194 / :453-461: k = sqrt((kx²+ky²)+kz²), bin by
 *                    binary search over the edges (histogram or digitize semantics)
 * Compile with -ffp-contract=off so no FMA changes a cell index, weight or |k|.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

static inline double pymod(double a, double b) {
  double r = fmod(a, b);
  if (r != 0.0) {
    if ((b < 0.0) != (r < 0.0)) r += b;
  } else {
    r = copysign(0.0, b);
  }
  return r;
}

static inline int64_t cell_index(double p, double cell, int64_t N) {
  int64_t t = (int64_t)(p / cell); /* C truncates toward zero like Python's int() */
  int64_t m = t % N;
  if (m < 0) m += N;
  return m;
}

int oc_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

/* bench.py's CPU arm sets the thread count itself: under torch.distributed.run every rank inherits
 * OMP_NUM_THREADS=1, which would time a single-threaded baseline. */
void oc_set_num_threads(int n) {
#ifdef _OPENMP
  if (n >= 1) omp_set_num_threads(n);
#else
  (void)n;
#endif
}

void oc_cic_deposit(const double* pos, int64_t np, double cell, int64_t N, double shift, double* grid) {
  for (int64_t p = 0; p < np; ++p) {
    const double x = pos[3 * p] + shift, y = pos[3 * p + 1] + shift, z = pos[3 * p + 2] + shift;
    const int64_t i = cell_index(x, cell, N), j = cell_index(y, cell, N), k = cell_index(z, cell, N);
    const double dx = pymod(x, cell) / cell, dy = pymod(y, cell) / cell, dz = pymod(z, cell) / cell;
    for (int di = 0; di < 2; ++di)
      for (int dj = 0; dj < 2; ++dj)
        for (int dk = 0; dk < 2; ++dk) {
          const double w = ((di ? dx : 1 - dx) * (dj ? dy : 1 - dy)) * (dk ? dz : 1 - dz);
          grid[(((i + di) % N) * N + (j + dj) % N) * N + (k + dk) % N] += w;
        }
  }
}

void oc_cic_gather(const double* field, int64_t N, const double* pos, int64_t np, double cell, double shift,
                   double* out) {
#pragma omp parallel for schedule(static)
  for (int64_t p = 0; p < np; ++p) {
    const double x = pos[3 * p] + shift, y = pos[3 * p + 1] + shift, z = pos[3 * p + 2] + shift;
    const int64_t i = cell_index(x, cell, N), j = cell_index(y, cell, N), k = cell_index(z, cell, N);
    const double dx = pymod(x, cell) / cell, dy = pymod(y, cell) / cell, dz = pymod(z, cell) / cell;
    double v = 0.0;
    for (int di = 0; di < 2; ++di)
      for (int dj = 0; dj < 2; ++dj)
        for (int dk = 0; dk < 2; ++dk) {
          const double w = ((di ? dx : 1 - dx) * (dj ? dy : 1 - dy)) * (dk ? dz : 1 - dz);
          v += w * field[(((i + di) % N) * N + (j + dj) % N) * N + (k + dk) % N];
        }
    out[p] = v;
  }
}

/* a_i = G m sum_j d/(|d|²+eps²)^{3/2}, d = x_j - x_i; i in [i0,i1) */
void oc_pp_accel(const double* pos, int64_t np, int64_t i0, int64_t i1, double G, double mass, double eps,
                 double box, double* acc) {
  const double eps2 = eps * eps;
#pragma omp parallel for schedule(dynamic, 16)
  for (int64_t i = i0; i < i1; ++i) {
    double ax = 0, ay = 0, az = 0;
    const double xi = pos[3 * i], yi = pos[3 * i + 1], zi = pos[3 * i + 2];
    for (int64_t j = 0; j < np; ++j) {
      double dx = pos[3 * j] - xi, dy = pos[3 * j + 1] - yi, dz = pos[3 * j + 2] - zi;
      if (box > 0) {
        dx -= box * nearbyint(dx / box);
        dy -= box * nearbyint(dy / box);
        dz -= box * nearbyint(dz / box);
      }
      const double r2 = dx * dx + dy * dy + dz * dz + eps2;
      if (r2 > 0) {
        const double w = 1.0 / (r2 * sqrt(r2));
        ax += w * dx; ay += w * dy; az += w * dz;
      }
    }
    acc[3 * (i - i0)] = G * mass * ax;
    acc[3 * (i - i0) + 1] = G * mass * ay;
    acc[3 * (i - i0) + 2] = G * mass * az;
  }
}

/* Short-range P3M force with a cell list (cells of side L/ncs >= r_cut), minimum image. */
int oc_p3m_short(const double* pos, int64_t np, double L, double r_s, double r_cut, double G, double mass,
                 double eps, double* acc, int64_t* n_inrange) {
  int64_t ncs = (int64_t)floor(L / r_cut);
  if (ncs < 3) return -1;
  const double cs = L / ncs;
  const int64_t nc = ncs * ncs * ncs;
  int64_t* start = (int64_t*)calloc(nc + 1, sizeof(int64_t));
  int64_t* cell = (int64_t*)malloc(sizeof(int64_t) * (np > 0 ? np : 1));
  int64_t* order = (int64_t*)malloc(sizeof(int64_t) * (np > 0 ? np : 1));
  if (!start || !cell || !order) return -2;
  for (int64_t p = 0; p < np; ++p) {
    int64_t c[3];
    for (int a = 0; a < 3; ++a) {
      double x = pos[3 * p + a];
      if (x < 0 || x >= L) x = pymod(x, L);
      c[a] = (int64_t)(x / cs);
      if (c[a] >= ncs) c[a] = ncs - 1;
    }
    cell[p] = (c[0] * ncs + c[1]) * ncs + c[2];
    start[cell[p] + 1]++;
  }
  for (int64_t c = 0; c < nc; ++c) start[c + 1] += start[c];
  int64_t* fill = (int64_t*)calloc(nc, sizeof(int64_t));
  for (int64_t p = 0; p < np; ++p) order[start[cell[p]] + fill[cell[p]]++] = p;
  free(fill);
  const double rc2 = r_cut * r_cut, eps2 = eps * eps, two_over_sqrtpi = 2.0 / sqrt(M_PI);
  int64_t total = 0;
#pragma omp parallel for schedule(dynamic, 64) reduction(+ : total)
  for (int64_t i = 0; i < np; ++i) {
    const int64_t c = cell[i];
    const int64_t cz = c % ncs, cy = (c / ncs) % ncs, cx = c / (ncs * ncs);
    double ax = 0, ay = 0, az = 0;
    for (int ox = -1; ox <= 1; ++ox)
      for (int oy = -1; oy <= 1; ++oy)
        for (int oz = -1; oz <= 1; ++oz) {
          const int64_t n = (((cx + ox + ncs) % ncs) * ncs + (cy + oy + ncs) % ncs) * ncs + (cz + oz + ncs) % ncs;
          for (int64_t s = start[n]; s < start[n + 1]; ++s) {
            const int64_t j = order[s];
            double dx = pos[3 * j] - pos[3 * i], dy = pos[3 * j + 1] - pos[3 * i + 1], dz = pos[3 * j + 2] - pos[3 * i + 2];
            dx -= L * nearbyint(dx / L);
            dy -= L * nearbyint(dy / L);
            dz -= L * nearbyint(dz / L);
            const double r2 = dx * dx + dy * dy + dz * dz;
            if (r2 < rc2 && r2 > 0) {
              const double r = sqrt(r2), u = r / (2 * r_s);
              const double f = erfc(u) + two_over_sqrtpi * u * exp(-u * u);
              const double rs2 = r2 + eps2;
              const double w = f / (rs2 * sqrt(rs2));
              ax += w * dx; ay += w * dy; az += w * dz;
              total++;
            }
          }
        }
    acc[3 * i] = G * mass * ax; acc[3 * i + 1] = G * mass * ay; acc[3 * i + 2] = G * mass * az;
  }
  if (n_inrange) *n_inrange = total;
  free(start); free(cell); free(order);
  return 0;
}

/* counts / sums of P over log shells; P given on the full N^3 grid */
void oc_shell_bin(const double* P, int64_t N, const double* kvals, const double* edges, int64_t n_edges,
                  int closed_last, int64_t* counts, double* sums) {
  const int64_t nb = n_edges - 1;
  memset(counts, 0, sizeof(int64_t) * nb);
  memset(sums, 0, sizeof(double) * nb);
  for (int64_t i = 0; i < N; ++i)
    for (int64_t j = 0; j < N; ++j)
      for (int64_t k = 0; k < N; ++k) {
        const double kk = sqrt((kvals[i] * kvals[i] + kvals[j] * kvals[j]) + kvals[k] * kvals[k]);
        if (!(kk > 0)) continue;
        /* idx = (#edges <= kk) - 1 */
        int64_t lo = 0, hi = n_edges;
        while (lo < hi) {
          const int64_t mid = (lo + hi) / 2;
          if (edges[mid] <= kk) lo = mid + 1; else hi = mid;
        }
        int64_t idx = lo - 1;
        if (closed_last && kk == edges[n_edges - 1]) idx = nb - 1;
        if (idx >= 0 && idx < nb) {
          counts[idx]++;
          sums[idx] += P[(i * N + j) * N + k];
        }
      }
}
