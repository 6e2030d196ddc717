// Developer microbenchmark: where do the 6.6 ms per 1024^3 C2R go?  Times cuFFT R2C / C2R
// out-of-place and in-place, and the plane-wise decomposition (1-D C2C along x + batched 2-D C2R).
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include <cufft.h>
#define CK(x) do { auto e = (x); if (e) { printf("error %d at line %d\n", (int)e, __LINE__); exit(1); } } while (0)

static float time_it(cudaStream_t st, int reps, const char* name, auto fn) {
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  fn(); fn();
  CK(cudaStreamSynchronize(st));
  cudaEventRecord(a, st);
  for (int i = 0; i < reps; ++i) fn();
  cudaEventRecord(b, st);
  CK(cudaStreamSynchronize(st));
  float ms; cudaEventElapsedTime(&ms, a, b);
  printf("%-44s %8.3f ms\n", name, ms / reps);
  return ms / reps;
}

int main(int argc, char** argv) {
  const int N = argc > 1 ? atoi(argv[1]) : 1024;
  const size_t nzh = N / 2 + 1;
  const size_t nreal = (size_t)N * N * N, ncplx = (size_t)N * N * nzh;
  float* real; cufftComplex* spec; cufftComplex* spec2;
  CK(cudaMalloc(&real, sizeof(float) * nreal));
  CK(cudaMalloc(&spec, sizeof(cufftComplex) * ncplx));
  CK(cudaMalloc(&spec2, sizeof(cufftComplex) * ncplx));
  CK(cudaMemset(real, 0, sizeof(float) * nreal));
  CK(cudaMemset(spec, 0, sizeof(cufftComplex) * ncplx));
  cudaStream_t st; cudaStreamCreate(&st);
  cufftHandle r2c, c2r, c2c1, c2r2;
  size_t ws;
  CK(cufftPlan3d(&r2c, N, N, N, CUFFT_R2C)); CK(cufftSetStream(r2c, st)); cufftGetSize(r2c, &ws); printf("R2C work area %.1f MB\n", ws / 1e6);
  CK(cufftPlan3d(&c2r, N, N, N, CUFFT_C2R)); CK(cufftSetStream(c2r, st)); cufftGetSize(c2r, &ws); printf("C2R work area %.1f MB\n", ws / 1e6);
  time_it(st, 5, "R2C 3-D out-of-place", [&] { cufftExecR2C(r2c, real, spec); });
  time_it(st, 5, "C2R 3-D out-of-place", [&] { cufftExecC2R(c2r, spec, real); });
  time_it(st, 5, "C2R 3-D in-place (padded)", [&] { cufftExecC2R(c2r, spec, (float*)spec); });
  time_it(st, 5, "R2C 3-D in-place (padded)", [&] { cufftExecR2C(r2c, (float*)spec, spec); });
  // decomposition: 1-D C2C along x (stride N*nzh, batch N*nzh), then N planes of 2-D C2R (y,z)
  int n1[1] = {N};
  int emb[1] = {N};
  CK(cufftPlanMany(&c2c1, 1, n1, emb, (int)(N * nzh), 1, emb, (int)(N * nzh), 1, CUFFT_C2C, (int)(N * nzh)));
  CK(cufftSetStream(c2c1, st));
  int n2[2] = {N, N};
  int cemb[2] = {N, (int)nzh}, remb[2] = {N, N};
  CK(cufftPlanMany(&c2r2, 2, n2, cemb, 1, (int)(N * nzh), remb, 1, N * N, CUFFT_C2R, N));
  CK(cufftSetStream(c2r2, st));
  time_it(st, 5, "1-D C2C along x (strided, in place)", [&] { cufftExecC2C(c2c1, spec, spec, CUFFT_INVERSE); });
  time_it(st, 5, "N x 2-D C2R (y,z) out-of-place", [&] { cufftExecC2R(c2r2, spec, real); });
  time_it(st, 5, "C2C x + 2-D C2R planes (sum)", [&] { cufftExecC2C(c2c1, spec, spec, CUFFT_INVERSE); cufftExecC2R(c2r2, spec, real); });
  // copy bandwidth reference
  time_it(st, 5, "cudaMemcpyAsync D2D of the real grid", [&] { cudaMemcpyAsync(spec2, real, sizeof(float) * nreal, cudaMemcpyDeviceToDevice, st); });
  return 0;
}
