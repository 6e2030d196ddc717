// Context: device selection, grow-only scratch, cuFFT plan cache, error string.
#include <stdarg.h>

#include "common.cuh"

static thread_local char g_err[1024] = "";

void nb_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

extern "C" const char* nb_last_error(void) { return g_err; }
extern "C" int nb_abi_version(void) { return NB_ABI_VERSION; }

extern "C" int nb_init(int device, nb_ctx** out) {
  NB_REQUIRE(out != nullptr, "out is NULL");
  int ndev = 0;
  NB_CHECK_CUDA(cudaGetDeviceCount(&ndev));
  NB_REQUIRE(device >= 0 && device < ndev, "no such CUDA device");
  NB_CHECK_CUDA(cudaSetDevice(device));
  nb_ctx* c = new nb_ctx();
  c->device = device;
  int sms = 0;
  if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device) == cudaSuccess && sms > 0)
    c->num_sms = sms;
  *out = c;
  return NB_OK;
}

extern "C" int nb_release_scratch(nb_ctx* ctx) {
  NB_REQUIRE(ctx != nullptr, "ctx is NULL");
  NB_CHECK_CUDA(cudaSetDevice(ctx->device));
  for (auto& kv : ctx->plans) cufftDestroy(kv.second);
  ctx->plans.clear();
  for (int s = 0; s < NB_SCR_NSLOTS; ++s) {
    if (ctx->scratch[s]) cudaFree(ctx->scratch[s]);
    ctx->scratch[s] = nullptr;
    ctx->scratch_bytes[s] = 0;
  }
  return NB_OK;
}

extern "C" void nb_destroy(nb_ctx* ctx) {
  if (!ctx) return;
  nb_release_scratch(ctx);
  delete ctx;
}

extern "C" int64_t nb_launch_count(nb_ctx* ctx) { return ctx ? ctx->launches : -1; }

int nb_scratch(nb_ctx* ctx, int slot, size_t bytes, void** out) {
  NB_REQUIRE(slot >= 0 && slot < NB_SCR_NSLOTS, "bad scratch slot");
  if (ctx->scratch_bytes[slot] < bytes) {
    if (ctx->scratch[slot]) {
      // a previous user of the slot may still be running on some stream
      NB_CHECK_CUDA(cudaDeviceSynchronize());
      NB_CHECK_CUDA(cudaFree(ctx->scratch[slot]));
      ctx->scratch[slot] = nullptr;
      ctx->scratch_bytes[slot] = 0;
    }
    size_t want = (bytes + 255) & ~(size_t)255;
    cudaError_t e = cudaMalloc(&ctx->scratch[slot], want);
    if (e != cudaSuccess) {
      nb_set_error("scratch slot %d: cudaMalloc(%zu) failed: %s", slot, want, cudaGetErrorString(e));
      return NB_ENOMEM;
    }
    ctx->scratch_bytes[slot] = want;
  }
  *out = ctx->scratch[slot];
  return NB_OK;
}

int nb_plan3d(nb_ctx* ctx, int N, cufftType type, cudaStream_t stream, cufftHandle* out) {
  auto key = std::make_tuple((int)type, N, N, N, 1, 0);
  auto it = ctx->plans.find(key);
  if (it == ctx->plans.end()) {
    cufftHandle h;
    NB_CHECK_CUFFT(cufftPlan3d(&h, N, N, N, type));
    it = ctx->plans.emplace(key, h).first;
  }
  NB_CHECK_CUFFT(cufftSetStream(it->second, stream));
  *out = it->second;
  return NB_OK;
}

int nb_plan_many(nb_ctx* ctx, int rank, const int* n, int batch, cufftType type, int layout_tag,
                 const int* inembed, int istride, int idist, const int* onembed, int ostride,
                 int odist, cudaStream_t stream, cufftHandle* out) {
  auto key = std::make_tuple((int)type, n[0], rank > 1 ? n[1] : 0, rank > 2 ? n[2] : 0, batch, layout_tag);
  auto it = ctx->plans.find(key);
  if (it == ctx->plans.end()) {
    cufftHandle h;
    NB_CHECK_CUFFT(cufftPlanMany(&h, rank, const_cast<int*>(n), const_cast<int*>(inembed), istride, idist,
                                 const_cast<int*>(onembed), ostride, odist, type, batch));
    it = ctx->plans.emplace(key, h).first;
  }
  NB_CHECK_CUFFT(cufftSetStream(it->second, stream));
  *out = it->second;
  return NB_OK;
}

int nb_exec_c2r(nb_ctx* ctx, cufftHandle plan, int prec, void* spec, void* dst, size_t real_bytes,
                cudaStream_t st) {
  void* out = dst;
  const bool misaligned = ((uintptr_t)dst & 15u) != 0;
  if (misaligned) {
    int rc = nb_scratch(ctx, NB_SCR_REAL1, real_bytes, &out);
    if (rc) return rc;
  }
  if (prec == NB_F64)
    NB_CHECK_CUFFT(cufftExecZ2D(plan, (cufftDoubleComplex*)spec, (cufftDoubleReal*)out));
  else
    NB_CHECK_CUFFT(cufftExecC2R(plan, (cufftComplex*)spec, (cufftReal*)out));
  if (misaligned) NB_CHECK_CUDA(cudaMemcpyAsync(dst, out, real_bytes, cudaMemcpyDeviceToDevice, st));
  return NB_OK;
}

extern "C" int nb_profile_enable(nb_ctx* ctx, int on) {
  NB_REQUIRE(ctx != nullptr, "ctx is NULL");
  ctx->prof_on = on != 0;
  return NB_OK;
}

// Sums the recorded phase durations: ms_sum_host[phase] (milliseconds) and
// count_host[phase] (number of recorded intervals) for phase < nphases; clears the records.
extern "C" int nb_profile_read(nb_ctx* ctx, double* ms_sum_host, int64_t* count_host, int nphases) {
  NB_REQUIRE(ctx && ms_sum_host && count_host && nphases > 0, "bad arguments");
  NB_CHECK_CUDA(cudaSetDevice(ctx->device));
  for (int p = 0; p < nphases; ++p) { ms_sum_host[p] = 0.0; count_host[p] = 0; }
  for (auto& t : ctx->prof_events) {
    const int ph = std::get<0>(t);
    cudaEvent_t e0 = std::get<1>(t), e1 = std::get<2>(t);
    float ms = 0.f;
    if (cudaEventSynchronize(e1) == cudaSuccess && cudaEventElapsedTime(&ms, e0, e1) == cudaSuccess &&
        ph >= 0 && ph < nphases) {
      ms_sum_host[ph] += ms;
      count_host[ph] += 1;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
  }
  ctx->prof_events.clear();
  return NB_OK;
}
