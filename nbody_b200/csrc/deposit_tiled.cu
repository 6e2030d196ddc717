// K1 tiled variant — placeholder until the first GPU measurement of the direct kernel.
#include "common.cuh"

int nb_deposit_tiled_try(nb_ctx* ctx, const double* pos, int64_t np, double cell, int N, float* grid,
                         double shift, cudaStream_t st, int* used) {
  (void)ctx; (void)pos; (void)np; (void)cell; (void)N; (void)grid; (void)shift; (void)st;
  *used = 0;
  return NB_OK;
}
