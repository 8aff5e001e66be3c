// minibatch.h — shuffled minibatch gather and advantage statistics (minibatch.cu).
#pragma once
#include "common.cuh"

namespace alegpu {

// staging for one gathered minibatch (same layouts as ai::ppo::train::Batch, train.h:34-57)
struct StagedBatch {
  uint8_t *obs = nullptr;     // [cap,4,84,84]
  int64_t *actions = nullptr; // [cap]
  float *logp = nullptr;      // [cap,A]
  float *adv = nullptr, *ret = nullptr;
  float *values = nullptr;    // [cap] old values (clipped value loss only)
  uint8_t *masks = nullptr;
};

// out row i = source row idx[i] (i < n); an index outside [0, n_rows_src) sets *error_flag
int launch_gather_rows(const int64_t *idx, int64_t n, int64_t n_rows_src, int A,
                       const uint8_t *obs, const int64_t *actions, const float *logp,
                       const float *adv, const float *ret, const float *values /* may be null */,
                       const uint8_t *masks, const StagedBatch &out, int *error_flag, cudaStream_t s);
// *error_flag = 1 if any idx[i] (i < n) lies outside [0, n_rows_src)
int launch_check_indices(const int64_t *idx, int64_t n, int64_t n_rows_src, int *error_flag,
                         cudaStream_t s);
// stats[0..2] = (count, sum, sum of squares) of adv over rows with masks != 0, in double
int launch_adv_stats(const float *adv, const uint8_t *masks, int64_t n, double *stats,
                     cudaStream_t s);

} // namespace alegpu
