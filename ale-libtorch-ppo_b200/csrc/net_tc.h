// net_tc.h — tcgen05/TMEM tensor-core Nature-CNN path (net_tc.cu).
#pragma once
#include "net.h"

namespace alegpu {

struct TcState {
  int64_t hidden = 0, actions = 0, cap = 0, Ppad = 0;
  int elem_bytes = 2;        // operand type of every buffer below: 2 = bf16 (kind::f16), 4 = fp32 rounded to tf32 (kind::tf32)
  void *w_shadow = nullptr;  // operand-type copy of the flat parameter buffer (written by Adam)
  void *c1w_perm = nullptr;  // bf16 conv1 weights [32][256] in the k' = (tap, y&3, x&3, c) order of the space-to-depth obs16
  // bf16 NHWC copy of the observations conv1 reads: one minibatch (cap rows) and, lazily, a whole
  // train() batch so the conversion runs once per train() instead of once per minibatch step
  void *obs16_mb = nullptr, *obs16_all = nullptr;
  int64_t obs16_all_rows = 0;
  void *act1 = nullptr, *act2 = nullptr, *act3 = nullptr;    // bf16 NHWC activations
  void *dact1 = nullptr, *dact2 = nullptr, *dact3 = nullptr; // NHWC activation grads (operand type)
  void *dact3p = nullptr, *dact2p = nullptr; // zero-bordered [n,10,10,64] copies (shift-convolution data gradients)
  void *fc16 = nullptr, *dfc16 = nullptr;                    // bf16 copies of fc in/out grads
  void *wt = nullptr;        // bf16 re-laid-out weights for dgrad
  float *stage = nullptr;    // fp32 staging for get_activations
  // weight-gradient branch: runs on its own stream next to the data-gradient chain (fork/join
  // with events, so it also becomes a parallel branch of a captured CUDA graph); every wgrad
  // has its own split-partials workspace because they may be in flight together
  cudaStream_t side = nullptr;
  cudaEvent_t ev_fork = nullptr, ev_d3 = nullptr, ev_d2 = nullptr, ev_d1 = nullptr, ev_join = nullptr;
  cudaEvent_t ev_fcw = nullptr; // recorded on the side branch once the FC + heads gradients exist
  float *ws_heads = nullptr, *ws_fc = nullptr, *ws_c3 = nullptr, *ws_c2 = nullptr, *ws_c1 = nullptr;
  size_t ws_heads_n = 0, ws_fc_n = 0, ws_c3_n = 0, ws_c2_n = 0, ws_c1_n = 0;
  // TMA im2col tensor maps (device array of CUtensorMap, see net_tc.cu: TmaSlot); nullptr = not in use
  void *tmaps = nullptr;
  int obs_all_map = 0;       // which of the round-robin obs16_all descriptor slots is current
  bool use_tma = true;       // ALEGPU_NO_TMA=1: gather conv operands with cp.async instead (A/B)
  bool use_tma_weights = true; // conv weights (conv2 forward, conv3 / conv2 data gradients) through TMA (ALEGPU_NO_TMA_WEIGHTS=1: cp.async)
  bool fc_tma = false;       // FC forward / data-gradient operands through tiled TMA (ALEGPU_NO_TMA_FC=1: cp.async)
  bool use_tma_wgrad = true; // ALEGPU_NO_TMA_WGRAD=1: the weight gradients' plain-matrix operand via cp.async (A/B)
  bool use_shift_dgrad = true; // conv3 / conv2 data gradients as shift convolutions (default: bf16 only; ALEGPU_SHIFT_DGRAD=0|1)
  bool use_shift = true;     // ALEGPU_NO_SHIFT=1: conv1 / conv3 forward through the im2col engine (A/B)
  bool shift_base_offset = false; // experiment switch, see net_tc_init: the field must stay 0
  bool serial = false;       // ALEGPU_TC_SERIAL=1: keep everything on one stream (profiling)
  bool conv1_u8 = false;     // ALEGPU_CONV1_U8=1: conv1 forward reads the u8 observations (A/B)
  bool conv1_wgrad_u8 = false; // ALEGPU_CONV1_WGRAD_U8=1: conv1 wgrad in im2col form over the u8 observations (A/B)
};

// dfc_f32: the learner's fp32 dfc buffer [cap][hidden] (the tf32 path's FC data-gradient operand)
int net_tc_init(TcState &T, int64_t hidden, int64_t actions, int64_t cap, int64_t Ppad,
                int elem_bytes, float *dfc_f32, cudaStream_t s);
void net_tc_destroy(TcState &T);
int net_tc_refresh_weights(TcState &T, const float *params, int64_t Ppad, cudaStream_t s);
// obs16 = the rows' bf16 NHWC copy when the caller already made one (net_tc_prepare_obs); nullptr:
// the forward converts `obs` into T.obs16_mb itself and the backward of the same rows reuses it
int net_tc_forward(const NetParams &P, TcState &T, const uint8_t *obs, int64_t n, NetActs &A,
                   cudaStream_t s, const void *obs16 = nullptr);
// reserve: (re)allocates T.obs16_all for n rows (skipped above 131 072 rows; never call it inside a
// stream capture).  prepare: converts rows [n,4,84,84] u8 into it; *obs16 = nullptr if not reserved
int net_tc_reserve_obs(TcState &T, int64_t n);
int net_tc_prepare_obs(TcState &T, const uint8_t *obs, int64_t n, const void **obs16, cudaStream_t s);
// includes the heads' weight gradients (heads_wgrad) — the caller must not launch them again
int net_tc_backward(const NetParams &P, const NetParams &G, TcState &T, const uint8_t *obs,
                    int64_t n, NetActs &A, cudaStream_t s, const void *obs16 = nullptr);
int net_tc_fetch_activation(TcState &T, int which, int64_t n, float *host, cudaStream_t s);

} // namespace alegpu
