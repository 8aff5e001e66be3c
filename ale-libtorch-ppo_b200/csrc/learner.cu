// learner.cu — the policy + optimiser + trainer behind the C ABI (include/alegpu.h).
//
// Host-side C++ mirror of ai::ppo::train::{train, mini_batch_update} (src/ai/ppo/train.h:114-157)
// over the hand-written kernels: per minibatch
//   forward (conv1..fc) -> heads+loss+dfc (one kernel) -> heads wgrad -> backward
//   -> [NCCL sum of the flat gradient over ranks] -> ||g||^2 partials -> clip+Adam -> metrics tail
// all enqueued on one stream with no host synchronisation; lr, the Adam step counter and the
// clip coefficient live in device memory, so the whole train() call can be captured once into
// a CUDA graph and replayed (the reference's capture_train_cuda_graph, train.h:163-195,
// freezes lr into the graph; this one does not).
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <vector>

#include "minibatch.h"
#include "net.h"
#include "net_tc.h"
#include "optim.h"

#ifdef ALEGPU_WITH_NCCL
#include <dlfcn.h>
#include <nccl.h>
#endif

namespace alegpu {

#ifdef ALEGPU_WITH_NCCL
// NCCL is bound lazily with dlopen so that (a) a host that already loaded a libnccl.so.2
// (e.g. the one bundled with torch, for torch.distributed plumbing) shares that copy instead of
// clashing with the system one, and (b) single-GPU users never load NCCL at all.
struct NcclApi {
  ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t,
                            cudaStream_t) = nullptr;
  ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*CommSplit)(ncclComm_t, int, int, ncclComm_t *, ncclConfig_t *) = nullptr;
  const char *(*GetErrorString)(ncclResult_t) = nullptr;
  bool ok = false;
};
static NcclApi &nccl() {
  static NcclApi api;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (h) {
      api.GetUniqueId = (decltype(api.GetUniqueId))dlsym(h, "ncclGetUniqueId");
      api.CommInitRank = (decltype(api.CommInitRank))dlsym(h, "ncclCommInitRank");
      api.AllReduce = (decltype(api.AllReduce))dlsym(h, "ncclAllReduce");
      api.AllGather = (decltype(api.AllGather))dlsym(h, "ncclAllGather");
      api.CommDestroy = (decltype(api.CommDestroy))dlsym(h, "ncclCommDestroy");
      api.CommSplit = (decltype(api.CommSplit))dlsym(h, "ncclCommSplit");
      api.GetErrorString = (decltype(api.GetErrorString))dlsym(h, "ncclGetErrorString");
      api.ok = api.GetUniqueId && api.CommInitRank && api.AllReduce && api.CommDestroy &&
               api.GetErrorString;
    }
  }
  return api;
}
#endif

thread_local int64_t g_launches = 0;
thread_local bool g_trace_on = false;
namespace {
struct TraceRec {
  const char *file;
  int line;
  cudaEvent_t ev;
};
thread_local std::vector<TraceRec> g_trace;
thread_local cudaStream_t g_trace_stream = nullptr;
} // namespace
void trace_launch(const char *file, int line) {
  cudaEvent_t e;
  if (cudaEventCreate(&e) != cudaSuccess)
    return;
  cudaEventRecord(e, g_trace_stream);
  g_trace.push_back({file, line, e});
}
namespace {
thread_local std::string g_error;
}
void set_error(const std::string &msg) { g_error = msg; }
int fail(int status, const std::string &msg) {
  g_error = msg;
  return status;
}

} // namespace alegpu

using namespace alegpu;

struct alegpu_learner {
  alegpu_learner_config cfg{};
  int64_t P = 0, Ppad = 0, cap = 0;
  cudaStream_t stream = nullptr;
  float *flat = nullptr; // params | grads | m | v, each Ppad floats
  float *params = nullptr, *grads = nullptr, *m = nullptr, *v = nullptr;
  NetParams Pv{}, Gv{};
  NetActs acts{};
  TcState tc{};
  float *ws = nullptr;
  size_t ws_floats = 0;
  float *norm_partials = nullptr, *loss_partials = nullptr, *counts = nullptr, *scalars = nullptr;
  int64_t counts_cap = 0;
  // opt-in minibatch preparation (alegpu_train_options): gather staging, advantage statistics
  StagedBatch staged;
  double *adv_stats = nullptr;
  int *gather_error = nullptr;
  unsigned int *loss_done = nullptr; // counter of the heads kernel's last-block loss reduction
  // device-side input errors (action index outside [0,A): the reference's gather would throw) land in
  // a MAPPED pinned int the host can read without synchronising; checked at the next learner call
  int *err_host = nullptr, *err_dev = nullptr;
  OptState *opt = nullptr;
  int64_t host_step = 0;
  double lr = 0.0;
  int64_t last_launches = 0;
  // cached CUDA graph of one train() call shape
  cudaGraphExec_t graph_exec = nullptr;
  std::vector<uint64_t> graph_key;
  int64_t graph_launches = 0;
#ifdef ALEGPU_WITH_NCCL
  ncclComm_t comm = nullptr;
  // second communicator + stream: the FC/heads gradient bucket is exchanged while the conv
  // data-gradient chain is still running (tensor-core path)
  ncclComm_t comm2 = nullptr;
  cudaStream_t comm_stream = nullptr;
  cudaEvent_t ev_comm = nullptr;
#endif
  // fused NVLink gradient exchange (optim.cu: adam_fused_allreduce_kernel): peers' gradient /
  // exchange / signal buffers mapped through CUDA IPC; false = the ncclAllReduce path
  bool fused_ar = false;
  ArParams ar{};
  void *xchg = nullptr;            // [ArSignals | reduced slice]
  std::vector<void *> ipc_opened;  // peer mappings to close
  std::string comm_mode = "single";
  cudaEvent_t ev_bridge_in = nullptr, ev_bridge_out = nullptr; // alegpu_learner_wait_stream / _stream_wait_learner
  std::vector<float> host_tmp;
};

namespace {

// reference tensor layout <-> internal layout (see net_f32.cu header) ------------------------
void permute_params(const float *src, float *dst, int64_t H, int64_t A, bool to_internal) {
  const int64_t P = net_param_count(H, A);
  std::memcpy(dst, src, sizeof(float) * P);
  int64_t off = 32 * 256 + 32;
  for (int co = 0; co < 64; ++co) // c2w: OIHW [64][32][4][4] <-> [64][ky][kx][ci]
    for (int ci = 0; ci < 32; ++ci)
      for (int ky = 0; ky < 4; ++ky)
        for (int kx = 0; kx < 4; ++kx) {
          const int64_t r = off + ((co * 32 + ci) * 4 + ky) * 4 + kx;
          const int64_t i = off + co * 512 + (ky * 4 + kx) * 32 + ci;
          if (to_internal) dst[i] = src[r]; else dst[r] = src[i];
        }
  off += 64 * 512 + 64;
  for (int co = 0; co < 64; ++co) // c3w: [64][64][3][3] <-> [64][ky][kx][ci]
    for (int ci = 0; ci < 64; ++ci)
      for (int ky = 0; ky < 3; ++ky)
        for (int kx = 0; kx < 3; ++kx) {
          const int64_t r = off + ((co * 64 + ci) * 3 + ky) * 3 + kx;
          const int64_t i = off + co * 576 + (ky * 3 + kx) * 64 + ci;
          if (to_internal) dst[i] = src[r]; else dst[r] = src[i];
        }
  off += 64 * 576 + 64;
  for (int64_t j = 0; j < H; ++j) // fcw: [H][c*49 + p] <-> [H][p*64 + c]
    for (int c = 0; c < 64; ++c)
      for (int p = 0; p < 49; ++p) {
        const int64_t r = off + j * 3136 + c * 49 + p;
        const int64_t i = off + j * 3136 + p * 64 + c;
        if (to_internal) dst[i] = src[r]; else dst[r] = src[i];
      }
}

int upload_flat(alegpu_learner *L, float *dev, const float *host_ref) {
  L->host_tmp.assign((size_t)L->Ppad, 0.0f);
  permute_params(host_ref, L->host_tmp.data(), L->cfg.hidden_size, L->cfg.action_size, true);
  ALEGPU_CUDA_CHECK(cudaMemcpyAsync(dev, L->host_tmp.data(), sizeof(float) * L->Ppad,
                                    cudaMemcpyHostToDevice, L->stream));
  ALEGPU_CUDA_CHECK(cudaStreamSynchronize(L->stream));
  return ALEGPU_OK;
}

int download_flat(alegpu_learner *L, const float *dev, float *host_ref) {
  L->host_tmp.assign((size_t)L->Ppad, 0.0f);
  ALEGPU_CUDA_CHECK(cudaMemcpyAsync(L->host_tmp.data(), dev, sizeof(float) * L->Ppad,
                                    cudaMemcpyDeviceToHost, L->stream));
  ALEGPU_CUDA_CHECK(cudaStreamSynchronize(L->stream));
  permute_params(L->host_tmp.data(), host_ref, L->cfg.hidden_size, L->cfg.action_size, false);
  return ALEGPU_OK;
}

int sync_opt_state(alegpu_learner *L) {
  OptState h{L->lr, (long long)L->host_step};
  ALEGPU_CUDA_CHECK(cudaMemcpyAsync(L->opt, &h, sizeof(h), cudaMemcpyHostToDevice, L->stream));
  ALEGPU_CUDA_CHECK(cudaStreamSynchronize(L->stream));
  return ALEGPU_OK;
}

bool use_tc(const alegpu_learner *L) {
  return L->cfg.math == ALEGPU_MATH_BF16 || L->cfg.math == ALEGPU_MATH_TF32;
}

// an input error a kernel of an EARLIER call reported through the mapped flag (no synchronisation)
int learner_input_error(alegpu_learner *L) {
  if (L->err_host && *(volatile int *)L->err_host) {
    const int code = *(volatile int *)L->err_host;
    *(volatile int *)L->err_host = 0;
    if (code == 2)
      return fail(ALEGPU_RUNTIME_ERROR, "train: a peer rank did not reach the gradient exchange in time");
    return fail(ALEGPU_OUT_OF_RANGE, "train: an action index was outside [0, action_size)");
  }
  return ALEGPU_OK;
}

int forward(alegpu_learner *L, const uint8_t *obs, int64_t n, const void *obs16 = nullptr) {
  if (use_tc(L))
    return net_tc_forward(L->Pv, L->tc, obs, n, L->acts, L->stream, obs16);
  return net_f32_forward(L->Pv, obs, n, L->acts, L->stream);
}

int backward(alegpu_learner *L, const uint8_t *obs, int64_t n, const void *obs16 = nullptr) {
  if (use_tc(L)) // launches the heads' weight gradients itself, on its side branch
    return net_tc_backward(L->Pv, L->Gv, L->tc, obs, n, L->acts, L->stream, obs16);
  if (int rc = heads_wgrad(L->Gv, L->acts, n, L->ws, L->ws_floats, L->stream))
    return rc;
  return net_f32_backward(L->Pv, L->Gv, obs, n, L->acts, L->ws, L->ws_floats, L->stream);
}

int allreduce_grads(alegpu_learner *L) {
#ifdef ALEGPU_WITH_NCCL
  if (L->cfg.world_size > 1) {
    float *g = L->grads;
    size_t count = (size_t)L->Ppad;
    const size_t head = (size_t)(L->Gv.fcw - L->Gv.c1w); // conv tensors come first in the flat buffer
    if (use_tc(L) && L->comm2 && !L->tc.serial) {
      // bucket 1 (fcw, fcb, heads): on the communication stream as soon as the side branch has them
      ALEGPU_CUDA_CHECK(cudaStreamWaitEvent(L->comm_stream, L->tc.ev_fcw, 0));
      ncclResult_t r = nccl().AllReduce(g + head, g + head, count - head, ncclFloat, ncclSum,
                                        L->comm2, L->comm_stream);
      if (r != ncclSuccess)
        return fail(ALEGPU_CUDA_ERROR, std::string("ncclAllReduce: ") + nccl().GetErrorString(r));
      ALEGPU_COUNT_LAUNCH();
      ALEGPU_CUDA_CHECK(cudaEventRecord(L->ev_comm, L->comm_stream));
      count = head; // bucket 2 (conv tensors, 312 KB) follows the backward on the main stream
    }
    ncclResult_t r = nccl().AllReduce(g, g, count, ncclFloat, ncclSum, L->comm, L->stream);
    if (r != ncclSuccess)
      return fail(ALEGPU_CUDA_ERROR, std::string("ncclAllReduce: ") + nccl().GetErrorString(r));
    ALEGPU_COUNT_LAUNCH();
    if (count != (size_t)L->Ppad)
      ALEGPU_CUDA_CHECK(cudaStreamWaitEvent(L->stream, L->ev_comm, 0));
  }
#endif
  return ALEGPU_OK;
}

int clip_and_step(alegpu_learner *L, float max_norm, float *total_norm_out,
                  const float *count = nullptr, float *m_loss = nullptr, int64_t mb = 0,
                  bool exchange = false) {
  static const bool no_fused = getenv("ALEGPU_NO_FUSED_ADAM") != nullptr; // A/B: the two-kernel form
  if (exchange && L->fused_ar) { // gradient exchange over NVLink inside the optimiser kernel
    L->host_step += 1;
    return launch_adam_fused_allreduce(L->params, L->m, L->v, use_tc(L) ? L->tc.w_shadow : nullptr,
                                       use_tc(L) ? L->tc.c1w_perm : nullptr,
                                       use_tc(L) ? L->tc.elem_bytes : 0, L->Ppad, L->P, L->norm_partials,
                                       L->opt, max_norm, total_norm_out, count, m_loss, mb, L->ar,
                                       L->stream);
  }
  if (!no_fused && adam_fused_fits(L->Ppad)) {
    L->host_step += 1;
    return launch_adam_fused(L->params, L->grads, L->m, L->v, use_tc(L) ? L->tc.w_shadow : nullptr,
                             use_tc(L) ? L->tc.c1w_perm : nullptr, use_tc(L) ? L->tc.elem_bytes : 0,
                             L->Ppad, L->P, L->norm_partials, L->opt, max_norm, total_norm_out,
                             L->grads + L->P, count, m_loss, mb, L->stream);
  }
  if (m_loss)
    if (int rc = finalize_metrics(L->grads + L->P, count, m_loss, mb, nullptr, L->stream))
      return rc;
  // the padding behind the P parameters carries the loss sum (see enqueue_train): not part of the norm
  if (int rc = launch_grad_sumsq(L->grads, L->Ppad, L->P, L->norm_partials, L->opt, max_norm,
                                 total_norm_out, L->stream))
    return rc;
  L->host_step += 1;
  return launch_adam_clip(L->params, L->grads, L->m, L->v, use_tc(L) ? L->tc.w_shadow : nullptr,
                          use_tc(L) ? L->tc.c1w_perm : nullptr, use_tc(L) ? L->tc.elem_bytes : 0,
                          L->Ppad, L->opt, L->stream);
}

// the body of train(): everything is stream-ordered, nothing synchronises
int ensure_staging(alegpu_learner *L) {
  if (L->staged.obs)
    return ALEGPU_OK;
  const size_t cap = (size_t)L->cap, A = (size_t)L->cfg.action_size;
  ALEGPU_CUDA_CHECK(cudaMalloc(&L->staged.obs, cap * 28224));
  ALEGPU_CUDA_CHECK(cudaMalloc(&L->staged.actions, cap * sizeof(int64_t)));
  ALEGPU_CUDA_CHECK(cudaMalloc(&L->staged.logp, cap * A * sizeof(float)));
  ALEGPU_CUDA_CHECK(cudaMalloc(&L->staged.adv, cap * sizeof(float)));
  ALEGPU_CUDA_CHECK(cudaMalloc(&L->staged.ret, cap * sizeof(float)));
  ALEGPU_CUDA_CHECK(cudaMalloc(&L->staged.values, cap * sizeof(float)));
  ALEGPU_CUDA_CHECK(cudaMalloc(&L->staged.masks, cap));
  ALEGPU_CUDA_CHECK(cudaMalloc(&L->gather_error, sizeof(int)));
  ALEGPU_CUDA_CHECK(cudaMemset(L->gather_error, 0, sizeof(int)));
  return ALEGPU_OK;
}

int enqueue_train(alegpu_learner *L, const alegpu_batch &b, const alegpu_metrics &mt,
                  int64_t epochs, int64_t nmb, const alegpu_hyperparameters &hp,
                  const alegpu_train_options &opt) {
  const int64_t mb = b.size / nmb, A = L->cfg.action_size;
  const bool shuffled = opt.indices != nullptr;
  // tensor-core path: conv1 reads a bf16 NHWC copy of the observations; with contiguous slices it
  // is made ONCE here for the whole batch (reused by every epoch), else per gathered minibatch
  const void *obs16_all = nullptr;
  if (use_tc(L) && !shuffled)
    if (int rc = net_tc_prepare_obs(L->tc, b.observations, b.size, &obs16_all, L->stream))
      return rc;
  // contiguous slices: the per-minibatch unmasked counts are known up front; shuffled: per step
  if (!shuffled)
    if (int rc = launch_mask_count(b.masks, mb, nmb, L->counts, L->stream))
      return rc;
#ifdef ALEGPU_WITH_NCCL
  if (L->cfg.world_size > 1 && !shuffled) { // the loss normaliser is the GLOBAL unmasked count (losses.cc:19)
    ncclResult_t r = nccl().AllReduce(L->counts, L->counts, (size_t)nmb, ncclFloat, ncclSum, L->comm,
                                   L->stream);
    if (r != ncclSuccess)
      return fail(ALEGPU_CUDA_ERROR, std::string("ncclAllReduce: ") + nccl().GetErrorString(r));
    ALEGPU_COUNT_LAUNCH();
  }
#endif
  for (int64_t ep = 0; ep < epochs; ++ep)
    for (int64_t k = 0; k < nmb; ++k) {
      const int64_t s = k * mb, o = (ep * nmb + k) * mb; // train.h:149-151 contiguous slice
      const uint8_t *obs = b.observations + s * 28224;
      LossArgs la{};
      la.actions = b.actions + s;
      la.old_log_probs = b.log_probabilities + s * A;
      la.advantages = b.advantages + s;
      la.returns = b.returns + s;
      la.masks = b.masks + s;
      la.old_values = opt.clip_value_loss ? b.values + s : nullptr;
      float *count = L->counts + k;
      if (shuffled) { // minibatch k of epoch ep = rows indices[ep*N + k*mb ...], gathered once
        if (int rc = launch_gather_rows(opt.indices + ep * b.size + s, mb, b.size, (int)A,
                                        b.observations, b.actions, b.log_probabilities,
                                        b.advantages, b.returns,
                                        opt.clip_value_loss ? b.values : nullptr, b.masks,
                                        L->staged, L->gather_error, L->stream))
          return rc;
        obs = L->staged.obs;
        la.actions = L->staged.actions;
        la.old_log_probs = L->staged.logp;
        la.advantages = L->staged.adv;
        la.returns = L->staged.ret;
        la.masks = L->staged.masks;
        if (opt.clip_value_loss)
          la.old_values = L->staged.values;
        count = L->counts + (ep * nmb + k) % L->counts_cap;
        if (int rc = launch_mask_count(la.masks, mb, 1, count, L->stream))
          return rc;
#ifdef ALEGPU_WITH_NCCL
        if (L->cfg.world_size > 1) {
          ncclResult_t r = nccl().AllReduce(count, count, 1, ncclFloat, ncclSum, L->comm, L->stream);
          if (r != ncclSuccess)
            return fail(ALEGPU_CUDA_ERROR, std::string("ncclAllReduce: ") + nccl().GetErrorString(r));
          ALEGPU_COUNT_LAUNCH();
        }
#endif
      }
      if (opt.normalize_advantages) {
        double *st = L->adv_stats + 3 * ((ep * nmb + k) % L->counts_cap);
        if (int rc = launch_adv_stats(la.advantages, la.masks, mb, st, L->stream))
          return rc;
#ifdef ALEGPU_WITH_NCCL
        if (L->cfg.world_size > 1) { // global statistics: the minibatch spans all ranks
          ncclResult_t r = nccl().AllReduce(st, st, 3, ncclDouble, ncclSum, L->comm, L->stream);
          if (r != ncclSuccess)
            return fail(ALEGPU_CUDA_ERROR, std::string("ncclAllReduce: ") + nccl().GetErrorString(r));
          ALEGPU_COUNT_LAUNCH();
        }
#endif
        la.adv_stats = st;
      }
      const void *obs16 =
          obs16_all ? (const void *)((const uint8_t *)obs16_all + (size_t)s * 28224 * L->tc.elem_bytes)
                    : nullptr;
      if (int rc = forward(L, obs, mb, obs16))
        return rc;
      la.clip = hp.clip_param;
      la.vcoef = hp.value_loss_coef;
      la.ecoef = hp.entropy_coef;
      la.count = count;
      la.m_clipped = mt.clipped_losses ? mt.clipped_losses + o : nullptr;
      la.m_value = mt.value_losses ? mt.value_losses + o : nullptr;
      la.m_entropy = mt.entropies ? mt.entropies + o : nullptr;
      la.m_total = mt.total_losses ? mt.total_losses + o : nullptr;
      la.m_ratio = mt.ratio ? mt.ratio + o : nullptr;
      la.m_adv = mt.advantages ? mt.advantages + o : nullptr;
      la.m_ret = mt.returns ? mt.returns + o : nullptr;
      la.m_masks = mt.masks ? mt.masks + o : nullptr;
      // the un-normalised loss sum of this rank's rows goes into the first padding slot of the flat
      // gradient: with world_size > 1 it is summed over ranks by the gradient exchange itself, so
      // Metrics.loss is the GLOBAL minibatch loss (losses.cc:19) on every rank
      la.loss_sum = L->grads + L->P;
      la.done_counter = L->loss_done;
      la.error_flag = L->err_dev;
      int nparts = 0;
      if (int rc = heads_loss_backward(L->Pv, L->acts, mb, la, L->loss_partials, &nparts, L->stream))
        return rc;
      if (int rc = backward(L, obs, mb, obs16))
        return rc;
      if (!L->fused_ar) // (the fused path sums the gradient inside the optimiser kernel below)
        if (int rc = allreduce_grads(L))
          return rc;
      // norm + clip + Adam + the Metrics.loss broadcast (train.h:98): one launch
      if (int rc = clip_and_step(L, hp.max_gradient_norm,
                                 mt.clipped_gradients ? mt.clipped_gradients + ep * nmb + k
                                                      : L->scalars,
                                 count, mt.loss ? mt.loss + o : nullptr, mb, /*exchange=*/true))
        return rc;
    }
  return ALEGPU_OK;
}

#ifdef ALEGPU_WITH_NCCL
// Map every peer's flat buffer and exchange buffer (CUDA IPC; the 2 x 64-byte handles travel through
// one ncclAllGather) so the optimiser kernel can pull gradients over NVLink.  Any rank that cannot
// (no peer access, IPC refused, net too large for the register-cached kernel) makes ALL ranks keep the
// ncclAllReduce path: the decision is an all-reduced minimum.  ALEGPU_NO_FUSED_AR=1 forces that path.
int setup_fused_allreduce(alegpu_learner *L) {
  const int world = L->cfg.world_size, rank = L->cfg.rank;
  L->comm_mode = "nccl";
  int ok = 1;
  if (getenv("ALEGPU_NO_FUSED_AR") || getenv("ALEGPU_NO_FUSED_ADAM") || !nccl().AllGather ||
      world > kArMaxRanks || !adam_fused_fits(L->Ppad))
    ok = 0;
  const int64_t slice4 = ceil_div(L->Ppad / 4, world);
  const size_t xbytes = 1024 + (size_t)slice4 * 16;
  struct Handles {
    cudaIpcMemHandle_t flat, xchg;
  } mine{};
  static_assert(sizeof(Handles) == 128, "two 64-byte IPC handles");
  if (ok) {
    ALEGPU_CUDA_CHECK(cudaMalloc(&L->xchg, xbytes));
    ALEGPU_CUDA_CHECK(cudaMemset(L->xchg, 0, xbytes));
    if (cudaIpcGetMemHandle(&mine.flat, L->flat) != cudaSuccess ||
        cudaIpcGetMemHandle(&mine.xchg, L->xchg) != cudaSuccess) {
      cudaGetLastError();
      ok = 0;
    }
  }
  char *d_buf = nullptr; // [mine | all ranks | ok flag]
  ALEGPU_CUDA_CHECK(cudaMalloc(&d_buf, 128 * (size_t)(world + 1) + 16));
  ALEGPU_CUDA_CHECK(cudaMemcpyAsync(d_buf, &mine, 128, cudaMemcpyHostToDevice, L->stream));
  ncclResult_t r = nccl().AllGather(d_buf, d_buf + 128, 128, ncclChar, L->comm, L->stream);
  if (r != ncclSuccess)
    return fail(ALEGPU_CUDA_ERROR, std::string("ncclAllGather: ") + nccl().GetErrorString(r));
  std::vector<Handles> all((size_t)world);
  ALEGPU_CUDA_CHECK(cudaMemcpyAsync(all.data(), d_buf + 128, 128 * (size_t)world, cudaMemcpyDeviceToHost,
                                    L->stream));
  ALEGPU_CUDA_CHECK(cudaStreamSynchronize(L->stream));
  ArParams ar{};
  ar.world = world;
  ar.rank = rank;
  ar.slice4 = slice4;
  ar.error_flag = L->err_dev;
  for (int p = 0; p < world && ok; ++p) {
    void *flat = L->flat, *xchg = L->xchg;
    if (p != rank) {
      if (cudaIpcOpenMemHandle(&flat, all[(size_t)p].flat, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
        cudaGetLastError();
        ok = 0;
        break;
      }
      L->ipc_opened.push_back(flat);
      if (cudaIpcOpenMemHandle(&xchg, all[(size_t)p].xchg, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
        cudaGetLastError();
        ok = 0;
        break;
      }
      L->ipc_opened.push_back(xchg);
    }
    ar.grads[p] = static_cast<float *>(flat) + L->Ppad; // params | grads | m | v
    ar.sig[p] = static_cast<ArSignals *>(xchg);
    ar.xbuf[p] = reinterpret_cast<float *>(static_cast<char *>(xchg) + 1024);
  }
  float *d_ok = reinterpret_cast<float *>(d_buf + 128 * (size_t)(world + 1));
  const float okf = (float)ok;
  ALEGPU_CUDA_CHECK(cudaMemcpyAsync(d_ok, &okf, sizeof(float), cudaMemcpyHostToDevice, L->stream));
  r = nccl().AllReduce(d_ok, d_ok, 1, ncclFloat, ncclMin, L->comm, L->stream);
  if (r != ncclSuccess)
    return fail(ALEGPU_CUDA_ERROR, std::string("ncclAllReduce: ") + nccl().GetErrorString(r));
  float all_ok = 0.0f;
  ALEGPU_CUDA_CHECK(cudaMemcpyAsync(&all_ok, d_ok, sizeof(float), cudaMemcpyDeviceToHost, L->stream));
  ALEGPU_CUDA_CHECK(cudaStreamSynchronize(L->stream));
  cudaFree(d_buf);
  if (all_ok >= 1.0f) {
    L->ar = ar;
    L->fused_ar = true;
    L->comm_mode = "fused-nvlink";
  } else {
    for (void *q : L->ipc_opened)
      cudaIpcCloseMemHandle(q);
    L->ipc_opened.clear();
    if (L->xchg)
      cudaFree(L->xchg);
    L->xchg = nullptr;
  }
  return ALEGPU_OK;
}
#endif

template <typename T> void key_push(std::vector<uint64_t> &k, T v) {
  uint64_t x = 0;
  std::memcpy(&x, &v, sizeof(T) < 8 ? sizeof(T) : 8);
  k.push_back(x);
}

} // namespace

extern "C" const char *alegpu_last_error(void) { return g_error.c_str(); }

extern "C" const char *alegpu_version(void) {
#ifdef ALEGPU_WITH_NCCL
  return "alegpu 0.1 sm_100a nccl=1";
#else
  return "alegpu 0.1 sm_100a nccl=0";
#endif
}

// ---- memory helpers ---------------------------------------------------------------------------
extern "C" int alegpu_device_count(int *count) {
  ALEGPU_CUDA_CHECK(cudaGetDeviceCount(count));
  return ALEGPU_OK;
}
extern "C" int alegpu_set_device(int device) {
  ALEGPU_CUDA_CHECK(cudaSetDevice(device));
  return ALEGPU_OK;
}
extern "C" int alegpu_malloc(void **dptr, size_t bytes) {
  ALEGPU_CUDA_CHECK(cudaMalloc(dptr, bytes ? bytes : 1));
  return ALEGPU_OK;
}
extern "C" int alegpu_free(void *dptr) {
  ALEGPU_CUDA_CHECK(cudaFree(dptr));
  return ALEGPU_OK;
}
extern "C" int alegpu_host_alloc(void **hptr, size_t bytes) {
  ALEGPU_CUDA_CHECK(cudaHostAlloc(hptr, bytes ? bytes : 1, cudaHostAllocDefault));
  return ALEGPU_OK;
}
extern "C" int alegpu_host_free(void *hptr) {
  ALEGPU_CUDA_CHECK(cudaFreeHost(hptr));
  return ALEGPU_OK;
}
extern "C" int alegpu_memcpy_h2d(void *dst, const void *src, size_t bytes, void *stream) {
  ALEGPU_CUDA_CHECK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, as_stream(stream)));
  return ALEGPU_OK;
}
extern "C" int alegpu_memcpy_d2h(void *dst, const void *src, size_t bytes, void *stream) {
  ALEGPU_CUDA_CHECK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, as_stream(stream)));
  return ALEGPU_OK;
}
extern "C" int alegpu_memcpy_d2d(void *dst, const void *src, size_t bytes, void *stream) {
  ALEGPU_CUDA_CHECK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, as_stream(stream)));
  return ALEGPU_OK;
}
extern "C" int alegpu_memset(void *dst, int value, size_t bytes, void *stream) {
  ALEGPU_CUDA_CHECK(cudaMemsetAsync(dst, value, bytes, as_stream(stream)));
  return ALEGPU_OK;
}
extern "C" int alegpu_stream_synchronize(void *stream) {
  ALEGPU_CUDA_CHECK(cudaStreamSynchronize(as_stream(stream)));
  return ALEGPU_OK;
}
extern "C" int alegpu_stream_create(void **stream) {
  cudaStream_t s;
  ALEGPU_CUDA_CHECK(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
  *stream = s;
  return ALEGPU_OK;
}
extern "C" int alegpu_stream_destroy(void *stream) {
  ALEGPU_CUDA_CHECK(cudaStreamDestroy(as_stream(stream)));
  return ALEGPU_OK;
}
extern "C" int alegpu_event_create(void **event) {
  cudaEvent_t e;
  ALEGPU_CUDA_CHECK(cudaEventCreate(&e));
  *event = e;
  return ALEGPU_OK;
}
extern "C" int alegpu_event_destroy(void *event) {
  ALEGPU_CUDA_CHECK(cudaEventDestroy(static_cast<cudaEvent_t>(event)));
  return ALEGPU_OK;
}
extern "C" int alegpu_event_record(void *event, void *stream) {
  ALEGPU_CUDA_CHECK(cudaEventRecord(static_cast<cudaEvent_t>(event), as_stream(stream)));
  return ALEGPU_OK;
}
extern "C" int alegpu_stream_wait_event(void *stream, void *event) {
  ALEGPU_CUDA_CHECK(cudaStreamWaitEvent(as_stream(stream), static_cast<cudaEvent_t>(event), 0));
  return ALEGPU_OK;
}
extern "C" int alegpu_event_synchronize(void *event) {
  ALEGPU_CUDA_CHECK(cudaEventSynchronize(static_cast<cudaEvent_t>(event)));
  return ALEGPU_OK;
}
extern "C" int alegpu_event_elapsed_ms(void *start, void *stop, float *ms) {
  ALEGPU_CUDA_CHECK(cudaEventElapsedTime(ms, static_cast<cudaEvent_t>(start),
                                         static_cast<cudaEvent_t>(stop)));
  return ALEGPU_OK;
}
extern "C" int64_t alegpu_launch_count(void) { return g_launches; }

// Launch tracer: between _begin and _end every kernel launch site on `stream` (this thread) is
// followed by a CUDA event; _end writes "file:line count total_us" lines (device time from the
// previous event to this one, i.e. kernel + launch gap) sorted by total time.
extern "C" int alegpu_trace_begin(void *stream) {
  for (auto &r : g_trace)
    cudaEventDestroy(r.ev);
  g_trace.clear();
  g_trace_stream = as_stream(stream);
  g_trace_on = true;
  trace_launch("begin", 0);
  return ALEGPU_OK;
}
extern "C" int alegpu_trace_end(char *out, size_t out_bytes) {
  g_trace_on = false;
  if (g_trace.empty())
    return ALEGPU_OK;
  ALEGPU_CUDA_CHECK(cudaEventSynchronize(g_trace.back().ev));
  struct Agg { std::string key; int n; double us; };
  std::vector<Agg> agg;
  for (size_t i = 1; i < g_trace.size(); ++i) {
    float ms = 0.0f;
    cudaEventElapsedTime(&ms, g_trace[i - 1].ev, g_trace[i].ev);
    std::string f = g_trace[i].file;
    const size_t slash = f.find_last_of('/');
    if (slash != std::string::npos)
      f = f.substr(slash + 1);
    std::string key = f + ":" + std::to_string(g_trace[i].line);
    auto it = std::find_if(agg.begin(), agg.end(), [&](const Agg &a) { return a.key == key; });
    if (it == agg.end())
      agg.push_back({key, 1, ms * 1e3});
    else {
      it->n += 1;
      it->us += ms * 1e3;
    }
  }
  std::sort(agg.begin(), agg.end(), [](const Agg &a, const Agg &b) { return a.us > b.us; });
  std::string text;
  for (auto &a : agg)
    text += a.key + " " + std::to_string(a.n) + " " + std::to_string(a.us) + "\n";
  if (out && out_bytes) {
    const size_t n = std::min(out_bytes - 1, text.size());
    std::memcpy(out, text.data(), n);
    out[n] = 0;
  }
  for (auto &r : g_trace)
    cudaEventDestroy(r.ev);
  g_trace.clear();
  return ALEGPU_OK;
}

extern "C" int alegpu_nccl_unique_id(void *out128) {
#ifdef ALEGPU_WITH_NCCL
  static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId size");
  if (!nccl().ok)
    return fail(ALEGPU_RUNTIME_ERROR, "libnccl.so.2 could not be loaded");
  ncclUniqueId id;
  ncclResult_t r = nccl().GetUniqueId(&id);
  if (r != ncclSuccess)
    return fail(ALEGPU_CUDA_ERROR, std::string("ncclGetUniqueId: ") + nccl().GetErrorString(r));
  std::memcpy(out128, &id, 128);
  return ALEGPU_OK;
#else
  (void)out128;
  return fail(ALEGPU_RUNTIME_ERROR, "built without NCCL");
#endif
}

// ---- learner ------------------------------------------------------------------------------------
extern "C" int alegpu_learner_create(const alegpu_learner_config *config, alegpu_learner **out) {
  ALEGPU_REQUIRE(config && out, ALEGPU_INVALID_ARGUMENT, "learner_create: null argument");
  ALEGPU_REQUIRE(config->hidden_size >= 1 && config->hidden_size <= 8192, ALEGPU_INVALID_ARGUMENT,
                 "learner_create: hidden_size out of range");
  ALEGPU_REQUIRE(config->action_size >= 1 && config->action_size <= 18, ALEGPU_INVALID_ARGUMENT,
                 "learner_create: action_size must be in [1,18]");
  ALEGPU_REQUIRE(config->max_minibatch >= 1, ALEGPU_INVALID_ARGUMENT,
                 "learner_create: max_minibatch must be >= 1");
  ALEGPU_REQUIRE(config->math == ALEGPU_MATH_FP32 || config->math == ALEGPU_MATH_BF16 ||
                     config->math == ALEGPU_MATH_TF32,
                 ALEGPU_INVALID_ARGUMENT, "learner_create: unknown math mode");
  if (config->math != ALEGPU_MATH_FP32)
    ALEGPU_REQUIRE(config->hidden_size % 16 == 0, ALEGPU_INVALID_ARGUMENT,
                   "learner_create: the tensor-core path needs hidden_size % 16 == 0");
  ALEGPU_CUDA_CHECK(cudaSetDevice(config->device));
  auto L = std::make_unique<alegpu_learner>();
  L->cfg = *config;
  if (L->cfg.world_size < 1)
    L->cfg.world_size = 1;
  const int64_t H = config->hidden_size, A = config->action_size, cap = config->max_minibatch;
  L->P = net_param_count(H, A);
  L->Ppad = ceil_div(L->P + 1, 8) * 8; // >= 1 padding slot: grads[P] carries the loss sum
  L->cap = cap;
  L->lr = config->learning_rate;
  ALEGPU_CUDA_CHECK(cudaStreamCreateWithFlags(&L->stream, cudaStreamNonBlocking));
  ALEGPU_CUDA_CHECK(cudaMalloc(&L->flat, sizeof(float) * 4 * L->Ppad));
  ALEGPU_CUDA_CHECK(cudaMemsetAsync(L->flat, 0, sizeof(float) * 4 * L->Ppad, L->stream));
  L->params = L->flat;
  L->grads = L->flat + L->Ppad;
  L->m = L->flat + 2 * L->Ppad;
  L->v = L->flat + 3 * L->Ppad;
  L->Pv = net_params_view(L->params, H, A);
  L->Gv = net_params_view(L->grads, H, A);
  NetActs &a = L->acts;
  a.cap = cap;
  const bool tc = config->math != ALEGPU_MATH_FP32;
  auto fmalloc = [&](float **p, size_t n) -> int {
    ALEGPU_CUDA_CHECK(cudaMalloc(p, sizeof(float) * (n ? n : 1)));
    return ALEGPU_OK;
  };
  if (!tc) {
    if (int rc = fmalloc(&a.act1, (size_t)cap * 12800)) return rc;
    if (int rc = fmalloc(&a.act2, (size_t)cap * 5184)) return rc;
    if (int rc = fmalloc(&a.act3, (size_t)cap * 3136)) return rc;
    if (int rc = fmalloc(&a.dact1, (size_t)cap * 12800)) return rc;
    if (int rc = fmalloc(&a.dact2, (size_t)cap * 5184)) return rc;
    if (int rc = fmalloc(&a.dact3, (size_t)cap * 3136)) return rc;
  }
  if (int rc = fmalloc(&a.fc, (size_t)cap * H)) return rc;
  if (int rc = fmalloc(&a.dfc, (size_t)cap * H)) return rc;
  if (int rc = fmalloc(&a.logits, (size_t)cap * A)) return rc;
  if (int rc = fmalloc(&a.values, (size_t)cap)) return rc;
  if (int rc = fmalloc(&a.dout, (size_t)cap * (A + 1))) return rc;
  if (tc) { // room for 24 reduction splits of up to 1024 rows
    a.fc_ws_floats = (size_t)std::min<int64_t>(cap, 1024) * H * 24;
    if (int rc = fmalloc(&a.fc_ws, a.fc_ws_floats)) return rc;
  }
  L->ws_floats = std::max<size_t>((size_t)4 << 20, (size_t)H * 3137 + 1024);
  if (int rc = fmalloc(&L->ws, L->ws_floats)) return rc;
  if (int rc = fmalloc(&L->norm_partials, kNormBlocks)) return rc;
  if (int rc = fmalloc(&L->loss_partials, kNumSMs * 4 + 8)) return rc;
  L->counts_cap = 4096;
  if (int rc = fmalloc(&L->counts, L->counts_cap)) return rc;
  ALEGPU_CUDA_CHECK(cudaMalloc(&L->adv_stats, sizeof(double) * 3 * (size_t)L->counts_cap));
  if (int rc = fmalloc(&L->scalars, 16)) return rc;
  ALEGPU_CUDA_CHECK(cudaMalloc(&L->opt, sizeof(OptState)));
  ALEGPU_CUDA_CHECK(cudaMalloc(&L->loss_done, sizeof(unsigned int)));
  ALEGPU_CUDA_CHECK(cudaMemset(L->loss_done, 0, sizeof(unsigned int)));
  ALEGPU_CUDA_CHECK(cudaHostAlloc(&L->err_host, sizeof(int), cudaHostAllocMapped));
  *L->err_host = 0;
  ALEGPU_CUDA_CHECK(cudaHostGetDevicePointer(&L->err_dev, L->err_host, 0));
  if (tc)
    if (int rc = net_tc_init(L->tc, H, A, cap, L->Ppad, config->math == ALEGPU_MATH_TF32 ? 4 : 2,
                             a.dfc, L->stream))
      return rc;
  a.dfc16 = L->tc.dfc16; // bf16 mode: the heads kernel also emits dfc in bf16 for the tensor-core backward
  a.dfc_round_tf32 = config->math == ALEGPU_MATH_TF32; // tf32 mode: it rounds dfc itself to tf32
#ifdef ALEGPU_WITH_NCCL
  if (L->cfg.world_size > 1) {
    ALEGPU_REQUIRE(config->nccl_unique_id, ALEGPU_INVALID_ARGUMENT,
                   "learner_create: world_size > 1 needs nccl_unique_id");
    ALEGPU_REQUIRE(nccl().ok, ALEGPU_RUNTIME_ERROR, "libnccl.so.2 could not be loaded");
    ncclUniqueId id;
    std::memcpy(&id, config->nccl_unique_id, 128);
    ncclResult_t r = nccl().CommInitRank(&L->comm, L->cfg.world_size, id, L->cfg.rank);
    if (r != ncclSuccess)
      return fail(ALEGPU_CUDA_ERROR, std::string("ncclCommInitRank: ") + nccl().GetErrorString(r));
    if (tc && nccl().CommSplit && !getenv("ALEGPU_NO_COMM_OVERLAP")) {
      r = nccl().CommSplit(L->comm, 0, L->cfg.rank, &L->comm2, nullptr);
      if (r != ncclSuccess)
        return fail(ALEGPU_CUDA_ERROR, std::string("ncclCommSplit: ") + nccl().GetErrorString(r));
      ALEGPU_CUDA_CHECK(cudaStreamCreateWithFlags(&L->comm_stream, cudaStreamNonBlocking));
      ALEGPU_CUDA_CHECK(cudaEventCreateWithFlags(&L->ev_comm, cudaEventDisableTiming));
    }
    if (int rc = setup_fused_allreduce(L.get()))
      return rc;
  }
#else
  ALEGPU_REQUIRE(L->cfg.world_size == 1, ALEGPU_RUNTIME_ERROR,
                 "learner_create: world_size > 1 but built without NCCL");
#endif
  if (int rc = sync_opt_state(L.get()))
    return rc;
  *out = L.release();
  return ALEGPU_OK;
}

extern "C" int alegpu_learner_destroy(alegpu_learner *L) {
  if (!L)
    return ALEGPU_OK;
  cudaSetDevice(L->cfg.device);
  cudaStreamSynchronize(L->stream);
  // A captured graph that holds NCCL kernel nodes keeps a reference on its communicator (NCCL hangs a
  // cudaUserObject on the graph) and ncclCommDestroy waits for that reference to drop: the graph
  // must die FIRST.  (Destroying the communicators first was the teardown hang of round 1.)
  if (L->graph_exec) {
    cudaGraphExecDestroy(L->graph_exec);
    L->graph_exec = nullptr;
  }
#ifdef ALEGPU_WITH_NCCL
  if (L->comm_stream)
    cudaStreamSynchronize(L->comm_stream);
  if (L->fused_ar) {
    // a peer's last optimiser kernel may still be pulling this rank's exchange buffer: every rank
    // reaches this point with ITS stream drained, so one tiny collective is the handshake after which
    // the mappings can go
    nccl().AllReduce(L->counts, L->counts, 1, ncclFloat, ncclSum, L->comm, L->stream);
    cudaStreamSynchronize(L->stream);
    for (void *q : L->ipc_opened)
      cudaIpcCloseMemHandle(q);
    L->ipc_opened.clear();
    // (freeing the exported exchange buffer happens below, after the communicators are gone)
  }
  if (L->comm2)
    nccl().CommDestroy(L->comm2);
  if (L->comm)
    nccl().CommDestroy(L->comm);
  if (L->comm_stream)
    cudaStreamDestroy(L->comm_stream);
  if (L->ev_comm)
    cudaEventDestroy(L->ev_comm);
#endif
  if (L->graph_exec)
    cudaGraphExecDestroy(L->graph_exec);
  NetActs &a = L->acts;
  for (float *p : {L->flat, a.act1, a.act2, a.act3, a.fc, a.dact1, a.dact2, a.dact3, a.dfc,
                   a.logits, a.values, a.dout, a.fc_ws, L->ws, L->norm_partials, L->loss_partials,
                   L->counts, L->scalars})
    if (p)
      cudaFree(p);
  if (L->opt)
    cudaFree(L->opt);
  if (L->loss_done)
    cudaFree(L->loss_done);
  for (cudaEvent_t e : {L->ev_bridge_in, L->ev_bridge_out})
    if (e)
      cudaEventDestroy(e);
  if (L->xchg)
    cudaFree(L->xchg);
  if (L->err_host)
    cudaFreeHost(L->err_host);
  net_tc_destroy(L->tc);
  for (void *p : {(void *)L->staged.obs, (void *)L->staged.actions, (void *)L->staged.logp,
                  (void *)L->staged.adv, (void *)L->staged.ret, (void *)L->staged.values,
                  (void *)L->staged.masks,
                  (void *)L->adv_stats, (void *)L->gather_error})
    if (p)
      cudaFree(p);
  cudaStreamDestroy(L->stream);
  delete L;
  return ALEGPU_OK;
}

extern "C" int64_t alegpu_learner_param_count(const alegpu_learner *L) { return L ? L->P : 0; }
extern "C" int alegpu_learner_wait_stream(alegpu_learner *L, void *stream) {
  ALEGPU_REQUIRE(L, ALEGPU_INVALID_ARGUMENT, "learner_wait_stream: null learner");
  if (as_stream(stream) == L->stream)
    return ALEGPU_OK;
  ALEGPU_CUDA_CHECK(cudaSetDevice(L->cfg.device));
  if (!L->ev_bridge_in)
    ALEGPU_CUDA_CHECK(cudaEventCreateWithFlags(&L->ev_bridge_in, cudaEventDisableTiming));
  ALEGPU_CUDA_CHECK(cudaEventRecord(L->ev_bridge_in, as_stream(stream)));
  ALEGPU_CUDA_CHECK(cudaStreamWaitEvent(L->stream, L->ev_bridge_in, 0));
  return ALEGPU_OK;
}
extern "C" int alegpu_stream_wait_learner(alegpu_learner *L, void *stream) {
  ALEGPU_REQUIRE(L, ALEGPU_INVALID_ARGUMENT, "stream_wait_learner: null learner");
  if (as_stream(stream) == L->stream)
    return ALEGPU_OK;
  ALEGPU_CUDA_CHECK(cudaSetDevice(L->cfg.device));
  if (!L->ev_bridge_out)
    ALEGPU_CUDA_CHECK(cudaEventCreateWithFlags(&L->ev_bridge_out, cudaEventDisableTiming));
  ALEGPU_CUDA_CHECK(cudaEventRecord(L->ev_bridge_out, L->stream));
  ALEGPU_CUDA_CHECK(cudaStreamWaitEvent(as_stream(stream), L->ev_bridge_out, 0));
  return ALEGPU_OK;
}
extern "C" const char *alegpu_learner_comm_mode(const alegpu_learner *L) {
  return L ? L->comm_mode.c_str() : "";
}
extern "C" void *alegpu_learner_stream(alegpu_learner *L) { return L ? L->stream : nullptr; }

extern "C" int alegpu_learner_set_params(alegpu_learner *L, const float *params_host) {
  ALEGPU_REQUIRE(L && params_host, ALEGPU_INVALID_ARGUMENT, "set_params: null argument");
  ALEGPU_CUDA_CHECK(cudaSetDevice(L->cfg.device));
  if (int rc = upload_flat(L, L->params, params_host))
    return rc;
  if (use_tc(L))
    if (int rc = net_tc_refresh_weights(L->tc, L->params, L->Ppad, L->stream))
      return rc;
  ALEGPU_CUDA_CHECK(cudaStreamSynchronize(L->stream));
  return ALEGPU_OK;
}

extern "C" int alegpu_learner_get_params(alegpu_learner *L, float *params_host) {
  ALEGPU_REQUIRE(L && params_host, ALEGPU_INVALID_ARGUMENT, "get_params: null argument");
  ALEGPU_CUDA_CHECK(cudaSetDevice(L->cfg.device));
  return download_flat(L, L->params, params_host);
}

extern "C" int alegpu_learner_get_adam(alegpu_learner *L, float *exp_avg_host,
                                       float *exp_avg_sq_host, int64_t *step) {
  ALEGPU_REQUIRE(L, ALEGPU_INVALID_ARGUMENT, "get_adam: null learner");
  ALEGPU_CUDA_CHECK(cudaSetDevice(L->cfg.device));
  if (exp_avg_host)
    if (int rc = download_flat(L, L->m, exp_avg_host))
      return rc;
  if (exp_avg_sq_host)
    if (int rc = download_flat(L, L->v, exp_avg_sq_host))
      return rc;
  if (step) {
    OptState h{};
    ALEGPU_CUDA_CHECK(cudaMemcpyAsync(&h, L->opt, sizeof(h), cudaMemcpyDeviceToHost, L->stream));
    ALEGPU_CUDA_CHECK(cudaStreamSynchronize(L->stream));
    *step = h.step;
  }
  return ALEGPU_OK;
}

// checkpoint restore: Adam moments (reference order/layout) and the optimiser step counter
extern "C" int alegpu_learner_set_adam(alegpu_learner *L, const float *exp_avg_host,
                                       const float *exp_avg_sq_host, int64_t step) {
  ALEGPU_REQUIRE(L && exp_avg_host && exp_avg_sq_host && step >= 0, ALEGPU_INVALID_ARGUMENT,
                 "set_adam: bad argument");
  ALEGPU_CUDA_CHECK(cudaSetDevice(L->cfg.device));
  if (int rc = upload_flat(L, L->m, exp_avg_host))
    return rc;
  if (int rc = upload_flat(L, L->v, exp_avg_sq_host))
    return rc;
  L->host_step = step;
  return sync_opt_state(L);
}

extern "C" int alegpu_learner_set_lr(alegpu_learner *L, double lr) {
  ALEGPU_REQUIRE(L, ALEGPU_INVALID_ARGUMENT, "set_lr: null learner");
  ALEGPU_CUDA_CHECK(cudaSetDevice(L->cfg.device));
  L->lr = lr;
  // stream-ordered 8-byte update of OptState.lr (a pageable source is staged by the runtime)
  ALEGPU_CUDA_CHECK(cudaMemcpyAsync(&L->opt->lr, &L->lr, sizeof(double), cudaMemcpyHostToDevice,
                                    L->stream));
  return ALEGPU_OK;
}

extern "C" int alegpu_learner_forward(alegpu_learner *L, const uint8_t *observations, int64_t n,
                                      float *logits, float *values) {
  ALEGPU_REQUIRE(L && observations, ALEGPU_INVALID_ARGUMENT, "forward: null argument");
  ALEGPU_REQUIRE(n >= 0 && n <= L->cap, ALEGPU_INVALID_ARGUMENT,
                 "forward: n exceeds max_minibatch");
  ALEGPU_CUDA_CHECK(cudaSetDevice(L->cfg.device));
  if (n == 0)
    return ALEGPU_OK;
  if (int rc = forward(L, observations, n))
    return rc;
  return heads_forward(L->Pv, L->acts.fc, n, logits ? logits : L->acts.logits,
                       values ? values : L->acts.values, L->stream);
}

// The rollout-side policy call (action selector lambda, src/bin/train.cc:367-379): forward of the
// E current observation stacks, softmax, one categorical draw per env; logits / values / actions are
// written with caller-given strides so they can land directly in column t of the rollout storage
// (Buffer::add, buffer.cc:41-56), and the actions also go to a dense [E] array (device, optionally
// mirrored into pinned host memory in ONE copy instead of E `.item()` syncs, rollout.cc:313).
extern "C" int alegpu_learner_act(alegpu_learner *L, const uint8_t *observations, int64_t E,
                                  const float *uniforms, uint64_t seed, uint64_t step,
                                  float *logits, int64_t logits_env_stride, float *values,
                                  int64_t values_env_stride, int64_t *actions,
                                  int64_t actions_env_stride, int64_t *actions_dense,
                                  int64_t *actions_host) {
  ALEGPU_REQUIRE(L && observations, ALEGPU_INVALID_ARGUMENT, "act: null argument");
  ALEGPU_REQUIRE(E >= 0, ALEGPU_INVALID_ARGUMENT, "act: negative env count");
  ALEGPU_REQUIRE(!actions_host || actions_dense, ALEGPU_INVALID_ARGUMENT,
                 "act: actions_host needs the dense device array it is copied from");
  ALEGPU_CUDA_CHECK(cudaSetDevice(L->cfg.device));
  const int A = (int)L->Pv.actions;
  for (int64_t e0 = 0; e0 < E; e0 += L->cap) { // chunks of at most max_minibatch rows
    const int64_t n = std::min<int64_t>(L->cap, E - e0);
    if (int rc = forward(L, observations + (size_t)e0 * 28224, n))
      return rc;
    if (int rc = heads_forward(L->Pv, L->acts.fc, n, L->acts.logits, L->acts.values, L->stream))
      return rc;
    if (int rc = launch_act_sample(
            L->acts.logits, L->acts.values, n, A, uniforms ? uniforms + e0 : nullptr, seed,
            step * 0x100000000ull + (uint64_t)e0, logits ? logits + e0 * logits_env_stride : nullptr,
            logits_env_stride, values ? values + e0 * values_env_stride : nullptr,
            values_env_stride, actions ? actions + e0 * actions_env_stride : nullptr,
            actions_env_stride, actions_dense ? actions_dense + e0 : nullptr, L->stream))
      return rc;
  }
  if (actions_host && E > 0)
    ALEGPU_CUDA_CHECK(cudaMemcpyAsync(actions_host, actions_dense, sizeof(int64_t) * (size_t)E,
                                      cudaMemcpyDeviceToHost, L->stream));
  return ALEGPU_OK;
}

extern "C" int alegpu_learner_get_activations(alegpu_learner *L, int64_t n, float *c1, float *c2,
                                              float *c3, float *fc) {
  ALEGPU_REQUIRE(L && n >= 0 && n <= L->cap, ALEGPU_INVALID_ARGUMENT, "get_activations: bad n");
  ALEGPU_CUDA_CHECK(cudaSetDevice(L->cfg.device));
  std::vector<float> tmp;
  auto fetch = [&](int which, int C, int HW, float *dst) -> int {
    if (!dst)
      return ALEGPU_OK;
    tmp.resize((size_t)n * C * HW);
    if (use_tc(L)) {
      if (int rc = net_tc_fetch_activation(L->tc, which, n, tmp.data(), L->stream))
        return rc;
    } else {
      const float *src = which == 1 ? L->acts.act1 : which == 2 ? L->acts.act2 : L->acts.act3;
      ALEGPU_CUDA_CHECK(cudaMemcpyAsync(tmp.data(), src, sizeof(float) * tmp.size(),
                                        cudaMemcpyDeviceToHost, L->stream));
      ALEGPU_CUDA_CHECK(cudaStreamSynchronize(L->stream));
    }
    for (int64_t i = 0; i < n; ++i) // NHWC -> NCHW
      for (int p = 0; p < HW; ++p)
        for (int c = 0; c < C; ++c)
          dst[(i * C + c) * HW + p] = tmp[(i * HW + p) * C + c];
    return ALEGPU_OK;
  };
  if (int rc = fetch(1, 32, 400, c1)) return rc;
  if (int rc = fetch(2, 64, 81, c2)) return rc;
  if (int rc = fetch(3, 64, 49, c3)) return rc;
  if (fc) {
    ALEGPU_CUDA_CHECK(cudaMemcpyAsync(fc, L->acts.fc, sizeof(float) * n * L->cfg.hidden_size,
                                      cudaMemcpyDeviceToHost, L->stream));
    ALEGPU_CUDA_CHECK(cudaStreamSynchronize(L->stream));
  }
  return ALEGPU_OK;
}

extern "C" int alegpu_learner_grads(alegpu_learner *L, const alegpu_batch *mbatch,
                                    const alegpu_hyperparameters *hp, float *grads_host,
                                    float *loss_host, float *logits, float *values) {
  ALEGPU_REQUIRE(L && mbatch && hp, ALEGPU_INVALID_ARGUMENT, "grads: null argument");
  const int64_t n = mbatch->size, A = L->cfg.action_size;
  ALEGPU_REQUIRE(n >= 1 && n <= L->cap, ALEGPU_INVALID_ARGUMENT, "grads: size exceeds max_minibatch");
  ALEGPU_CUDA_CHECK(cudaSetDevice(L->cfg.device));
  if (int rc = launch_mask_count(mbatch->masks, n, 1, L->counts, L->stream))
    return rc;
  if (int rc = forward(L, mbatch->observations, n))
    return rc;
  LossArgs la{};
  la.actions = mbatch->actions;
  la.old_log_probs = mbatch->log_probabilities;
  la.advantages = mbatch->advantages;
  la.returns = mbatch->returns;
  la.masks = mbatch->masks;
  la.clip = hp->clip_param;
  la.vcoef = hp->value_loss_coef;
  la.ecoef = hp->entropy_coef;
  la.count = L->counts;
  la.loss_sum = L->scalars + 2;
  la.done_counter = L->loss_done;
  la.error_flag = L->err_dev;
  int nparts = 0;
  if (int rc = heads_loss_backward(L->Pv, L->acts, n, la, L->loss_partials, &nparts, L->stream))
    return rc;
  if (int rc = backward(L, mbatch->observations, n))
    return rc;
  if (int rc = finalize_metrics(L->scalars + 2, L->counts, nullptr, 0, L->scalars + 1, L->stream))
    return rc;
  if (logits)
    ALEGPU_CUDA_CHECK(cudaMemcpyAsync(logits, L->acts.logits, sizeof(float) * n * A,
                                      cudaMemcpyDeviceToDevice, L->stream));
  if (values)
    ALEGPU_CUDA_CHECK(cudaMemcpyAsync(values, L->acts.values, sizeof(float) * n,
                                      cudaMemcpyDeviceToDevice, L->stream));
  if (loss_host) {
    ALEGPU_CUDA_CHECK(cudaMemcpyAsync(loss_host, L->scalars + 1, sizeof(float),
                                      cudaMemcpyDeviceToHost, L->stream));
    ALEGPU_CUDA_CHECK(cudaStreamSynchronize(L->stream));
  }
  if (grads_host)
    return download_flat(L, L->grads, grads_host);
  ALEGPU_CUDA_CHECK(cudaStreamSynchronize(L->stream));
  return ALEGPU_OK;
}

extern "C" int alegpu_learner_apply_grads(alegpu_learner *L, const float *grads_host,
                                          double max_norm, float *total_norm_host) {
  ALEGPU_REQUIRE(L && grads_host, ALEGPU_INVALID_ARGUMENT, "apply_grads: null argument");
  ALEGPU_CUDA_CHECK(cudaSetDevice(L->cfg.device));
  if (int rc = upload_flat(L, L->grads, grads_host))
    return rc;
  if (int rc = clip_and_step(L, (float)max_norm, L->scalars))
    return rc;
  float tn = 0.0f;
  ALEGPU_CUDA_CHECK(cudaMemcpyAsync(&tn, L->scalars, sizeof(float), cudaMemcpyDeviceToHost,
                                    L->stream));
  ALEGPU_CUDA_CHECK(cudaStreamSynchronize(L->stream));
  if (total_norm_host)
    *total_norm_host = tn;
  return ALEGPU_OK;
}

extern "C" int alegpu_learner_train(alegpu_learner *L, const alegpu_batch *batch,
                                    const alegpu_metrics *metrics, int64_t num_epochs,
                                    int64_t num_mini_batches, const alegpu_hyperparameters *hp,
                                    int flags) {
  return alegpu_learner_train_ex(L, batch, metrics, num_epochs, num_mini_batches, hp, nullptr,
                                 flags);
}

extern "C" int alegpu_learner_train_ex(alegpu_learner *L, const alegpu_batch *batch,
                                       const alegpu_metrics *metrics, int64_t num_epochs,
                                       int64_t num_mini_batches, const alegpu_hyperparameters *hp,
                                       const alegpu_train_options *options, int flags) {
  ALEGPU_REQUIRE(L && batch && hp, ALEGPU_INVALID_ARGUMENT, "train: null argument");
  ALEGPU_REQUIRE(num_mini_batches >= 1 && num_epochs >= 0, ALEGPU_INVALID_ARGUMENT,
                 "train: bad epoch / minibatch count");
  ALEGPU_REQUIRE(batch->size % num_mini_batches == 0, ALEGPU_RUNTIME_ERROR,
                 "Batch size must be divisible by num_mini_batches"); // train.h:140-143
  const int64_t mb = batch->size / num_mini_batches;
  ALEGPU_REQUIRE(mb >= 1 && mb <= L->cap, ALEGPU_INVALID_ARGUMENT,
                 "train: minibatch exceeds max_minibatch");
  ALEGPU_REQUIRE(num_mini_batches <= L->counts_cap, ALEGPU_INVALID_ARGUMENT,
                 "train: too many minibatches");
  ALEGPU_REQUIRE(batch->observations && batch->actions && batch->log_probabilities &&
                     batch->advantages && batch->returns && batch->masks,
                 ALEGPU_INVALID_ARGUMENT, "train: null batch pointer");
  ALEGPU_CUDA_CHECK(cudaSetDevice(L->cfg.device));
  alegpu_train_options opt{};
  if (options)
    opt = *options;
  ALEGPU_REQUIRE(!opt.clip_value_loss || batch->values, ALEGPU_INVALID_ARGUMENT,
                 "train: clip_value_loss needs batch.values (the values stored at collection time)");
  if (int rc = learner_input_error(L))
    return rc;
  if (opt.indices) {
    if (int rc = ensure_staging(L))
      return rc;
    // every index is validated BEFORE anything trains on it (an index_select in the reference
    // would throw before the first optimiser step): one pass over the index vector + one 4-byte read
    if (int rc = launch_check_indices(opt.indices, num_epochs * batch->size, batch->size,
                                      L->gather_error, L->stream))
      return rc;
    int bad = 0;
    ALEGPU_CUDA_CHECK(cudaMemcpyAsync(&bad, L->gather_error, sizeof(int), cudaMemcpyDeviceToHost,
                                      L->stream));
    ALEGPU_CUDA_CHECK(cudaStreamSynchronize(L->stream));
    if (bad) {
      ALEGPU_CUDA_CHECK(cudaMemsetAsync(L->gather_error, 0, sizeof(int), L->stream));
      return fail(ALEGPU_OUT_OF_RANGE, "train: a minibatch index was outside [0, batch size)");
    }
  }
  if (use_tc(L) && !opt.indices)
    if (int rc = net_tc_reserve_obs(L->tc, batch->size))
      return rc;
  alegpu_metrics mt{};
  if (metrics)
    mt = *metrics;
  const int64_t before = g_launches;
  if (!(flags & 1)) {
    int rc = enqueue_train(L, *batch, mt, num_epochs, num_mini_batches, *hp, opt);
    L->last_launches = g_launches - before;
    return rc;
  }
  std::vector<uint64_t> key;
  key_push(key, batch->observations); key_push(key, batch->actions);
  key_push(key, batch->log_probabilities); key_push(key, batch->advantages);
  key_push(key, batch->returns); key_push(key, batch->masks); key_push(key, batch->size);
  key_push(key, mt.loss); key_push(key, mt.clipped_losses); key_push(key, mt.value_losses);
  key_push(key, mt.entropies); key_push(key, mt.ratio); key_push(key, mt.total_losses);
  key_push(key, mt.advantages); key_push(key, mt.returns); key_push(key, mt.masks);
  key_push(key, mt.clipped_gradients); key_push(key, num_epochs); key_push(key, num_mini_batches);
  key_push(key, hp->clip_param); key_push(key, hp->value_loss_coef);
  key_push(key, hp->entropy_coef); key_push(key, hp->max_gradient_norm);
  key_push(key, opt.indices); key_push(key, opt.normalize_advantages);
  key_push(key, opt.clip_value_loss); key_push(key, batch->values);
  if (!L->graph_exec || key != L->graph_key) {
    if (L->graph_exec) {
      cudaGraphExecDestroy(L->graph_exec);
      L->graph_exec = nullptr;
    }
    const int64_t step_before = L->host_step;
    ALEGPU_CUDA_CHECK(cudaStreamBeginCapture(L->stream, cudaStreamCaptureModeThreadLocal));
    int rc = enqueue_train(L, *batch, mt, num_epochs, num_mini_batches, *hp, opt);
    cudaGraph_t graph = nullptr;
    cudaError_t ce = cudaStreamEndCapture(L->stream, &graph);
    L->host_step = step_before; // capture enqueued nothing; the replay below advances it
    if (rc) {
      if (graph)
        cudaGraphDestroy(graph);
      return rc;
    }
    if (ce != cudaSuccess)
      return fail(ALEGPU_CUDA_ERROR, std::string("cudaStreamEndCapture: ") + cudaGetErrorString(ce));
    ce = cudaGraphInstantiate(&L->graph_exec, graph, 0);
    cudaGraphDestroy(graph);
    if (ce != cudaSuccess)
      return fail(ALEGPU_CUDA_ERROR, std::string("cudaGraphInstantiate: ") + cudaGetErrorString(ce));
    L->graph_key = key;
    L->graph_launches = g_launches - before;
  }
  ALEGPU_CUDA_CHECK(cudaGraphLaunch(L->graph_exec, L->stream));
  L->host_step += num_epochs * num_mini_batches;
  g_launches = before + L->graph_launches;
  L->last_launches = L->graph_launches;
  return ALEGPU_OK;
}

extern "C" int64_t alegpu_learner_last_launch_count(const alegpu_learner *L) {
  return L ? L->last_launches : 0;
}
