// net_tc.cu — placeholder until the tcgen05 kernels land (next commit).
#include "net_tc.h"
namespace alegpu {
static int nyi() { return fail(ALEGPU_RUNTIME_ERROR, "tensor-core path not built yet"); }
int net_tc_init(TcState &, int64_t, int64_t, int64_t, int64_t, cudaStream_t) { return nyi(); }
void net_tc_destroy(TcState &) {}
int net_tc_refresh_weights(TcState &, const float *, int64_t, cudaStream_t) { return nyi(); }
int net_tc_forward(const NetParams &, TcState &, const uint8_t *, int64_t, NetActs &, cudaStream_t) { return nyi(); }
int net_tc_backward(const NetParams &, const NetParams &, TcState &, const uint8_t *, int64_t, NetActs &, float *, size_t, cudaStream_t) { return nyi(); }
int net_tc_fetch_activation(TcState &, int, int64_t, float *, cudaStream_t) { return nyi(); }
} // namespace alegpu
