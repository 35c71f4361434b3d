// G16: parameter-server push / pull over NVLink peer memory -- the B200-native form of the gRPC
// RecvTensor traffic that tf's ParameterServerStrategy generates per worker step
// (mnist_keras_distributed.py:241-242, tf2_mnist_distributed.py:189).
//
// The owner ("ps") GPU holds the authoritative fp32 shard in symmetric memory; a worker
//   push:  adds its scaled gradient into the owner's shard with reduction atomics
//          (atomicAdd with unused result -> RED.ADD.F32 over NVLink), no barrier, no lock: updates from
//          different workers interleave exactly like TF's unsynchronised async SGD;
//   pull:  copies the owner's current shard into its local weights with plain peer loads.
// Both are one kernel each and fully asynchronous on the caller's stream.
#include "common.cuh"

namespace {
__global__ void ps_push_kernel(float* __restrict__ w_owner, const float* __restrict__ g, long n, float neg_lr) {
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x)
    atomicAdd(w_owner + i, neg_lr * __ldg(g + i));
}
__global__ void ps_pull_kernel(float* __restrict__ w_local, const float* w_owner, long n) {
  const long n4 = n >> 2;
  const float4* s4 = reinterpret_cast<const float4*>(w_owner);
  float4* d4 = reinterpret_cast<float4*>(w_local);
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long)gridDim.x * blockDim.x)
    d4[i] = __ldcv(s4 + i);            // volatile: always fetch the owner's latest value
  for (long i = (n4 << 2) + (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x)
    w_local[i] = __ldcv(w_owner + i);
}
int ps_grid(long n) {
  long b = (n + 255) / 256;
  if (b > 148L * 4) b = 148L * 4;
  return (int)(b < 1 ? 1 : b);
}
}  // namespace

// w_owner[i] -= lr * g[i] for a shard of n floats that lives on the owner GPU (peer-mapped pointer).
TFDE_API int tfde_ps_push(float* w_owner, const float* g, long n, float lr, cudaStream_t stream) {
  ps_push_kernel<<<ps_grid(n), 256, 0, stream>>>(w_owner, g, n, -lr);
  TFDE_RETURN_LAUNCH_STATUS();
}
// w_local[i] = w_owner[i]; both 16-byte aligned.
TFDE_API int tfde_ps_pull(float* w_local, const float* w_owner, long n, cudaStream_t stream) {
  ps_pull_kernel<<<ps_grid(n), 256, 0, stream>>>(w_local, w_owner, n);
  TFDE_RETURN_LAUNCH_STATUS();
}
