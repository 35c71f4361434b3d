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
// One launch for a whole model: a table of segments {owner pointer, local pointer, n, kind}, blockIdx.y = segment.
struct PsSeg { float* owner; float* local; long n; long kind; };      // kind 0: trainable, 1: non-trainable state
__global__ void ps_pull_multi_kernel(const PsSeg* __restrict__ table) {
  const PsSeg sg = table[blockIdx.y];
  const bool vec = (((uintptr_t)sg.owner | (uintptr_t)sg.local) & 15) == 0;
  const long n4 = vec ? (sg.n >> 2) : 0;
  const float4* s4 = reinterpret_cast<const float4*>(sg.owner);
  float4* d4 = reinterpret_cast<float4*>(sg.local);
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long)gridDim.x * blockDim.x) d4[i] = __ldcv(s4 + i);
  for (long i = (n4 << 2) + (long)blockIdx.x * blockDim.x + threadIdx.x; i < sg.n; i += (long)gridDim.x * blockDim.x)
    sg.local[i] = __ldcv(sg.owner + i);
}
// trainable segments: owner -= lr * g (reduction atomics over NVLink, unsynchronised between workers);
// non-trainable segments (batch-norm moving statistics): the worker's value is stored (last writer wins, as in TF).
// Block (0,0) then bumps the owner-side global_step (a device counter on ps 0) and reports the new value.
__global__ void ps_push_multi_kernel(const PsSeg* __restrict__ table, float neg_lr, unsigned* step_owner,
                                     unsigned* __restrict__ step_out) {
  const PsSeg sg = table[blockIdx.y];
  if (sg.kind == 0) {
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < sg.n; i += (long)gridDim.x * blockDim.x)
      atomicAdd(sg.owner + i, neg_lr * __ldg(sg.local + i));
  } else {
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < sg.n; i += (long)gridDim.x * blockDim.x)
      sg.owner[i] = __ldg(sg.local + i);
  }
  if (step_owner != nullptr && blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) {
    const unsigned now = atomicAdd_system(step_owner, 1u) + 1u;
    if (step_out) *step_out = now;
  }
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

struct TfdePsSeg { void* owner; void* local; long n; long kind; };
TFDE_API size_t tfde_ps_seg_bytes() { return sizeof(TfdePsSeg); }
// table: device array of nseg TfdePsSeg; max_n: the longest segment (sizes the grid).
TFDE_API int tfde_ps_pull_multi(const void* table, int nseg, long max_n, cudaStream_t stream) {
  if (nseg < 1) return 0;
  int gx = ps_grid(max_n); if (gx > 64) gx = 64;
  ps_pull_multi_kernel<<<dim3(gx, nseg), 256, 0, stream>>>((const PsSeg*)table);
  TFDE_RETURN_LAUNCH_STATUS();
}
// table entries of kind 0 name the LOCAL GRADIENT as `local`; kind 1 entries name the local state value.
TFDE_API int tfde_ps_push_multi(const void* table, int nseg, long max_n, float lr, unsigned* step_owner,
                                unsigned* step_out, cudaStream_t stream) {
  if (nseg < 1) return 0;
  int gx = ps_grid(max_n); if (gx > 64) gx = 64;
  ps_push_multi_kernel<<<dim3(gx, nseg), 256, 0, stream>>>((const PsSeg*)table, -lr, step_owner, step_out);
  TFDE_RETURN_LAUNCH_STATUS();
}
