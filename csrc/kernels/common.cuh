// Shared device helpers for the sm_100a kernels of tfde_b200.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 (see build.py).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define TFDE_API extern "C" __attribute__((visibility("default")))

// Every entry point returns a cudaError_t as int (0 == success). Launch errors are fetched with
// cudaPeekAtLastError so a sticky error from an earlier asynchronous failure is not cleared.
#define TFDE_RETURN_LAUNCH_STATUS() return (int)cudaPeekAtLastError()

#define TFDE_ERR_INVALID 1  // cudaErrorInvalidValue

// Size (in unsigned ints) of the self-cleaning ticket buffer every "last block finalises" kernel shares.
constexpr int TFDE_MAX_COUNTERS = 4096;

// ---------------------------------------------------------------------------------------------
// Programmatic dependent launch (PDL). Every step kernel starts with
//   pdl_launch_dependents();   let the NEXT kernel of the stream be scheduled right away ...
//   pdl_wait();                ... and block here until the PREVIOUS kernel has completed and flushed.
// so consecutive kernels of the captured step overlap their launch latency and prologue (2-3 us per edge
// on a step that is ~9 us per kernel) while every data dependency stays exactly as strict as before.
// Both instructions are no-ops when the kernel was not launched with the PDL attribute.
// ---------------------------------------------------------------------------------------------
static __device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
static __device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
#define TFDE_PDL_PROLOGUE() do { pdl_launch_dependents(); pdl_wait(); } while (0)

#include <cstdlib>
#include <utility>
static inline bool tfde_pdl_enabled() {
  static int cached = -1;
  if (cached < 0) { const char* e = getenv("TFDE_PDL"); cached = (e == nullptr || e[0] != '0') ? 1 : 0; }
  return cached == 1;
}
// <<<>>> replacement that adds the programmatic-stream-serialization attribute (TFDE_PDL=0 disables).
template <typename... KArgs, typename... Args>
static inline cudaError_t tfde_launch(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                                      Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  const bool on = tfde_pdl_enabled();
  cfg.attrs = on ? attr : nullptr;
  cfg.numAttrs = on ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) applies to the CURRENT device only, so the "already raised" note a
// launcher keeps must be per device (a process may drive several GPUs, e.g. mirrored evaluation). `done` is the
// launcher's own static array of TFDE_MAX_DEVICES entries.
constexpr int TFDE_MAX_DEVICES = 64;
template <typename K>
static inline cudaError_t tfde_ensure_dyn_smem(K kernel, size_t smem, size_t (&done)[TFDE_MAX_DEVICES]) {
  if (smem <= 48 * 1024) return cudaSuccess;
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  size_t& d = done[dev & (TFDE_MAX_DEVICES - 1)];
  if (smem > d) {
    e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    d = smem;
  }
  return cudaSuccess;
}

static __device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
static __device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// Block-wide sum for blockDim.x <= 1024 (multiple of 32). `red` is >= 32 floats of shared memory.
static __device__ __forceinline__ float block_sum(float v, float* red) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  v = warp_sum(v);
  __syncthreads();  // protect `red` from the previous use
  if (lane == 0) red[wid] = v;
  __syncthreads();
  const int nw = (blockDim.x + 31) >> 5;
  v = (threadIdx.x < nw) ? red[threadIdx.x] : 0.f;
  if (wid == 0) v = warp_sum(v);
  if (threadIdx.x == 0) red[0] = v;
  __syncthreads();
  return red[0];
}

// ---------------------------------------------------------------------------------------------
// Philox4x32-10, identical to ops/reference.py::philox4x32_safe.
// counter = (idx/4, step, stream, 0), key = (seed_lo, seed_hi).
// ---------------------------------------------------------------------------------------------
struct Philox4 {
  uint32_t v[4];
};
static __device__ __forceinline__ Philox4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2,
                                                        uint32_t c3, uint32_t k0, uint32_t k1) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    const uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
    c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  Philox4 out;
  out.v[0] = c0; out.v[1] = c1; out.v[2] = c2; out.v[3] = c3;
  return out;
}

// Device-resident hyper-parameter / state block shared by all kernels of one model so that a
// captured CUDA graph keeps working while the host changes the learning rate or the step advances.
// Layout mirrored in ops/native.py (HPARAMS_*). 16 floats.
struct TfdeHParams {
  float lr;        // 0
  float momentum;  // 1
  float beta1;     // 2
  float beta2;     // 3
  float eps;       // 4
  float grad_scale;  // 5  multiplied into the reduced gradient (1 when the loss is pre-scaled)
  float beta1_pow;   // 6  beta1^t, maintained on device by the optimizer kernel
  float beta2_pow;   // 7
  uint32_t step;     // 8  optimizer steps applied so far (device-side global_step)
  uint32_t seed_lo;  // 9
  uint32_t seed_hi;  // 10
  uint32_t pad[5];
};
