// Small diagnostics callable through the C ABI.
#include "common.cuh"

// 0 = not capturing, 1 = capture active, 2 = capture invalidated, negative = query failed.
TFDE_API int tfde_capture_status(cudaStream_t stream) {
  cudaStreamCaptureStatus st;
  const cudaError_t e = cudaStreamIsCapturing(stream, &st);
  if (e != cudaSuccess) { cudaGetLastError(); return -(int)e; }
  return st == cudaStreamCaptureStatusNone ? 0 : (st == cudaStreamCaptureStatusActive ? 1 : 2);
}
// Clears and returns the calling thread's last CUDA runtime error of this library's runtime instance.
TFDE_API int tfde_clear_last_error() { return (int)cudaGetLastError(); }

// dst[i] = src[i] for a handful of floats (n <= 1024). Used as the last node of the step graph to drop the
// 4-float metrics tail into device-mapped pinned host memory: a 2 us kernel node instead of a ~6 us memcpy
// node, and the host reads it after waiting on the step's event.
__global__ void copy_small_kernel(float* __restrict__ dst, const float* __restrict__ src, int n) {
  TFDE_PDL_PROLOGUE();
  for (int i = threadIdx.x; i < n; i += blockDim.x) dst[i] = src[i];
}
TFDE_API int tfde_copy_small(float* dst, const float* src, int n, cudaStream_t stream) {
  if (n < 0 || n > 1024) return TFDE_ERR_INVALID;
  tfde_launch(copy_small_kernel, dim3(1), dim3(64), 0, stream, dst, src, n);
  TFDE_RETURN_LAUNCH_STATUS();
}
