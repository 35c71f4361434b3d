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
