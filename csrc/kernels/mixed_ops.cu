// Glue kernels for the mixed-precision (bf16 tensor-core) dense path: masked fp32->bf16 casts and the
// bias gradient. The GEMMs themselves are gemm_tcgen05.cu; these keep ReLU-backward fused into the cast
// (G8: never a stand-alone elementwise pass) and feed the K-major operands tcgen05 wants.
#include "common.cuh"
#include <cuda_bf16.h>

namespace {

// y = bf16(x * (mask > 0)), 4 elements per thread where alignment allows
__global__ void cast_bf16_masked_kernel(const float* __restrict__ x, const float* __restrict__ mask,
                                        __nv_bfloat16* __restrict__ y, long n) {
  TFDE_PDL_PROLOGUE();
  const long n4 = n >> 2;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long)gridDim.x * blockDim.x) {
    float4 v = __ldg(reinterpret_cast<const float4*>(x) + i);
    if (mask != nullptr) {
      const float4 m = __ldg(reinterpret_cast<const float4*>(mask) + i);
      if (!(m.x > 0.f)) v.x = 0.f; if (!(m.y > 0.f)) v.y = 0.f;
      if (!(m.z > 0.f)) v.z = 0.f; if (!(m.w > 0.f)) v.w = 0.f;
    }
    __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y), b = __floats2bfloat162_rn(v.z, v.w);
    uint2 o; o.x = *reinterpret_cast<uint32_t*>(&a); o.y = *reinterpret_cast<uint32_t*>(&b);
    reinterpret_cast<uint2*>(y)[i] = o;
  }
  for (long i = n4 * 4 + (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
    float v = __ldg(x + i);
    if (mask != nullptr && !(__ldg(mask + i) > 0.f)) v = 0.f;
    y[i] = __float2bfloat16(v);
  }
}

// out[c] = sum_r x[r,c] * (mask[r,c] > 0). Grid = (column tiles of 32, row blocks); every block reduces its
// row range (32 columns x 8 row lanes), writes a partial, and the last block of a column tile adds the
// partials in row-block order (deterministic, same ticket scheme as the batch-norm statistics).
__global__ void __launch_bounds__(256)
colsum_masked_kernel(const float* __restrict__ x, const float* __restrict__ mask, float* __restrict__ out,
                     long R, int C, float* __restrict__ partials, unsigned* __restrict__ counters) {
  TFDE_PDL_PROLOGUE();
  __shared__ float s[8][33];
  __shared__ int s_last;
  const int cx = threadIdx.x & 31, ry = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + cx;
  float a = 0.f;
  if (c < C) {
    for (long r = (long)blockIdx.y * 8 + ry; r < R; r += (long)gridDim.y * 8) {
      float v = __ldg(x + r * C + c);
      if (mask != nullptr && !(__ldg(mask + r * C + c) > 0.f)) v = 0.f;
      a += v;
    }
  }
  s[ry][cx] = a;
  __syncthreads();
  float t = 0.f;
  if (ry == 0) {
#pragma unroll
    for (int i = 0; i < 8; ++i) t += s[i][cx];
  }
  if (gridDim.y == 1) {
    if (ry == 0 && c < C) out[c] = t;
    return;
  }
  float* part = partials + ((size_t)blockIdx.x * gridDim.y + blockIdx.y) * 32;
  if (ry == 0) __stcg(part + cx, t);
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_last = (atomicAdd(counters + blockIdx.x, 1u) == gridDim.y - 1);
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  if (ry == 0 && c < C) {
    float tot = 0.f;
    for (unsigned rb = 0; rb < gridDim.y; ++rb)
      tot += __ldcg(partials + ((size_t)blockIdx.x * gridDim.y + rb) * 32 + cx);
    out[c] = tot;
  }
  if (threadIdx.x == 0) counters[blockIdx.x] = 0u;
}

}  // namespace

TFDE_API int tfde_cast_bf16_masked(const float* x, const float* mask, void* y, long n, cudaStream_t stream) {
  long grid = (n / 4 + 255) / 256;
  if (grid > 148L * 8) grid = 148L * 8;
  if (grid < 1) grid = 1;
  tfde_launch(cast_bf16_masked_kernel, dim3((int)grid), dim3(256), 0, stream, x, mask, (__nv_bfloat16*)y, n);
  TFDE_RETURN_LAUNCH_STATUS();
}

// partials: >= ceil(C/32) * row_blocks * 32 floats (row_blocks <= 128); counters: ceil(C/32) zeroed tickets.
TFDE_API int tfde_colsum_masked(const float* x, const float* mask, float* out, long R, int C, float* partials,
                                unsigned* counters, cudaStream_t stream) {
  const int ct = (C + 31) / 32;
  if (ct > TFDE_MAX_COUNTERS) return TFDE_ERR_INVALID;
  long rb = (R + 8 * 16 - 1) / (8 * 16);                 // >= 16 rows per thread
  const long cap = (148L * 4 + ct - 1) / ct;
  if (rb > cap) rb = cap;
  if (rb > 128) rb = 128;
  if (rb < 1 || partials == nullptr || counters == nullptr) rb = 1;
  tfde_launch(colsum_masked_kernel, dim3(ct, (unsigned)rb), dim3(256), 0, stream, x, mask, out, R, C, partials, counters);
  TFDE_RETURN_LAUNCH_STATUS();
}
