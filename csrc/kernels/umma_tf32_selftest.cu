// Self-test of the shared-memory-operand TF32 tcgen05 path (umma_tf32.cuh): the three GEMM forms of one CTA of the
// whole-step kernel, on one 64x32 activation tile, one 32x64 weight slice and one 64x64 gradient tile:
//   D1[b, n] = sum_c act[b, c] * W[c, n]
//   D2[c, n] = sum_b act[b, c] * dh[b, n]       (computed as D2^T[n, c])
//   D3[b, c] = sum_n dh[b, n] * W[c, n]
// every operand K-major (MN-major SWIZZLE_NONE tf32 operands produced all-zero accumulators on this part during
// bring-up, so each tile that is needed in both orientations is simply written twice)
// Used by tests/test_umma_tf32_gpu.py against an fp32 matmul (TF32 tolerance).
#include "common.cuh"
#include "umma_tf32.cuh"

namespace {
__global__ void __launch_bounds__(256)
umma_tf32_selftest_kernel(const float* __restrict__ act, const float* __restrict__ W, const float* __restrict__ dh,
                          float* __restrict__ D1, float* __restrict__ D2, float* __restrict__ D3, int variant) {
  extern __shared__ __align__(128) float sm[];
  // all operands K-major (rows = the GEMM's M or N index, 4 consecutive K values per 16 bytes)
  float* actK = sm;                 // [b 64][c 32]  RB 256   A of D1          (2048 floats)
  float* wtK = sm + 2048;           // [n 64][c 32]  RB 256   B of D1
  float* actB = sm + 4096;          // [c 32][b 64]  RB 512   B of D2
  float* wsB = sm + 6144;           // [c 32][n 64]  RB 512   B of D3
  float* dhB = sm + 8192;           // [b 64][n 64]  RB 512   A of D3          (4096 floats)
  float* dhT = sm + 12288;          // [n 64][b 64]  RB 512   A of D2          (4096 floats, + 4096 of slack after it)
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  for (int i = t; i < 64 * 32; i += 256) {
    const int b = i >> 5, c = i & 31;
    actK[umma::blk(b, c, 256)] = act[i]; actB[umma::blk(c, b, 512)] = act[i];
  }
  for (int i = t; i < 32 * 64; i += 256) {
    const int c = i >> 6, n = i & 63;
    wsB[umma::blk(c, n, 512)] = W[i]; wtK[umma::blk(n, c, 256)] = W[i];
  }
  for (int i = t; i < 64 * 64; i += 256) {
    const int b = i >> 6, n = i & 63;
    dhB[umma::blk(b, n, 512)] = dh[i]; dhT[umma::blk(n, b, 512)] = dh[i];
  }
  for (int i = t; i < 4096; i += 256) sm[16384 + i] = 0.f;
  if (warp == 0) umma::tmem_alloc(&tmem_slot, 128);
  if (t == 0) umma::mbar_init(&bar, 1);
  umma::fence_operands();
  umma::fence_before_sync();
  __syncthreads();
  umma::fence_after_sync();
  const uint32_t tmem = tmem_slot;
  if (t == 0) {
    const uint32_t i64 = umma::idesc_tf32(128, 64, 0, 0), i32 = umma::idesc_tf32(128, 32, 0, 0);
    (void)variant;
    for (int k = 0; k < 4; ++k)     // D1: K = c (32)
      umma::mma_tf32(tmem + 0, umma::desc(umma::smem_u32(actK) + 2 * k * 128, 128, 1024),
                     umma::desc(umma::smem_u32(wtK) + 2 * k * 128, 128, 1024), i64, k > 0);
    for (int k = 0; k < 8; ++k)     // D2^T[n, c]: K = b (64)
      umma::mma_tf32(tmem + 64, umma::desc(umma::smem_u32(dhT) + 2 * k * 128, 128, 2048),
                     umma::desc(umma::smem_u32(actB) + 2 * k * 128, 128, 2048), i32, k > 0);
    for (int k = 0; k < 8; ++k)     // D3[b, c]: K = n (64)
      umma::mma_tf32(tmem + 96, umma::desc(umma::smem_u32(dhB) + 2 * k * 128, 128, 2048),
                     umma::desc(umma::smem_u32(wsB) + 2 * k * 128, 128, 2048), i32, k > 0);
    umma::commit(&bar);
  }
  umma::mbar_wait(&bar, 0);
  umma::fence_after_sync();
  if (warp < 2) {
    const int row = warp * 32 + lane;
    const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16);
    float r[32];
    umma::tmem_ld32(taddr + 0, r);
    for (int j = 0; j < 32; ++j) D1[row * 64 + j] = r[j];
    umma::tmem_ld32(taddr + 32, r);
    for (int j = 0; j < 32; ++j) D1[row * 64 + 32 + j] = r[j];
    umma::tmem_ld32(taddr + 64, r);          // D2^T: row = n, columns = c
    for (int j = 0; j < 32; ++j) D2[j * 64 + row] = r[j];
    umma::tmem_ld32(taddr + 96, r);          // D3: row = b, columns = c
    for (int j = 0; j < 32; ++j) D3[row * 32 + j] = r[j];
  }
  umma::fence_before_sync();
  __syncthreads();
  if (warp == 0) { umma::fence_after_sync(); umma::tmem_dealloc(tmem, 128); }
}
}  // namespace

// act [64][32], W [32][64], dh [64][64] -> D1 [64][64], D2 [32][64], D3 [64][32] (device pointers, fp32).
TFDE_API int tfde_umma_tf32_selftest(const float* act, const float* W, const float* dh, float* D1, float* D2, float* D3,
                                     int variant, cudaStream_t stream) {
  const size_t smem = (16384 + 4096) * sizeof(float) + 1024;
  static size_t done[TFDE_MAX_DEVICES] = {};
  const cudaError_t e = tfde_ensure_dyn_smem(umma_tf32_selftest_kernel, smem, done);
  if (e != cudaSuccess) return (int)e;
  umma_tf32_selftest_kernel<<<1, 256, smem, stream>>>(act, W, dh, D1, D2, D3, variant);
  TFDE_RETURN_LAUNCH_STATUS();
}
