// Fused first layer of "Model A" (distributed_with_keras.py:34-35):
//   Conv2D(32, 3, activation='relu', input_shape=(28,28,1)) -> MaxPooling2D()
// Forward writes only the pooled activation [B,13,13,32] and a 2-bit arg-max per element (G1+G6 in
// SURVEY.md section 2.7); the 26x26x32 conv output (5.5 MB at B=64) never exists in memory.
// Backward fuses pool routing + ReLU mask + conv filter/bias gradient (G2+G6+G8); as this is the
// first layer there is no data gradient.
//
// Mapping: Cin == 1 makes K = 9, far below any tensor-core tile and below TMA's 16-byte inner
// box rule, so this is a CUDA-core kernel: one warp per pooled pixel, lane == output channel
// (Cout must be 32). The 4x4 input patch is read once by 16 lanes and broadcast with shuffles.
// The input may be uint8 (raw MNIST) -- the /255 scaling of distributed_with_keras.py:20-23 is
// folded into the load (G15).
#include "common.cuh"

namespace {

constexpr int WARPS = 8;

// IT: pixel-index type. `long` at the batch-64 operating point (measured 1.2 us/step FASTER there than the
// 32-bit variant on the same box: the compiler keeps a deeper load pipeline), `unsigned` for large batches where
// the emulated 64-bit div/mod dominates the instruction count (batch 8192: 1.49 -> 1.21 ms/step).
template <bool U8, typename IT>
__global__ void __launch_bounds__(WARPS * 32)
conv3x3c1_relu_pool_fwd_kernel(const void* __restrict__ xin, const float* __restrict__ w,
                               const float* __restrict__ bias, float* __restrict__ pooled,
                               uint8_t* __restrict__ idx, int Bn, int H, int W, int PH, int PW,
                               float in_scale) {
  TFDE_PDL_PROLOGUE();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float wr[9];
#pragma unroll
  for (int i = 0; i < 9; ++i) wr[i] = __ldg(w + i * 32 + lane);
  const float b = bias ? __ldg(bias + lane) : 0.f;
  // 32-bit index arithmetic (the host rejects Bn*PH*PW >= 2^31): emulated 64-bit div/mod was most of the
  // instruction count of this loop at large batch
  const IT total = (IT)Bn * (IT)PH * (IT)PW, uPW = (IT)PW, uPH = (IT)PH;
  for (IT p = (IT)blockIdx.x * WARPS + warp; p < total; p += (IT)gridDim.x * WARPS) {
    const IT t = p / uPW, pw = p - t * uPW;
    const IT bi = t / uPH, ph = t - bi * uPH;
    float v = 0.f;
    if (lane < 16) {
      const size_t off = ((size_t)bi * H + (2 * ph + (lane >> 2))) * W + (2 * pw + (lane & 3));
      if (U8) v = (float)__ldg(reinterpret_cast<const uint8_t*>(xin) + off) * in_scale;
      else    v = __ldg(reinterpret_cast<const float*>(xin) + off);
    }
    float patch[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) patch[i] = __shfl_sync(0xffffffffu, v, i);
    float best = -1.f;
    int besti = 0;
#pragma unroll
    for (int dy = 0; dy < 2; ++dy)
#pragma unroll
      for (int dx = 0; dx < 2; ++dx) {
        float o = b;
#pragma unroll
        for (int kh = 0; kh < 3; ++kh)
#pragma unroll
          for (int kw = 0; kw < 3; ++kw) o = fmaf(wr[kh * 3 + kw], patch[(dy + kh) * 4 + dx + kw], o);
        o = fmaxf(o, 0.f);
        if (o > best) { best = o; besti = dy * 2 + dx; }
      }
    pooled[(size_t)p * 32 + lane] = best;
    idx[(size_t)p * 32 + lane] = (uint8_t)besti;
  }
}

// Partials layout: [gridDim.x][320] floats (288 filter taps, then 32 bias entries).
template <bool U8, typename IT>
__global__ void __launch_bounds__(WARPS * 32)
conv3x3c1_relu_pool_bwd_kernel(const void* __restrict__ xin, const float* __restrict__ pooled,
                               const uint8_t* __restrict__ idx, const float* __restrict__ dpooled,
                               float* __restrict__ dw, float* __restrict__ db,
                               float* __restrict__ partials, unsigned* __restrict__ counter,
                               int Bn, int H, int W, int PH, int PW, float in_scale) {
  TFDE_PDL_PROLOGUE();
  __shared__ float s_patch[WARPS][16];
  __shared__ __align__(16) float s_red[WARPS][10][32];
  __shared__ int s_last;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float acc[10];
#pragma unroll
  for (int i = 0; i < 10; ++i) acc[i] = 0.f;
  const IT total = (IT)Bn * (IT)PH * (IT)PW, uPW = (IT)PW, uPH = (IT)PH;
  const IT stride = (IT)gridDim.x * WARPS;
  // Software-pipelined: the loads of pixel p+stride are issued before pixel p is consumed.
  auto fetch = [&](IT p, float& xv, float& g, int& bsel) {
    xv = 0.f; g = 0.f; bsel = 0;
    if (p >= total) return;
    const IT t = p / uPW, pw = p - t * uPW;
    const IT bi = t / uPH, ph = t - bi * uPH;
    if (lane < 16) {
      const size_t off = ((size_t)bi * H + (2 * ph + (lane >> 2))) * W + (2 * pw + (lane & 3));
      if (U8) xv = (float)__ldg(reinterpret_cast<const uint8_t*>(xin) + off) * in_scale;
      else    xv = __ldg(reinterpret_cast<const float*>(xin) + off);
    }
    g = __ldg(dpooled + (size_t)p * 32 + lane);
    if (!(__ldg(pooled + (size_t)p * 32 + lane) > 0.f)) g = 0.f;
    bsel = idx[(size_t)p * 32 + lane];
  };
  float xv_n, g_n; int b_n;
  IT p = (IT)blockIdx.x * WARPS + warp;
  fetch(p, xv_n, g_n, b_n);
  for (; p < total; p += stride) {
    const float xv = xv_n, g = g_n; const int bsel = b_n;
    fetch(p + stride, xv_n, g_n, b_n);
    __syncwarp();
    if (lane < 16) s_patch[warp][lane] = xv;
    __syncwarp();
    const int dy = bsel >> 1, dx = bsel & 1;
#pragma unroll
    for (int kh = 0; kh < 3; ++kh)
#pragma unroll
      for (int kw = 0; kw < 3; ++kw)
        acc[kh * 3 + kw] = fmaf(g, s_patch[warp][(dy + kh) * 4 + dx + kw], acc[kh * 3 + kw]);
    acc[9] += g;
  }
#pragma unroll
  for (int i = 0; i < 10; ++i) s_red[warp][i][lane] = acc[i];
  __syncthreads();
  // fixed-order reductions => bit-reproducible gradients
  for (int e = threadIdx.x; e < 320; e += blockDim.x) {
    const int i = e >> 5, c = e & 31;
    float s = 0.f;
#pragma unroll
    for (int wv = 0; wv < WARPS; ++wv) s += s_red[wv][i][c];
    __stcg(partials + (size_t)blockIdx.x * 320 + e, s);
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_last = (atomicAdd(counter, 1u) == gridDim.x - 1);
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  // Final reduction over the per-block partials: 80 float4 columns x 3 interleaved row groups (240
  // threads), 8 independent 16-byte L2 loads in flight per thread; groups are then combined in a fixed
  // order, so the result does not depend on which block arrived last.
  float4* s_grp = reinterpret_cast<float4*>(&s_red[0][0][0]);   // [3][80] float4, re-uses s_red
  __syncthreads();
  if (threadIdx.x < 240) {
    const unsigned grp = threadIdx.x / 80, c4 = threadIdx.x % 80;
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    for (unsigned b0 = grp; b0 < gridDim.x; b0 += 24) {
      float4 v[8];
#pragma unroll
      for (unsigned j = 0; j < 8; ++j) {
        const unsigned blk = b0 + 3 * j;
        v[j] = blk < gridDim.x ? __ldcg(reinterpret_cast<const float4*>(partials + (size_t)blk * 320) + c4)
                               : make_float4(0.f, 0.f, 0.f, 0.f);
      }
#pragma unroll
      for (unsigned j = 0; j < 8; ++j) { s.x += v[j].x; s.y += v[j].y; s.z += v[j].z; s.w += v[j].w; }
    }
    s_grp[grp * 80 + c4] = s;
  }
  __syncthreads();
  if (threadIdx.x < 80) {
    const float4 a = s_grp[threadIdx.x], b = s_grp[80 + threadIdx.x], c = s_grp[160 + threadIdx.x];
    const float4 s = make_float4((a.x + b.x) + c.x, (a.y + b.y) + c.y, (a.z + b.z) + c.z, (a.w + b.w) + c.w);
    const int e = threadIdx.x * 4;   // [kh][kw][ci=0][co] == e = (kh*3+kw)*32 + co ; e >= 288 is the bias
    if (e < 288) *reinterpret_cast<float4*>(dw + e) = s;
    else if (db != nullptr) *reinterpret_cast<float4*>(db + (e - 288)) = s;
  }
  if (threadIdx.x == 0) *counter = 0u;
}

int grid_for(long total_pixels) {
  long blocks = (total_pixels + WARPS - 1) / WARPS;
  const long cap = 148L * 4;
  return (int)(blocks < cap ? (blocks < 1 ? 1 : blocks) : cap);
}

}  // namespace

TFDE_API int tfde_conv3x3c1_relu_pool_fwd(const void* x, int x_is_u8, float in_scale,
                                          const float* w, const float* bias, float* pooled,
                                          uint8_t* idx, int Bn, int H, int W, int Cout,
                                          cudaStream_t stream) {
  if (Cout != 32 || H < 4 || W < 4) return TFDE_ERR_INVALID;
  const int PH = (H - 2) / 2, PW = (W - 2) / 2;
  if ((long)Bn * PH * PW >= (1L << 31) - 148L * 64 * WARPS) return TFDE_ERR_INVALID;           // 32-bit pixel index
  const int grid = grid_for((long)Bn * PH * PW);
  const bool big = (long)Bn * PH * PW > 200000L;      // see the IT note on the kernel
  if (x_is_u8 && big)
    tfde_launch(conv3x3c1_relu_pool_fwd_kernel<true, unsigned>, dim3(grid), dim3(WARPS * 32), 0, stream, x, w, bias, pooled, idx, Bn, H, W, PH, PW, in_scale);
  else if (x_is_u8)
    tfde_launch(conv3x3c1_relu_pool_fwd_kernel<true, long>, dim3(grid), dim3(WARPS * 32), 0, stream, x, w, bias, pooled, idx, Bn, H, W, PH, PW, in_scale);
  else if (big)
    tfde_launch(conv3x3c1_relu_pool_fwd_kernel<false, unsigned>, dim3(grid), dim3(WARPS * 32), 0, stream, x, w, bias, pooled, idx, Bn, H, W, PH, PW, 1.f);
  else
    tfde_launch(conv3x3c1_relu_pool_fwd_kernel<false, long>, dim3(grid), dim3(WARPS * 32), 0, stream, x, w, bias, pooled, idx, Bn, H, W, PH, PW, 1.f);
  TFDE_RETURN_LAUNCH_STATUS();
}

// Number of floats of scratch the backward needs (partials), independent of the batch size.
TFDE_API size_t tfde_conv3x3c1_relu_pool_bwd_ws_floats() { return (size_t)148 * 4 * 320; }

TFDE_API int tfde_conv3x3c1_relu_pool_bwd(const void* x, int x_is_u8, float in_scale,
                                          const float* pooled, const uint8_t* idx,
                                          const float* dpooled, float* dw, float* db,
                                          float* partials, unsigned* counter, int Bn, int H, int W,
                                          int Cout, cudaStream_t stream) {
  if (Cout != 32 || H < 4 || W < 4) return TFDE_ERR_INVALID;
  const int PH = (H - 2) / 2, PW = (W - 2) / 2;
  long blocks = ((long)Bn * PH * PW + WARPS * 4 - 1) / (WARPS * 4);  // >= 4 pixels per warp
  if (blocks < 1) blocks = 1;
  // latency-bound streaming reduction: large batches get 4 resident blocks per SM (the partials buffer holds
  // 592 rows); the batch-64 operating point keeps 2 (shorter last-block finalise)
  long cap = ((long)Bn * PH * PW > 200000L) ? 148L * 4 : 148L * 2;
  {
    static long forced = -1;     // TFDE_CONVBWD_BLOCKS=<n> (1..592) overrides the cap (tuning experiments)
    if (forced < 0) { const char* e = getenv("TFDE_CONVBWD_BLOCKS"); forced = e ? atol(e) : 0; }
    if (forced >= 1 && forced <= 148L * 4) cap = forced;
  }
  if (blocks > cap) blocks = cap;
  if ((long)Bn * PH * PW >= (1L << 31) - 148L * 4 * WARPS) return TFDE_ERR_INVALID;   // 32-bit pixel index
  const int grid = (int)blocks;
  const bool big = (long)Bn * PH * PW > 200000L;
  if (x_is_u8 && big)
    tfde_launch(conv3x3c1_relu_pool_bwd_kernel<true, unsigned>, dim3(grid), dim3(WARPS * 32), 0, stream, x, pooled, idx, dpooled, dw, db, partials, counter, Bn, H, W, PH, PW, in_scale);
  else if (x_is_u8)
    tfde_launch(conv3x3c1_relu_pool_bwd_kernel<true, long>, dim3(grid), dim3(WARPS * 32), 0, stream, x, pooled, idx, dpooled, dw, db, partials, counter, Bn, H, W, PH, PW, in_scale);
  else if (big)
    tfde_launch(conv3x3c1_relu_pool_bwd_kernel<false, unsigned>, dim3(grid), dim3(WARPS * 32), 0, stream, x, pooled, idx, dpooled, dw, db, partials, counter, Bn, H, W, PH, PW, 1.f);
  else
    tfde_launch(conv3x3c1_relu_pool_bwd_kernel<false, long>, dim3(grid), dim3(WARPS * 32), 0, stream, x, pooled, idx, dpooled, dw, db, partials, counter, Bn, H, W, PH, PW, 1.f);
  TFDE_RETURN_LAUNCH_STATUS();
}
