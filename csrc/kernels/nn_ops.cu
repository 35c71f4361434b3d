// Normalisation / pooling / activation / dropout kernels (channels-last), sm_100a.
// Covers SURVEY.md section 2.7 rows G6 (generic max-pool), G7 (batch-norm train/infer/backward, fused
// with ReLU), G8 (stand-alone Activation layer, mnist_keras_distributed.py:88), G10 (Philox dropout,
// mnist_keras_distributed.py:106), plus softmax, residual add and global average pooling for the
// ResNet-50 config. All of these are bandwidth-bound: one pass over the data each, float4 where the
// channel count allows, deterministic fixed-order reductions.
#include "common.cuh"

namespace {

// ------------------------------------------------------------------------------------------
// Batch normalisation. x viewed as [M, C]. Block = 32 (channels) x 8 (rows); grid = (ctiles, RB).
// Phase 1 writes per-row-block partial sums; the last block per channel tile finalises.
// ------------------------------------------------------------------------------------------
constexpr int BN_ROWS = 8;

__global__ void __launch_bounds__(32 * BN_ROWS)
bn_stats_kernel(const float* __restrict__ x, long M, int C, float* __restrict__ partials,
                unsigned* __restrict__ counters, float* __restrict__ save_mean,
                float* __restrict__ save_invstd, float* __restrict__ moving_mean,
                float* __restrict__ moving_var, float momentum, float eps, int unbiased_moving) {
  __shared__ float s_a[BN_ROWS][32], s_b[BN_ROWS][32];
  __shared__ int s_last;
  const int lane = threadIdx.x, wy = threadIdx.y;
  const int c = blockIdx.x * 32 + lane;
  float s = 0.f, q = 0.f;
  if (c < C) {
    for (long m = (long)blockIdx.y * BN_ROWS + wy; m < M; m += (long)gridDim.y * BN_ROWS) {
      const float v = __ldg(x + m * C + c);
      s += v; q = fmaf(v, v, q);
    }
  }
  s_a[wy][lane] = s; s_b[wy][lane] = q;
  __syncthreads();
  float* part = partials + ((size_t)blockIdx.x * gridDim.y + blockIdx.y) * 64;
  if (wy == 0) {
#pragma unroll
    for (int r = 1; r < BN_ROWS; ++r) { s += s_a[r][lane]; q += s_b[r][lane]; }
    __stcg(part + lane, s); __stcg(part + 32 + lane, q);
  }
  __threadfence();
  __syncthreads();
  if (lane == 0 && wy == 0) s_last = (atomicAdd(counters + blockIdx.x, 1u) == gridDim.y - 1);
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  if (wy == 0 && c < C) {
    double ds = 0.0, dq = 0.0;
    for (unsigned rb = 0; rb < gridDim.y; ++rb) {
      const float* pp = partials + ((size_t)blockIdx.x * gridDim.y + rb) * 64;
      ds += (double)__ldcg(pp + lane); dq += (double)__ldcg(pp + 32 + lane);
    }
    const double mean = ds / (double)M;
    double var = dq / (double)M - mean * mean;
    if (var < 0.0) var = 0.0;
    save_mean[c] = (float)mean;
    save_invstd[c] = rsqrtf((float)var + eps);
    const double vm = (unbiased_moving && M > 1) ? var * ((double)M / (double)(M - 1)) : var;
    moving_mean[c] = moving_mean[c] * momentum + (float)mean * (1.f - momentum);
    moving_var[c] = moving_var[c] * momentum + (float)vm * (1.f - momentum);
  }
  if (lane == 0 && wy == 0) counters[blockIdx.x] = 0u;
}

// y = ((x - mean) * invstd) * gamma + beta, optional ReLU. `stat_is_var`: second stat is a variance
// (inference with moving statistics) instead of an inverse standard deviation.
__device__ __forceinline__ float bn_apply_one(float xv, float mu, float inv, const float* gamma, const float* beta,
                                              int c, int relu) {
  float v = (xv - mu) * inv;
  if (gamma) v *= __ldg(gamma + c);
  if (beta) v += __ldg(beta + c);
  return relu ? fmaxf(v, 0.f) : v;
}
__global__ void bn_apply_kernel(const float* __restrict__ x, float* __restrict__ y, long total,
                                int C, const float* __restrict__ mean,
                                const float* __restrict__ stat, int stat_is_var, float eps,
                                const float* __restrict__ gamma, const float* __restrict__ beta,
                                int relu) {
  if ((C & 3) == 0) {   // 16-byte path: 4 consecutive channels per thread
    const long n4 = total >> 2;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long)gridDim.x * blockDim.x) {
      const int c = (int)((i << 2) % C);
      const float4 xv = __ldg(reinterpret_cast<const float4*>(x) + i);
      const float4 mu = __ldg(reinterpret_cast<const float4*>(mean + c));
      float4 st = __ldg(reinterpret_cast<const float4*>(stat + c));
      if (stat_is_var) { st.x = rsqrtf(st.x + eps); st.y = rsqrtf(st.y + eps); st.z = rsqrtf(st.z + eps); st.w = rsqrtf(st.w + eps); }
      float4 o;
      o.x = bn_apply_one(xv.x, mu.x, st.x, gamma, beta, c, relu);
      o.y = bn_apply_one(xv.y, mu.y, st.y, gamma, beta, c + 1, relu);
      o.z = bn_apply_one(xv.z, mu.z, st.z, gamma, beta, c + 2, relu);
      o.w = bn_apply_one(xv.w, mu.w, st.w, gamma, beta, c + 3, relu);
      reinterpret_cast<float4*>(y)[i] = o;
    }
    return;
  }
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    const float inv = stat_is_var ? rsqrtf(__ldg(stat + c) + eps) : __ldg(stat + c);
    y[i] = bn_apply_one(__ldg(x + i), __ldg(mean + c), inv, gamma, beta, c, relu);
  }
}

// Backward phase 1: dbeta = sum dy', dgamma = sum dy' * xhat, with dy' = dy * (y > 0) when fused ReLU.
__global__ void __launch_bounds__(32 * BN_ROWS)
bn_bwd_reduce_kernel(const float* __restrict__ x, const float* __restrict__ y,
                     const float* __restrict__ dy, long M, int C, const float* __restrict__ mean,
                     const float* __restrict__ invstd, int relu, float* __restrict__ partials,
                     unsigned* __restrict__ counters, float* __restrict__ dbeta,
                     float* __restrict__ dgamma_full) {
  __shared__ float s_a[BN_ROWS][32], s_b[BN_ROWS][32];
  __shared__ int s_last;
  const int lane = threadIdx.x, wy = threadIdx.y;
  const int c = blockIdx.x * 32 + lane;
  float s = 0.f, q = 0.f;
  if (c < C) {
    const float mu = __ldg(mean + c), inv = __ldg(invstd + c);
    for (long m = (long)blockIdx.y * BN_ROWS + wy; m < M; m += (long)gridDim.y * BN_ROWS) {
      float g = __ldg(dy + m * C + c);
      if (relu && !(__ldg(y + m * C + c) > 0.f)) g = 0.f;
      s += g; q = fmaf(g, (__ldg(x + m * C + c) - mu) * inv, q);
    }
  }
  s_a[wy][lane] = s; s_b[wy][lane] = q;
  __syncthreads();
  float* part = partials + ((size_t)blockIdx.x * gridDim.y + blockIdx.y) * 64;
  if (wy == 0) {
#pragma unroll
    for (int r = 1; r < BN_ROWS; ++r) { s += s_a[r][lane]; q += s_b[r][lane]; }
    __stcg(part + lane, s); __stcg(part + 32 + lane, q);
  }
  __threadfence();
  __syncthreads();
  if (lane == 0 && wy == 0) s_last = (atomicAdd(counters + blockIdx.x, 1u) == gridDim.y - 1);
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  if (wy == 0 && c < C) {
    double ds = 0.0, dq = 0.0;
    for (unsigned rb = 0; rb < gridDim.y; ++rb) {
      const float* pp = partials + ((size_t)blockIdx.x * gridDim.y + rb) * 64;
      ds += (double)__ldcg(pp + lane); dq += (double)__ldcg(pp + 32 + lane);
    }
    dbeta[c] = (float)ds;
    dgamma_full[c] = (float)dq;
  }
  if (lane == 0 && wy == 0) counters[blockIdx.x] = 0u;
}

// Backward phase 2: dx = gamma*invstd/M * (M*dy' - dbeta - xhat*dgamma)
__device__ __forceinline__ float bn_bwd_one(float xv, float yv, float g, float mu, float inv, float gm, float dbt,
                                            float dgm, float fM, float invM, int relu) {
  if (relu && !(yv > 0.f)) g = 0.f;
  const float xhat = (xv - mu) * inv;
  return gm * inv * invM * (fM * g - dbt - xhat * dgm);
}
__global__ void bn_bwd_apply_kernel(const float* __restrict__ x, const float* __restrict__ y,
                                    const float* __restrict__ dy, float* __restrict__ dx,
                                    long total, long M, int C, const float* __restrict__ mean,
                                    const float* __restrict__ invstd,
                                    const float* __restrict__ gamma,
                                    const float* __restrict__ dbeta,
                                    const float* __restrict__ dgamma_full, int relu) {
  const float invM = 1.f / (float)M, fM = (float)M;
  if ((C & 3) == 0) {
    const long n4 = total >> 2;
    const float4 one4 = make_float4(1.f, 1.f, 1.f, 1.f);
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long)gridDim.x * blockDim.x) {
      const int c = (int)((i << 2) % C);
      const float4 xv = __ldg(reinterpret_cast<const float4*>(x) + i);
      const float4 yv = relu ? __ldg(reinterpret_cast<const float4*>(y) + i) : one4;
      const float4 g = __ldg(reinterpret_cast<const float4*>(dy) + i);
      const float4 mu = __ldg(reinterpret_cast<const float4*>(mean + c));
      const float4 iv = __ldg(reinterpret_cast<const float4*>(invstd + c));
      const float4 gm = gamma ? __ldg(reinterpret_cast<const float4*>(gamma + c)) : one4;
      const float4 db = __ldg(reinterpret_cast<const float4*>(dbeta + c));
      const float4 dg = __ldg(reinterpret_cast<const float4*>(dgamma_full + c));
      float4 o;
      o.x = bn_bwd_one(xv.x, yv.x, g.x, mu.x, iv.x, gm.x, db.x, dg.x, fM, invM, relu);
      o.y = bn_bwd_one(xv.y, yv.y, g.y, mu.y, iv.y, gm.y, db.y, dg.y, fM, invM, relu);
      o.z = bn_bwd_one(xv.z, yv.z, g.z, mu.z, iv.z, gm.z, db.z, dg.z, fM, invM, relu);
      o.w = bn_bwd_one(xv.w, yv.w, g.w, mu.w, iv.w, gm.w, db.w, dg.w, fM, invM, relu);
      reinterpret_cast<float4*>(dx)[i] = o;
    }
    return;
  }
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    dx[i] = bn_bwd_one(__ldg(x + i), relu ? __ldg(y + i) : 1.f, __ldg(dy + i), __ldg(mean + c), __ldg(invstd + c),
                       gamma ? __ldg(gamma + c) : 1.f, __ldg(dbeta + c), __ldg(dgamma_full + c), fM, invM, relu);
  }
}

// ------------------------------------------------------------------------------------------
// Generic max-pool (NHWC). idx = kh*PW + kw of the first maximum.
// ------------------------------------------------------------------------------------------
__global__ void maxpool_fwd_kernel(const float* __restrict__ x, float* __restrict__ y,
                                   uint8_t* __restrict__ idx, int Bn, int H, int W, int C, int PH,
                                   int PW, int sh, int sw, int pt, int pl, int Ho, int Wo) {
  const long total = (long)Bn * Ho * Wo * C;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C); long t = i / C;
    const int ow = (int)(t % Wo); t /= Wo;
    const int oh = (int)(t % Ho); const long b = t / Ho;
    float best = -INFINITY; int bi = 0;
    for (int kh = 0; kh < PH; ++kh) {
      const int ih = oh * sh - pt + kh;
      if ((unsigned)ih >= (unsigned)H) continue;
      for (int kw = 0; kw < PW; ++kw) {
        const int iw = ow * sw - pl + kw;
        if ((unsigned)iw >= (unsigned)W) continue;
        const float v = __ldg(x + ((b * H + ih) * W + iw) * C + c);
        if (v > best) { best = v; bi = kh * PW + kw; }
      }
    }
    y[i] = best; idx[i] = (uint8_t)bi;
  }
}

// Gather formulation (no atomics, no zero-fill): each input element sums the windows that chose it.
__global__ void maxpool_bwd_kernel(const float* __restrict__ dy, const uint8_t* __restrict__ idx,
                                   float* __restrict__ dx, int Bn, int H, int W, int C, int PH,
                                   int PW, int sh, int sw, int pt, int pl, int Ho, int Wo) {
  const long total = (long)Bn * H * W * C;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C); long t = i / C;
    const int w = (int)(t % W); t /= W;
    const int h = (int)(t % H); const long b = t / H;
    float s = 0.f;
    const int hh = h + pt, ww = w + pl;
    int oh_lo = (hh - PH + sh) / sh; if (hh - PH + 1 <= 0) oh_lo = 0;
    int ow_lo = (ww - PW + sw) / sw; if (ww - PW + 1 <= 0) ow_lo = 0;
    const int oh_hi = min(hh / sh, Ho - 1), ow_hi = min(ww / sw, Wo - 1);
    for (int oh = oh_lo; oh <= oh_hi; ++oh) {
      const int kh = hh - oh * sh;
      if (kh < 0 || kh >= PH) continue;
      for (int ow = ow_lo; ow <= ow_hi; ++ow) {
        const int kw = ww - ow * sw;
        if (kw < 0 || kw >= PW) continue;
        const long o = ((b * Ho + oh) * Wo + ow) * C + c;
        if (idx[o] == (uint8_t)(kh * PW + kw)) s += __ldg(dy + o);
      }
    }
    dx[i] = s;
  }
}

// ------------------------------------------------------------------------------------------
// Elementwise
// ------------------------------------------------------------------------------------------
__global__ void relu_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, long n) {
  const long n4 = n >> 2;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long)gridDim.x * blockDim.x) {
    float4 v = __ldg(reinterpret_cast<const float4*>(x) + i);
    v.x = fmaxf(v.x, 0.f); v.y = fmaxf(v.y, 0.f); v.z = fmaxf(v.z, 0.f); v.w = fmaxf(v.w, 0.f);
    reinterpret_cast<float4*>(y)[i] = v;
  }
  for (long i = (n4 << 2) + (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x)
    y[i] = fmaxf(__ldg(x + i), 0.f);
}
__global__ void relu_bwd_kernel(const float* __restrict__ y, const float* __restrict__ dy,
                                float* __restrict__ dx, long n) {
  const long n4 = n >> 2;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long)gridDim.x * blockDim.x) {
    const float4 m = __ldg(reinterpret_cast<const float4*>(y) + i);
    float4 g = __ldg(reinterpret_cast<const float4*>(dy) + i);
    if (!(m.x > 0.f)) g.x = 0.f; if (!(m.y > 0.f)) g.y = 0.f; if (!(m.z > 0.f)) g.z = 0.f; if (!(m.w > 0.f)) g.w = 0.f;
    reinterpret_cast<float4*>(dx)[i] = g;
  }
  for (long i = (n4 << 2) + (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x)
    dx[i] = (__ldg(y + i) > 0.f) ? __ldg(dy + i) : 0.f;
}
__global__ void add_kernel(const float* __restrict__ a, const float* __restrict__ b,
                           float* __restrict__ y, long n, int relu) {
  const long n4 = n >> 2;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long)gridDim.x * blockDim.x) {
    const float4 u = __ldg(reinterpret_cast<const float4*>(a) + i), w = __ldg(reinterpret_cast<const float4*>(b) + i);
    float4 v = make_float4(u.x + w.x, u.y + w.y, u.z + w.z, u.w + w.w);
    if (relu) { v.x = fmaxf(v.x, 0.f); v.y = fmaxf(v.y, 0.f); v.z = fmaxf(v.z, 0.f); v.w = fmaxf(v.w, 0.f); }
    reinterpret_cast<float4*>(y)[i] = v;
  }
  for (long i = (n4 << 2) + (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
    float v = __ldg(a + i) + __ldg(b + i);
    y[i] = relu ? fmaxf(v, 0.f) : v;
  }
}
__global__ void u8_to_f32_kernel(const uint8_t* __restrict__ x, float* __restrict__ y, long n,
                                 float scale) {
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x)
    y[i] = (float)__ldg(x + i) * scale;
}

// Inverted dropout; the mask is regenerated from (seed, step, stream, index) in the backward pass, so
// nothing is stored. One Philox call covers 4 consecutive elements.
__global__ void dropout_kernel(const float* __restrict__ x, float* __restrict__ y, long n,
                               float rate, const TfdeHParams* __restrict__ hp, uint32_t stream_id,
                               int training) {
  const float scale = 1.f / (1.f - rate);
  const double thr_d = (double)rate * 4294967296.0;
  const uint32_t thr = thr_d >= 4294967295.0 ? 0xFFFFFFFFu : (uint32_t)thr_d;
  const uint32_t step = hp->step, k0 = hp->seed_lo, k1 = hp->seed_hi;
  const long n4 = (n + 3) / 4;
  for (long q = (long)blockIdx.x * blockDim.x + threadIdx.x; q < n4; q += (long)gridDim.x * blockDim.x) {
    if (!training) {
#pragma unroll
      for (int j = 0; j < 4; ++j) { const long i = q * 4 + j; if (i < n) y[i] = __ldg(x + i); }
      continue;
    }
    const Philox4 r = philox4x32_10((uint32_t)q, step, stream_id, 0u, k0, k1);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const long i = q * 4 + j;
      if (i < n) y[i] = (r.v[j] >= thr) ? __ldg(x + i) * scale : 0.f;
    }
  }
}

// Row softmax forward / backward: one warp per row, any C.
__global__ void softmax_fwd_kernel(const float* __restrict__ z, float* __restrict__ p, int rows, int C) {
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const float* zr = z + (long)row * C;
  float mx = -INFINITY;
  for (int c = lane; c < C; c += 32) mx = fmaxf(mx, __ldg(zr + c));
  mx = warp_max(mx);
  float s = 0.f;
  for (int c = lane; c < C; c += 32) s += __expf(__ldg(zr + c) - mx);
  s = warp_sum(s);
  const float inv = 1.f / s;
  for (int c = lane; c < C; c += 32) p[(long)row * C + c] = __expf(__ldg(zr + c) - mx) * inv;
}
__global__ void softmax_bwd_kernel(const float* __restrict__ p, const float* __restrict__ dp,
                                   float* __restrict__ dz, int rows, int C) {
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const long o = (long)row * C;
  float s = 0.f;
  for (int c = lane; c < C; c += 32) s = fmaf(__ldg(dp + o + c), __ldg(p + o + c), s);
  s = warp_sum(s);
  for (int c = lane; c < C; c += 32) dz[o + c] = __ldg(p + o + c) * (__ldg(dp + o + c) - s);
}

// Global average pool [B,H,W,C] -> [B,C] and its backward.
__global__ void gap_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, int Bn, int HW, int C) {
  const long total = (long)Bn * C;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C); const long b = i / C;
    float s = 0.f;
    for (int t = 0; t < HW; ++t) s += __ldg(x + (b * HW + t) * C + c);
    y[i] = s / (float)HW;
  }
}
__global__ void gap_bwd_kernel(const float* __restrict__ dy, float* __restrict__ dx, int Bn, int HW, int C) {
  const long total = (long)Bn * HW * C;
  const float inv = 1.f / (float)HW;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C); const long b = i / ((long)HW * C);
    dx[i] = __ldg(dy + b * C + c) * inv;
  }
}

inline int ew_grid(long n, int threads = 256) {
  long g = (n + threads - 1) / threads;
  const long cap = 148L * 8;
  return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}
// ---- small channel counts (C <= 32: Model B's 6/12/24-channel layers) -------------------------------------------
// With lane == channel only C of 32 lanes work (6 of 32 after the first convolution: 48 us for a 2.4 MB tensor).
// Here the tensor is read FLAT and fully coalesced: the launch has T = gridDim.x * 256 threads with T % C == 0, thread
// gt reads elements gt, gt + T, gt + 2T, ... which all belong to channel gt % C. Per block, C threads then add the
// per-thread sums of their channel in a fixed order; the last block finalises exactly like the wide kernels.
constexpr int BNS_THREADS = 256;

template <bool BWD>
__global__ void __launch_bounds__(BNS_THREADS)
bn_reduce_smallc_kernel(const float* __restrict__ x, const float* __restrict__ y, const float* __restrict__ dy, long M,
                        int C, const float* __restrict__ mean, const float* __restrict__ invstd, int relu,
                        float* __restrict__ partials, unsigned* __restrict__ counters,
                        // forward outputs
                        float* __restrict__ save_mean, float* __restrict__ save_invstd, float* __restrict__ moving_mean,
                        float* __restrict__ moving_var, float momentum, float eps, int unbiased_moving,
                        // backward outputs
                        float* __restrict__ dbeta, float* __restrict__ dgamma_full) {
  __shared__ float s_a[BNS_THREADS], s_b[BNS_THREADS];
  __shared__ int s_last;
  const int t = threadIdx.x;
  const long T = (long)gridDim.x * BNS_THREADS, gt = (long)blockIdx.x * BNS_THREADS + t, total = M * C;
  const int c = (int)(gt % C);
  float s = 0.f, q = 0.f;
  if (BWD) {
    const float mu = __ldg(mean + c), inv = __ldg(invstd + c);
    for (long i = gt; i < total; i += T) {
      float g = __ldg(dy + i);
      if (relu && !(__ldg(y + i) > 0.f)) g = 0.f;
      s += g; q = fmaf(g, (__ldg(x + i) - mu) * inv, q);
    }
  } else {
    for (long i = gt; i < total; i += T) {
      const float v = __ldg(x + i);
      s += v; q = fmaf(v, v, q);
    }
  }
  s_a[t] = s; s_b[t] = q;
  __syncthreads();
  float* part = partials + (size_t)blockIdx.x * 64;
  if (t < C) {
    // thread t sums the entries of channel (block_base + t) % C: t, t + C, t + 2C, ...
    float a = 0.f, b = 0.f;
    for (int j = t; j < BNS_THREADS; j += C) { a += s_a[j]; b += s_b[j]; }
    const int ch = (int)(((long)blockIdx.x * BNS_THREADS + t) % C);
    __stcg(part + ch, a); __stcg(part + 32 + ch, b);
  }
  __threadfence();
  __syncthreads();
  if (t == 0) s_last = (atomicAdd(counters, 1u) == gridDim.x - 1);
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  // finalise: 8 groups of 32 lanes (lane = channel) sum interleaved blocks, then the groups are combined in a fixed
  // order -- gridDim.x / 8 dependent L2 loads per thread instead of gridDim.x
  __shared__ double s_ds[8][32], s_dq[8][32];
  {
    const int lane = t & 31, grp = t >> 5;
    double ds = 0.0, dq = 0.0;
    if (lane < C)
      for (unsigned rb = grp; rb < gridDim.x; rb += 8) {
        ds += (double)__ldcg(partials + (size_t)rb * 64 + lane);
        dq += (double)__ldcg(partials + (size_t)rb * 64 + 32 + lane);
      }
    s_ds[grp][lane] = ds; s_dq[grp][lane] = dq;
  }
  __syncthreads();
  if (t < C) {
    double ds = 0.0, dq = 0.0;
#pragma unroll
    for (int g8 = 0; g8 < 8; ++g8) { ds += s_ds[g8][t]; dq += s_dq[g8][t]; }
    if (BWD) {
      dbeta[t] = (float)ds;
      dgamma_full[t] = (float)dq;
    } else {
      const double m = ds / (double)M;
      double var = dq / (double)M - m * m;
      if (var < 0.0) var = 0.0;
      save_mean[t] = (float)m;
      save_invstd[t] = rsqrtf((float)var + eps);
      const double vm = (unbiased_moving && M > 1) ? var * ((double)M / (double)(M - 1)) : var;
      moving_mean[t] = moving_mean[t] * momentum + (float)m * (1.f - momentum);
      moving_var[t] = moving_var[t] * momentum + (float)vm * (1.f - momentum);
    }
  }
  if (t == 0) *counters = 0u;
}

inline int bn_gcd(int a, int b) { while (b) { const int r = a % b; a = b; b = r; } return a; }
// blocks of the small-C reduction: ~16 elements per thread, at most 4 per SM, a multiple of C / gcd(C, 256) so that
// (blocks * 256) % C == 0
inline int bn_smallc_blocks(long M, int C) {
  const int unit = C / bn_gcd(C, BNS_THREADS);
  long want = (M * C + (long)BNS_THREADS * 16 - 1) / ((long)BNS_THREADS * 16);
  if (want > 148L * 2) want = 148L * 2;
  want = (want + unit - 1) / unit * unit;
  return (int)(want < unit ? unit : want);
}
inline bool bn_use_smallc(long M, int C) { return C <= 32 && M * C >= 8192; }

inline int bn_row_blocks(long M, int ctiles) {
  long want = (M + BN_ROWS * 8 - 1) / (BN_ROWS * 8);  // >= 8 rows per thread
  long cap = (148L * 4 + ctiles - 1) / ctiles;
  if (want > cap) want = cap;
  if (want < 1) want = 1;
  return (int)want;
}

}  // namespace

TFDE_API size_t tfde_bn_ws_floats(long M, int C) {
  const int ct = (C + 31) / 32;
  const size_t wide = (size_t)ct * bn_row_blocks(M, ct) * 64;
  const size_t small_c = C <= 32 ? (size_t)bn_smallc_blocks(M, C) * 64 : 0;
  return wide > small_c ? wide : small_c;
}

TFDE_API int tfde_bn_fwd_train(const float* x, float* y, long M, int C, const float* gamma,
                               const float* beta, float* moving_mean, float* moving_var,
                               float* save_mean, float* save_invstd, float momentum, float eps,
                               int relu, int unbiased_moving, float* partials, unsigned* counters,
                               cudaStream_t stream) {
  const int ct = (C + 31) / 32;
  if (ct > TFDE_MAX_COUNTERS) return TFDE_ERR_INVALID;
  if (bn_use_smallc(M, C)) {
    bn_reduce_smallc_kernel<false><<<bn_smallc_blocks(M, C), BNS_THREADS, 0, stream>>>(
        x, nullptr, nullptr, M, C, nullptr, nullptr, 0, partials, counters, save_mean, save_invstd, moving_mean,
        moving_var, momentum, eps, unbiased_moving, nullptr, nullptr);
  } else {
    dim3 grid(ct, bn_row_blocks(M, ct)), block(32, BN_ROWS);
    bn_stats_kernel<<<grid, block, 0, stream>>>(x, M, C, partials, counters, save_mean, save_invstd,
                                                moving_mean, moving_var, momentum, eps, unbiased_moving);
  }
  const long total = M * C;
  bn_apply_kernel<<<ew_grid(total), 256, 0, stream>>>(x, y, total, C, save_mean, save_invstd, 0, eps,
                                                      gamma, beta, relu);
  TFDE_RETURN_LAUNCH_STATUS();
}

TFDE_API int tfde_bn_fwd_infer(const float* x, float* y, long M, int C, const float* gamma,
                               const float* beta, const float* moving_mean, const float* moving_var,
                               float eps, int relu, cudaStream_t stream) {
  const long total = M * C;
  bn_apply_kernel<<<ew_grid(total), 256, 0, stream>>>(x, y, total, C, moving_mean, moving_var, 1, eps,
                                                      gamma, beta, relu);
  TFDE_RETURN_LAUNCH_STATUS();
}

TFDE_API int tfde_bn_bwd(const float* x, const float* y, const float* dy, float* dx, long M, int C,
                         const float* gamma, const float* save_mean, const float* save_invstd,
                         int relu, float* dbeta, float* dgamma_full, float* partials,
                         unsigned* counters, cudaStream_t stream) {
  const int ct = (C + 31) / 32;
  if (ct > TFDE_MAX_COUNTERS) return TFDE_ERR_INVALID;
  if (bn_use_smallc(M, C)) {
    bn_reduce_smallc_kernel<true><<<bn_smallc_blocks(M, C), BNS_THREADS, 0, stream>>>(
        x, y, dy, M, C, save_mean, save_invstd, relu, partials, counters, nullptr, nullptr, nullptr, nullptr, 0.f, 0.f,
        0, dbeta, dgamma_full);
  } else {
    dim3 grid(ct, bn_row_blocks(M, ct)), block(32, BN_ROWS);
    bn_bwd_reduce_kernel<<<grid, block, 0, stream>>>(x, y, dy, M, C, save_mean, save_invstd, relu,
                                                     partials, counters, dbeta, dgamma_full);
  }
  const long total = M * C;
  bn_bwd_apply_kernel<<<ew_grid(total), 256, 0, stream>>>(x, y, dy, dx, total, M, C, save_mean,
                                                          save_invstd, gamma, dbeta, dgamma_full, relu);
  TFDE_RETURN_LAUNCH_STATUS();
}

TFDE_API int tfde_maxpool_fwd(const float* x, float* y, uint8_t* idx, int Bn, int H, int W, int C,
                              int PH, int PW, int sh, int sw, int pt, int pl, int Ho, int Wo,
                              cudaStream_t stream) {
  if (PH * PW > 255) return TFDE_ERR_INVALID;
  maxpool_fwd_kernel<<<ew_grid((long)Bn * Ho * Wo * C), 256, 0, stream>>>(x, y, idx, Bn, H, W, C, PH, PW, sh, sw, pt, pl, Ho, Wo);
  TFDE_RETURN_LAUNCH_STATUS();
}
TFDE_API int tfde_maxpool_bwd(const float* dy, const uint8_t* idx, float* dx, int Bn, int H, int W,
                              int C, int PH, int PW, int sh, int sw, int pt, int pl, int Ho, int Wo,
                              cudaStream_t stream) {
  maxpool_bwd_kernel<<<ew_grid((long)Bn * H * W * C), 256, 0, stream>>>(dy, idx, dx, Bn, H, W, C, PH, PW, sh, sw, pt, pl, Ho, Wo);
  TFDE_RETURN_LAUNCH_STATUS();
}
TFDE_API int tfde_relu_fwd(const float* x, float* y, long n, cudaStream_t stream) {
  relu_fwd_kernel<<<ew_grid(n), 256, 0, stream>>>(x, y, n);
  TFDE_RETURN_LAUNCH_STATUS();
}
TFDE_API int tfde_relu_bwd(const float* y, const float* dy, float* dx, long n, cudaStream_t stream) {
  relu_bwd_kernel<<<ew_grid(n), 256, 0, stream>>>(y, dy, dx, n);
  TFDE_RETURN_LAUNCH_STATUS();
}
TFDE_API int tfde_add(const float* a, const float* b, float* y, long n, int relu, cudaStream_t stream) {
  add_kernel<<<ew_grid(n), 256, 0, stream>>>(a, b, y, n, relu);
  TFDE_RETURN_LAUNCH_STATUS();
}
TFDE_API int tfde_u8_to_f32(const uint8_t* x, float* y, long n, float scale, cudaStream_t stream) {
  u8_to_f32_kernel<<<ew_grid(n), 256, 0, stream>>>(x, y, n, scale);
  TFDE_RETURN_LAUNCH_STATUS();
}
TFDE_API int tfde_dropout(const float* x, float* y, long n, float rate, const void* hparams,
                          unsigned stream_id, int training, cudaStream_t stream) {
  if (!(rate >= 0.f && rate < 1.f)) return TFDE_ERR_INVALID;
  dropout_kernel<<<ew_grid((n + 3) / 4), 256, 0, stream>>>(x, y, n, rate, (const TfdeHParams*)hparams, stream_id, training);
  TFDE_RETURN_LAUNCH_STATUS();
}
TFDE_API int tfde_softmax_fwd(const float* z, float* p, int rows, int C, cudaStream_t stream) {
  softmax_fwd_kernel<<<(rows + 7) / 8, 256, 0, stream>>>(z, p, rows, C);
  TFDE_RETURN_LAUNCH_STATUS();
}
TFDE_API int tfde_softmax_bwd(const float* p, const float* dp, float* dz, int rows, int C, cudaStream_t stream) {
  softmax_bwd_kernel<<<(rows + 7) / 8, 256, 0, stream>>>(p, dp, dz, rows, C);
  TFDE_RETURN_LAUNCH_STATUS();
}
TFDE_API int tfde_gap_fwd(const float* x, float* y, int Bn, int HW, int C, cudaStream_t stream) {
  gap_fwd_kernel<<<ew_grid((long)Bn * C), 256, 0, stream>>>(x, y, Bn, HW, C);
  TFDE_RETURN_LAUNCH_STATUS();
}
TFDE_API int tfde_gap_bwd(const float* dy, float* dx, int Bn, int HW, int C, cudaStream_t stream) {
  gap_bwd_kernel<<<ew_grid((long)Bn * HW * C), 256, 0, stream>>>(dy, dx, Bn, HW, C);
  TFDE_RETURN_LAUNCH_STATUS();
}
