// Direct convolution for SMALL channel counts: the three Conv2D layers of "Model B"
// (mnist_keras_distributed.py:83-98 / tf2_mnist_distributed.py:109-124): 3x3x1->6 'same', 6x6x6->12 'same' stride 2,
// 6x6x12->24 'same' stride 2, all without bias. With Cout = 6/12/24 and Cin = 1/6/12 an implicit GEMM wastes its
// tiles on padding (the generic fp32 GEMM path measured 29/57/28 us forward, 55/135/254 us filter gradient and
// 253/563 us data gradient per layer at batch 128 -- 1.3 of the 1.6 ms step) and a stride-2 data gradient that
// walks all 36 taps multiplies three zeros out of four. Here:
//   * one CTA per image: the zero-padded input image (channel-planar) and the whole filter live in shared memory;
//   * forward: thread = (output pixel, group of CG output channels), accumulators in registers, weights are warp-
//     uniform broadcast loads;
//   * data gradient: thread = (input pixel, group of CG input channels); only the taps whose stride phase matches the
//     pixel are visited (9 of 36 at stride 2), pixels are enumerated by stride phase so a warp shares its taps;
//   * filter gradient: thread = (tap, input channel, group of output channels), reduction over the image's output
//     pixels in registers, one partial per image, then a fixed-order reduction over the batch (deterministic).
// fp32 FFMA, exact like the reference. NHWC activations, HWIO filters (the Keras layouts).
#include "common.cuh"

namespace {

struct SCParams {
  const float* x; const float* w; const float* bias; float* y;      // forward
  const float* dy; float* dx; float* part;                          // backward (part: [B][KH*KW*Cin*Cout])
  int B, H, W, Cin, KH, KW, Ho, Wo, sh, sw, pt, pl, Cout, relu;
  int Hp, Wp, WpS;       // padded input extent covered by the output grid, odd row stride in shared memory
  int plane;             // Hp * WpS (+1 when even): floats per input-channel plane
};

constexpr int SC_THREADS = 256;

__device__ __forceinline__ void sc_load_filter(float* ws, const float* __restrict__ w, int n) {
  for (int i = threadIdx.x; i < n; i += SC_THREADS) ws[i] = __ldg(w + i);
}

// Zero-padded channel-planar copy of image b: xs[ci][iy][ix], iy/ix in padded coordinates.
__device__ __forceinline__ void sc_load_image(float* xs, const SCParams& p, int b) {
  const int per = p.Hp * p.Wp * p.Cin;
  const float* xb = p.x + (size_t)b * p.H * p.W * p.Cin;
  for (int i = threadIdx.x; i < per; i += SC_THREADS) {
    const int ci = i % p.Cin, r = i / p.Cin, ix = r % p.Wp, iy = r / p.Wp;
    const int gy = iy - p.pt, gx = ix - p.pl;
    float v = 0.f;
    if ((unsigned)gy < (unsigned)p.H && (unsigned)gx < (unsigned)p.W) v = __ldg(xb + ((size_t)gy * p.W + gx) * p.Cin + ci);
    xs[ci * p.plane + iy * p.WpS + ix] = v;
  }
}

template <int CG>
__global__ void __launch_bounds__(SC_THREADS)
smallconv_fwd_kernel(const SCParams p) {
  TFDE_PDL_PROLOGUE();
  extern __shared__ __align__(16) float sm[];
  float* ws = sm;                                        // [KH*KW*Cin][Cout]
  float* xs = sm + ((p.KH * p.KW * p.Cin * p.Cout + 3) & ~3);
  const int b = blockIdx.x;
  sc_load_filter(ws, p.w, p.KH * p.KW * p.Cin * p.Cout);
  sc_load_image(xs, p, b);
  __syncthreads();
  const int npix = p.Ho * p.Wo, groups = p.Cout / CG;
  float* yb = p.y + (size_t)b * npix * p.Cout;
  for (int item = threadIdx.x; item < npix * groups; item += SC_THREADS) {
    const int g = item / npix, pix = item - g * npix;     // a warp shares g (npix >= 32 in practice): broadcast weights
    const int oy = pix / p.Wo, ox = pix - oy * p.Wo;
    float acc[CG];
#pragma unroll
    for (int j = 0; j < CG; ++j) acc[j] = p.bias ? __ldg(p.bias + g * CG + j) : 0.f;
    const float* wp = ws + g * CG;
    for (int kh = 0; kh < p.KH; ++kh) {
      for (int kw = 0; kw < p.KW; ++kw) {
        const float* xp = xs + (oy * p.sh + kh) * p.WpS + ox * p.sw + kw;
        for (int ci = 0; ci < p.Cin; ++ci) {
          const float xv = xp[ci * p.plane];
#pragma unroll
          for (int j = 0; j < CG; j += 2) {
            const float2 w2 = *reinterpret_cast<const float2*>(wp + j);
            acc[j] = fmaf(xv, w2.x, acc[j]);
            acc[j + 1] = fmaf(xv, w2.y, acc[j + 1]);
          }
          wp += p.Cout;
        }
      }
    }
    float* yo = yb + (size_t)pix * p.Cout + g * CG;
#pragma unroll
    for (int j = 0; j < CG; ++j) yo[j] = p.relu ? fmaxf(acc[j], 0.f) : acc[j];
  }
}

// ---------------------------------------------------------------------------------------------------------------
// data gradient: dx[b, iy, ix, ci] = sum_{kh, kw, co} dy[b, oy, ox, co] * w[kh, kw, ci, co],  oy*sh + kh - pt == iy
// ---------------------------------------------------------------------------------------------------------------
template <int CG>
__global__ void __launch_bounds__(SC_THREADS)
smallconv_dgrad_kernel(const SCParams p) {
  TFDE_PDL_PROLOGUE();
  extern __shared__ __align__(16) float sm[];
  const int taps = p.KH * p.KW;
  float* wt = sm;                                        // [taps][Cout][Cin]  (transposed: a thread's CG input channels are contiguous)
  float* dys = sm + ((taps * p.Cin * p.Cout + 3) & ~3);  // [Cout][Ho][WoS] channel-planar
  const int b = blockIdx.x;
  const int WoS = p.Wo | 1, dplane = p.Ho * WoS + ((p.Ho * WoS) & 1 ? 0 : 1);
  for (int i = threadIdx.x; i < taps * p.Cin * p.Cout; i += SC_THREADS) {
    const int co = i % p.Cout, r = i / p.Cout, ci = r % p.Cin, tap = r / p.Cin;
    wt[(tap * p.Cout + co) * p.Cin + ci] = __ldg(p.w + i);
  }
  const int npo = p.Ho * p.Wo;
  const float* dyb = p.dy + (size_t)b * npo * p.Cout;
  for (int i = threadIdx.x; i < npo * p.Cout; i += SC_THREADS) {
    const int co = i % p.Cout, pix = i / p.Cout, oy = pix / p.Wo, ox = pix - oy * p.Wo;
    dys[co * dplane + oy * WoS + ox] = __ldg(dyb + i);
  }
  __syncthreads();
  const int npix = p.H * p.W, groups = p.Cin / CG;
  // enumerate pixels by stride phase so that the threads of a warp visit the same taps
  const bool phased = (p.H % p.sh == 0) && (p.W % p.sw == 0);
  const int hq = p.H / p.sh, wq = p.W / p.sw, per_phase = hq * wq;
  float* dxb = p.dx + (size_t)b * npix * p.Cin;
  for (int item = threadIdx.x; item < npix * groups; item += SC_THREADS) {
    const int g = item / npix, q = item - g * npix;
    int iy, ix;
    if (phased) {
      const int ph = q / per_phase, r = q - ph * per_phase;
      iy = (r / wq) * p.sh + ph / p.sw; ix = (r % wq) * p.sw + ph % p.sw;
    } else { iy = q / p.W; ix = q - iy * p.W; }
    float acc[CG];
#pragma unroll
    for (int j = 0; j < CG; ++j) acc[j] = 0.f;
    for (int kh = 0; kh < p.KH; ++kh) {
      const int ty = iy + p.pt - kh;
      if (ty < 0 || ty % p.sh) continue;
      const int oy = ty / p.sh;
      if (oy >= p.Ho) continue;
      for (int kw = 0; kw < p.KW; ++kw) {
        const int tx = ix + p.pl - kw;
        if (tx < 0 || tx % p.sw) continue;
        const int ox = tx / p.sw;
        if (ox >= p.Wo) continue;
        const float* dp = dys + oy * WoS + ox;
        const float* wp = wt + (size_t)(kh * p.KW + kw) * p.Cout * p.Cin + g * CG;
        for (int co = 0; co < p.Cout; ++co) {
          const float dv = dp[co * dplane];
#pragma unroll
          for (int j = 0; j < CG; j += 2) {
            const float2 w2 = *reinterpret_cast<const float2*>(wp + j);
            acc[j] = fmaf(dv, w2.x, acc[j]);
            acc[j + 1] = fmaf(dv, w2.y, acc[j + 1]);
          }
          wp += p.Cin;
        }
      }
    }
    float* o = dxb + ((size_t)iy * p.W + ix) * p.Cin + g * CG;
#pragma unroll
    for (int j = 0; j < CG; ++j) o[j] = acc[j];
  }
}

// ---------------------------------------------------------------------------------------------------------------
// filter gradient: dW[kh, kw, ci, co] = sum_{b, oy, ox} x[b, oy*sh+kh-pt, ox*sw+kw-pl, ci] * dy[b, oy, ox, co]
// one partial per image (PCH pixel chunks per image reduced through shared memory), then a fixed-order sum over b
// ---------------------------------------------------------------------------------------------------------------
template <int CG>
__global__ void __launch_bounds__(SC_THREADS)
smallconv_wgrad_kernel(const SCParams p, int pch) {
  TFDE_PDL_PROLOGUE();
  extern __shared__ __align__(16) float sm[];
  const int npo = p.Ho * p.Wo, nW = p.KH * p.KW * p.Cin * p.Cout;
  float* dys = sm;                                        // [pix][Cout]
  float* xs = dys + ((npo * p.Cout + 3) & ~3);            // [Cin][Hp][WpS]
  float* red = xs + ((p.Cin * p.plane + 3) & ~3);         // [pch][nW] when pch > 1
  const int b = blockIdx.x;
  const float* dyb = p.dy + (size_t)b * npo * p.Cout;
  for (int i = threadIdx.x; i < npo * p.Cout; i += SC_THREADS) dys[i] = __ldg(dyb + i);
  sc_load_image(xs, p, b);
  __syncthreads();
  const int groups = p.Cout / CG, base = p.KH * p.KW * p.Cin * groups;
  float* out = p.part + (size_t)b * nW;
  for (int item = threadIdx.x; item < base * pch; item += SC_THREADS) {
    const int chunk = item / base, e = item - chunk * base;
    const int g = e % groups, tc = e / groups, ci = tc % p.Cin, tap = tc / p.Cin, kh = tap / p.KW, kw = tap - kh * p.KW;
    const int p0 = (int)((long)npo * chunk / pch), p1 = (int)((long)npo * (chunk + 1) / pch);
    float acc[CG];
#pragma unroll
    for (int j = 0; j < CG; ++j) acc[j] = 0.f;
    const float* xc = xs + ci * p.plane + kh * p.WpS + kw;
    int oy = p0 / p.Wo, ox = p0 - oy * p.Wo;
    for (int pix = p0; pix < p1; ++pix) {
      const float xv = xc[oy * p.sh * p.WpS + ox * p.sw];
      const float* dp = dys + pix * p.Cout + g * CG;
#pragma unroll
      for (int j = 0; j < CG; j += 2) {
        const float2 d2 = *reinterpret_cast<const float2*>(dp + j);
        acc[j] = fmaf(xv, d2.x, acc[j]);
        acc[j + 1] = fmaf(xv, d2.y, acc[j + 1]);
      }
      if (++ox == p.Wo) { ox = 0; ++oy; }
    }
    float* dst = (pch > 1 ? red + (size_t)chunk * nW : out) + tc * p.Cout + g * CG;
#pragma unroll
    for (int j = 0; j < CG; ++j) dst[j] = acc[j];
  }
  if (pch > 1) {
    __syncthreads();
    for (int i = threadIdx.x; i < nW; i += SC_THREADS) {
      float a = 0.f;
      for (int c = 0; c < pch; ++c) a += red[(size_t)c * nW + i];
      out[i] = a;
    }
  }
}

// dw[i] = sum_b part[b][i] in a fixed order: block = 32 elements x 8 batch groups (each thread adds B/8 partials with
// 8 loads in flight), the groups are then combined in order through shared memory.
__global__ void __launch_bounds__(256)
smallconv_wgrad_reduce_kernel(const float* __restrict__ part, float* __restrict__ dw, int nW, int B) {
  TFDE_PDL_PROLOGUE();
  __shared__ float s_p[8][32];
  const int lane = threadIdx.x & 31, grp = threadIdx.x >> 5;
  const int i = blockIdx.x * 32 + lane;
  float a = 0.f;
  if (i < nW) {
    const int per = (B + 7) / 8, b0 = grp * per, b1 = min(B, b0 + per);
    int b = b0;
    for (; b + 8 <= b1; b += 8) {
      float v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) v[u] = __ldcg(part + (size_t)(b + u) * nW + i);
#pragma unroll
      for (int u = 0; u < 8; ++u) a += v[u];
    }
    for (; b < b1; ++b) a += __ldcg(part + (size_t)b * nW + i);
  }
  s_p[grp][lane] = a;
  __syncthreads();
  if (grp == 0 && i < nW) {
    float t = 0.f;
#pragma unroll
    for (int g8 = 0; g8 < 8; ++g8) t += s_p[g8][lane];
    dw[i] = t;
  }
}

}  // namespace

// ---------------------------------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------------------------------
namespace {
constexpr size_t SC_SMEM_MAX = 200 * 1024;

bool sc_fill(SCParams& p, int B, int H, int W, int Cin, int KH, int KW, int Ho, int Wo, int sh, int sw, int pt, int pl,
             int Cout) {
  p.B = B; p.H = H; p.W = W; p.Cin = Cin; p.KH = KH; p.KW = KW; p.Ho = Ho; p.Wo = Wo; p.sh = sh; p.sw = sw;
  p.pt = pt; p.pl = pl; p.Cout = Cout;
  p.Hp = (Ho - 1) * sh + KH; p.Wp = (Wo - 1) * sw + KW;
  p.WpS = p.Wp | 1;
  p.plane = p.Hp * p.WpS; if (!(p.plane & 1)) p.plane += 1;
  return B >= 1 && H >= 1 && W >= 1 && Cin >= 1 && Cout >= 2 && KH >= 1 && KW >= 1 && Ho >= 1 && Wo >= 1 && sh >= 1 &&
         sw >= 1 && pt >= 0 && pl >= 0;
}
int sc_group(int c) { return c % 12 == 0 ? 12 : (c % 6 == 0 ? 6 : (c % 8 == 0 ? 8 : (c % 4 == 0 ? 4 : (c % 2 == 0 ? 2 : 0)))); }
int sc_group_small(int c) { return c % 6 == 0 ? 6 : (c % 8 == 0 ? 8 : (c % 4 == 0 ? 4 : (c % 2 == 0 ? 2 : 0))); }
size_t sc_smem_fwd(const SCParams& p) {
  return (size_t)(((p.KH * p.KW * p.Cin * p.Cout + 3) & ~3) + p.Cin * p.plane) * 4;
}
size_t sc_smem_dgrad(const SCParams& p) {
  const int WoS = p.Wo | 1; int dplane = p.Ho * WoS; if (!(dplane & 1)) dplane += 1;
  return (size_t)(((p.KH * p.KW * p.Cin * p.Cout + 3) & ~3) + p.Cout * dplane) * 4;
}
int sc_pch(const SCParams& p, int cg) {
  const int base = p.KH * p.KW * p.Cin * (p.Cout / cg), nW = p.KH * p.KW * p.Cin * p.Cout, npo = p.Ho * p.Wo;
  int pch = (2 * SC_THREADS + base - 1) / base;            // aim at >= 2 items per thread slot
  if (pch > npo / 8) pch = npo / 8;
  while (pch > 1 && (size_t)pch * nW * 4 > 48 * 1024) --pch;
  return pch < 1 ? 1 : pch;
}
size_t sc_smem_wgrad(const SCParams& p, int pch) {
  const int nW = p.KH * p.KW * p.Cin * p.Cout;
  return (size_t)(((p.Ho * p.Wo * p.Cout + 3) & ~3) + ((p.Cin * p.plane + 3) & ~3) + (pch > 1 ? pch * nW : 0)) * 4;
}

template <typename K>
int sc_launch(K kernel, size_t smem, size_t (&done)[TFDE_MAX_DEVICES], int grid, cudaStream_t stream, const SCParams& p) {
  const cudaError_t e = tfde_ensure_dyn_smem(kernel, smem, done);
  if (e != cudaSuccess) return (int)e;
  tfde_launch(kernel, dim3(grid), dim3(SC_THREADS), smem, stream, p);
  return 0;
}
}  // namespace

// 1 when the direct small-channel kernels cover this convolution (forward; `bwd` != 0: backward too, `need_dx` says
// whether a data gradient is wanted). Otherwise the caller uses the implicit-GEMM path (tfde_conv2d_*).
TFDE_API int tfde_smallconv_supported(int B, int H, int W, int Cin, int KH, int KW, int Ho, int Wo, int sh, int sw,
                                      int pt, int pl, int Cout, int bwd, int need_dx) {
  SCParams p{};
  if (!sc_fill(p, B, H, W, Cin, KH, KW, Ho, Wo, sh, sw, pt, pl, Cout)) return 0;
  if (Cout > 32 || Cin > 32 || (Cout & 1) || sc_group_small(Cout) == 0) return 0;
  if (sc_smem_fwd(p) > SC_SMEM_MAX) return 0;
  if (bwd) {
    const int cg = sc_group(Cout);
    if (sc_smem_wgrad(p, sc_pch(p, cg)) > SC_SMEM_MAX) return 0;
    if (need_dx && ((Cin & 1) || sc_group_small(Cin) == 0 || sc_smem_dgrad(p) > SC_SMEM_MAX)) return 0;
  }
  return 1;
}
TFDE_API size_t tfde_smallconv_bwd_ws_floats(int B, int Cin, int KH, int KW, int Cout) {
  return (size_t)B * KH * KW * Cin * Cout;
}

// y = conv(x, w) (+ bias) (+ ReLU). ref: Conv2D layers of mnist_keras_distributed.py:83-98 (cuDNN in TF).
TFDE_API int tfde_smallconv_fwd(const float* x, const float* w, const float* bias, float* y, int B, int H, int W, int Cin,
                                int KH, int KW, int Ho, int Wo, int sh, int sw, int pt, int pl, int Cout, int relu,
                                cudaStream_t stream) {
  if (!tfde_smallconv_supported(B, H, W, Cin, KH, KW, Ho, Wo, sh, sw, pt, pl, Cout, 0, 0)) return TFDE_ERR_INVALID;
  SCParams p{};
  sc_fill(p, B, H, W, Cin, KH, KW, Ho, Wo, sh, sw, pt, pl, Cout);
  p.x = x; p.w = w; p.bias = bias; p.y = y; p.relu = relu;
  const size_t smem = sc_smem_fwd(p);
  static size_t d2[TFDE_MAX_DEVICES] = {}, d4[TFDE_MAX_DEVICES] = {}, d6[TFDE_MAX_DEVICES] = {}, d8[TFDE_MAX_DEVICES] = {};
  int rc;
  switch (sc_group_small(Cout)) {
    case 6: rc = sc_launch(smallconv_fwd_kernel<6>, smem, d6, B, stream, p); break;
    case 8: rc = sc_launch(smallconv_fwd_kernel<8>, smem, d8, B, stream, p); break;
    case 4: rc = sc_launch(smallconv_fwd_kernel<4>, smem, d4, B, stream, p); break;
    default: rc = sc_launch(smallconv_fwd_kernel<2>, smem, d2, B, stream, p); break;
  }
  if (rc) return rc;
  TFDE_RETURN_LAUNCH_STATUS();
}

// dw = filter gradient (no bias, no fused activation), dx = data gradient (NULL: not wanted, e.g. first layer).
// ws: tfde_smallconv_bwd_ws_floats() floats of scratch (per-image partials).
TFDE_API int tfde_smallconv_bwd(const float* x, const float* dy, const float* w, float* dx, float* dw, float* ws, int B,
                                int H, int W, int Cin, int KH, int KW, int Ho, int Wo, int sh, int sw, int pt, int pl,
                                int Cout, cudaStream_t stream) {
  if (!tfde_smallconv_supported(B, H, W, Cin, KH, KW, Ho, Wo, sh, sw, pt, pl, Cout, 1, dx != nullptr)) return TFDE_ERR_INVALID;
  SCParams p{};
  sc_fill(p, B, H, W, Cin, KH, KW, Ho, Wo, sh, sw, pt, pl, Cout);
  p.x = x; p.w = w; p.dy = dy; p.dx = dx; p.part = ws;
  const int nW = KH * KW * Cin * Cout;
  {
    const int cg = sc_group(Cout), pch = sc_pch(p, cg);
    const size_t smem = sc_smem_wgrad(p, pch);
    static size_t d2[TFDE_MAX_DEVICES] = {}, d4[TFDE_MAX_DEVICES] = {}, d6[TFDE_MAX_DEVICES] = {}, d8[TFDE_MAX_DEVICES] = {},
                  d12[TFDE_MAX_DEVICES] = {};
    cudaError_t e = cudaSuccess;
    switch (cg) {
      case 12: e = tfde_ensure_dyn_smem(smallconv_wgrad_kernel<12>, smem, d12); if (!e) tfde_launch(smallconv_wgrad_kernel<12>, dim3(B), dim3(SC_THREADS), smem, stream, p, pch); break;
      case 6: e = tfde_ensure_dyn_smem(smallconv_wgrad_kernel<6>, smem, d6); if (!e) tfde_launch(smallconv_wgrad_kernel<6>, dim3(B), dim3(SC_THREADS), smem, stream, p, pch); break;
      case 8: e = tfde_ensure_dyn_smem(smallconv_wgrad_kernel<8>, smem, d8); if (!e) tfde_launch(smallconv_wgrad_kernel<8>, dim3(B), dim3(SC_THREADS), smem, stream, p, pch); break;
      case 4: e = tfde_ensure_dyn_smem(smallconv_wgrad_kernel<4>, smem, d4); if (!e) tfde_launch(smallconv_wgrad_kernel<4>, dim3(B), dim3(SC_THREADS), smem, stream, p, pch); break;
      default: e = tfde_ensure_dyn_smem(smallconv_wgrad_kernel<2>, smem, d2); if (!e) tfde_launch(smallconv_wgrad_kernel<2>, dim3(B), dim3(SC_THREADS), smem, stream, p, pch); break;
    }
    if (e != cudaSuccess) return (int)e;
    tfde_launch(smallconv_wgrad_reduce_kernel, dim3((nW + 31) / 32), dim3(256), 0, stream, (const float*)ws, dw, nW, B);
  }
  if (dx != nullptr) {
    const size_t smem = sc_smem_dgrad(p);
    static size_t d2[TFDE_MAX_DEVICES] = {}, d4[TFDE_MAX_DEVICES] = {}, d6[TFDE_MAX_DEVICES] = {}, d8[TFDE_MAX_DEVICES] = {};
    int rc;
    switch (sc_group_small(Cin)) {
      case 6: rc = sc_launch(smallconv_dgrad_kernel<6>, smem, d6, B, stream, p); break;
      case 8: rc = sc_launch(smallconv_dgrad_kernel<8>, smem, d8, B, stream, p); break;
      case 4: rc = sc_launch(smallconv_dgrad_kernel<4>, smem, d4, B, stream, p); break;
      default: rc = sc_launch(smallconv_dgrad_kernel<2>, smem, d2, B, stream, p); break;
    }
    if (rc) return rc;
  }
  TFDE_RETURN_LAUNCH_STATUS();
}
