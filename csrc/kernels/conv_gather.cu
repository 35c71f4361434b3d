// Operand builders for running NHWC x HWIO convolutions on the tcgen05 bf16 GEMMs (gemm_tcgen05*.cu).
// A convolution is three GEMMs; each needs its operands K-major in
// bf16, which these gather kernels produce straight from the fp32 activations / gradients / weights, with
// the fp32->bf16 cast and the ReLU-backward mask folded in:
//
//   forward  y[M, Cout]    = im2col(x)[M, Kc]        . W^T[Cout, Kc]^T          M = B*Ho*Wo, Kc = KH*KW*Cin
//   wgrad    dW[Kc, Cout]  = im2col(x)^T[Kc, M]      . dy^T[Cout, M]^T
//   dgrad    dx[Min, Cin]  = gather(dy)[Min, Kd]     . Wrot[Cin, Kd]^T          Min = B*H*W, Kd = KH*KW*Cout
//
// Row strides (`ld*`) are in elements and multiples of 8 (TMA needs 16-byte global strides); padding
// columns are written as zeros so they contribute nothing to the reduction.
// Reference ops replaced: Conv2D forward/backward (cuDNN in TF): distributed_with_keras.py:34,
// mnist_keras_distributed.py:83-98; used by the ResNet-50 config of BASELINE.json.
#include "common.cuh"
#include <cuda_bf16.h>

namespace {

struct CG {
  int B, H, W, C, KH, KW, Ho, Wo, sh, sw, pt, pl, Cout;
};

__device__ __forceinline__ uint32_t pack2(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}

// out[m, k] (ld = ldo) ; one thread = 8 consecutive k of one row m
__global__ void im2col_bf16_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ out, const CG g,
                                   long M, int Kc, int ldo) {
  TFDE_PDL_PROLOGUE();
  const int k8n = ldo >> 3;
  const long total = M * k8n;
  if (total < (1L << 31) && (g.C & 7) != 0) {
    // Few-channel input (the 3-channel network stem): 32-bit index arithmetic and an incrementally advanced
    // (kh, kw, ci) tap instead of four divisions per element.
    const unsigned k8 = (unsigned)k8n, Wo = (unsigned)g.Wo, Ho = (unsigned)g.Ho, Cu = (unsigned)g.C, KW = (unsigned)g.KW;
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < (unsigned)total; i += gridDim.x * blockDim.x) {
      const unsigned m = i / k8, k0 = (i - m * k8) << 3;
      const unsigned t = m / Wo, ow = m - t * Wo;
      const unsigned b = t / Ho, oh = t - b * Ho;
      const int ih0 = (int)(oh * g.sh) - g.pt, iw0 = (int)(ow * g.sw) - g.pl;
      const float* xb = x + (size_t)b * g.H * g.W * g.C;
      unsigned tap = k0 / Cu, ci = k0 - tap * Cu;
      unsigned kh = tap / KW, kw = tap - kh * KW;
      float v[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float e = 0.f;
        const int ih = ih0 + (int)kh, iw = iw0 + (int)kw;
        if (k0 + j < (unsigned)Kc && (unsigned)ih < (unsigned)g.H && (unsigned)iw < (unsigned)g.W)
          e = __ldg(xb + ((size_t)ih * g.W + iw) * g.C + ci);
        v[j] = e;
        if (++ci == Cu) { ci = 0; if (++kw == KW) { kw = 0; ++kh; } }
      }
      uint4 o; o.x = pack2(v[0], v[1]); o.y = pack2(v[2], v[3]); o.z = pack2(v[4], v[5]); o.w = pack2(v[6], v[7]);
      *reinterpret_cast<uint4*>(out + (size_t)m * ldo + k0) = o;
    }
    return;
  }
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const long m = i / k8n;
    const int k0 = (int)(i % k8n) * 8;
    const int ow = (int)(m % g.Wo); const long t = m / g.Wo;
    const int oh = (int)(t % g.Ho); const long b = t / g.Ho;
    const int ih0 = oh * g.sh - g.pt, iw0 = ow * g.sw - g.pl;
    float v[8];
    if ((g.C & 7) == 0 && k0 < Kc) {
      // the 8 taps share (kh, kw) and are contiguous in x
      const int ci = k0 % g.C; const int tt = k0 / g.C;
      const int kw = tt % g.KW, kh = tt / g.KW;
      const int ih = ih0 + kh, iw = iw0 + kw;
      if ((unsigned)ih < (unsigned)g.H && (unsigned)iw < (unsigned)g.W) {
        const float4* src = reinterpret_cast<const float4*>(x + ((b * g.H + ih) * g.W + iw) * (long)g.C + ci);
        const float4 a = __ldg(src), c = __ldg(src + 1);
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = c.x; v[5] = c.y; v[6] = c.z; v[7] = c.w;
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = 0.f;
      }
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int k = k0 + j;
        float e = 0.f;
        if (k < Kc) {
          const int ci = k % g.C; const int tt = k / g.C;
          const int kw = tt % g.KW, kh = tt / g.KW;
          const int ih = ih0 + kh, iw = iw0 + kw;
          if ((unsigned)ih < (unsigned)g.H && (unsigned)iw < (unsigned)g.W)
            e = __ldg(x + ((b * g.H + ih) * g.W + iw) * (long)g.C + ci);
        }
        v[j] = e;
      }
    }
    uint4 o; o.x = pack2(v[0], v[1]); o.y = pack2(v[2], v[3]); o.z = pack2(v[4], v[5]); o.w = pack2(v[6], v[7]);
    *reinterpret_cast<uint4*>(out + m * ldo + k0) = o;
  }
}

// out[k, m] (ld = ldo >= M): 32x32 tiles through shared memory; reads coalesced along k, writes along m.
__global__ void im2colT_bf16_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ out, const CG g,
                                    long M, int Kc, long ldo) {
  TFDE_PDL_PROLOGUE();
  __shared__ float tile[32][33];
  const long m0 = (long)blockIdx.x * 32;
  const int k0 = blockIdx.y * 32;
  for (int r = threadIdx.y; r < 32; r += blockDim.y) {
    const long m = m0 + r;
    const int k = k0 + threadIdx.x;
    float e = 0.f;
    if (m < M && k < Kc) {
      const int ow = (int)(m % g.Wo); const long t = m / g.Wo;
      const int oh = (int)(t % g.Ho); const long b = t / g.Ho;
      const int ci = k % g.C; const int tt = k / g.C;
      const int kw = tt % g.KW, kh = tt / g.KW;
      const int ih = oh * g.sh - g.pt + kh, iw = ow * g.sw - g.pl + kw;
      if ((unsigned)ih < (unsigned)g.H && (unsigned)iw < (unsigned)g.W)
        e = __ldg(x + ((b * g.H + ih) * g.W + iw) * (long)g.C + ci);
    }
    tile[r][threadIdx.x] = e;
  }
  __syncthreads();
  for (int r = threadIdx.y; r < 32; r += blockDim.y) {
    const int k = k0 + r;
    const long m = m0 + threadIdx.x;
    if (k < Kc && m < ldo) out[(long)k * ldo + m] = __float2bfloat16(m < M ? tile[threadIdx.x][r] : 0.f);
  }
}

// out[(b,h,w), (kh,kw,co)] = dy[b, (h+pt-kh)/sh, (w+pl-kw)/sw, co] * (mask > 0); one thread = 8 k
__global__ void dgrad_gather_bf16_kernel(const float* __restrict__ dy, const float* __restrict__ mask,
                                         __nv_bfloat16* __restrict__ out, const CG g, long Min, int Kd, int ldo) {
  TFDE_PDL_PROLOGUE();
  const int k8n = ldo >> 3;
  const long total = Min * k8n;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const long m = i / k8n;
    const int k0 = (int)(i % k8n) * 8;
    const int w = (int)(m % g.W); const long t = m / g.W;
    const int h = (int)(t % g.H); const long b = t / g.H;
    float v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = 0.f;
    if ((g.Cout & 7) == 0 && k0 < Kd) {
      const int co = k0 % g.Cout; const int tt = k0 / g.Cout;
      const int kw = tt % g.KW, kh = tt / g.KW;
      const int th = h + g.pt - kh, tw = w + g.pl - kw;
      if (th >= 0 && tw >= 0) {
        const int oh = th / g.sh, ow = tw / g.sw;
        if (oh * g.sh == th && ow * g.sw == tw && oh < g.Ho && ow < g.Wo) {
          const long off = ((b * g.Ho + oh) * g.Wo + ow) * (long)g.Cout + co;
          const float4 a = __ldg(reinterpret_cast<const float4*>(dy + off)), c = __ldg(reinterpret_cast<const float4*>(dy + off) + 1);
          v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = c.x; v[5] = c.y; v[6] = c.z; v[7] = c.w;
          if (mask != nullptr) {
            const float4 ma = __ldg(reinterpret_cast<const float4*>(mask + off)), mc = __ldg(reinterpret_cast<const float4*>(mask + off) + 1);
            if (!(ma.x > 0.f)) v[0] = 0.f; if (!(ma.y > 0.f)) v[1] = 0.f; if (!(ma.z > 0.f)) v[2] = 0.f; if (!(ma.w > 0.f)) v[3] = 0.f;
            if (!(mc.x > 0.f)) v[4] = 0.f; if (!(mc.y > 0.f)) v[5] = 0.f; if (!(mc.z > 0.f)) v[6] = 0.f; if (!(mc.w > 0.f)) v[7] = 0.f;
          }
        }
      }
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int k = k0 + j;
        if (k >= Kd) continue;
        const int co = k % g.Cout; const int tt = k / g.Cout;
        const int kw = tt % g.KW, kh = tt / g.KW;
        const int th = h + g.pt - kh, tw = w + g.pl - kw;
        if (th < 0 || tw < 0) continue;
        const int oh = th / g.sh, ow = tw / g.sw;
        if (oh * g.sh != th || ow * g.sw != tw || oh >= g.Ho || ow >= g.Wo) continue;
        const long off = ((b * g.Ho + oh) * g.Wo + ow) * (long)g.Cout + co;
        float e = __ldg(dy + off);
        if (mask != nullptr && !(__ldg(mask + off) > 0.f)) e = 0.f;
        v[j] = e;
      }
    }
    uint4 o; o.x = pack2(v[0], v[1]); o.y = pack2(v[2], v[3]); o.z = pack2(v[4], v[5]); o.w = pack2(v[6], v[7]);
    *reinterpret_cast<uint4*>(out + m * ldo + k0) = o;
  }
}

// out[ci, (kh,kw,co)] = w[kh,kw,ci,co]  (ld = ldo), zero padded
__global__ void wrot_bf16_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ out, int KHW, int Cin,
                                 int Cout, int ldo) {
  TFDE_PDL_PROLOGUE();
  const long total = (long)Cin * ldo;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int k = (int)(i % ldo); const int ci = (int)(i / ldo);
    float e = 0.f;
    if (k < KHW * Cout) {
      const int co = k % Cout, t = k / Cout;
      e = __ldg(w + ((long)t * Cin + ci) * Cout + co);
    }
    out[i] = __float2bfloat16(e);
  }
}

// y[c, r] (ld = ldy >= R, zero padded) = bf16(x[r, c] * (mask > 0)) for fp32 x[R, C]
__global__ void cast_transpose_pad_bf16_kernel(const float* __restrict__ x, const float* __restrict__ mask,
                                               __nv_bfloat16* __restrict__ y, long R, int C, long ldy) {
  TFDE_PDL_PROLOGUE();
  __shared__ float tile[32][33];
  const int c0 = blockIdx.x * 32; const long r0 = (long)blockIdx.y * 32;
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const long r = r0 + i; const int c = c0 + threadIdx.x;
    float v = 0.f;
    if (r < R && c < C) {
      v = __ldg(x + r * C + c);
      if (mask != nullptr && !(__ldg(mask + r * C + c) > 0.f)) v = 0.f;
    }
    tile[i][threadIdx.x] = v;
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int c = c0 + i; const long r = r0 + threadIdx.x;
    if (c < C && r < ldy) y[(long)c * ldy + r] = __float2bfloat16(r < R ? tile[threadIdx.x][i] : 0.f);
  }
}

CG mk(int B, int H, int W, int C, int KH, int KW, int Ho, int Wo, int sh, int sw, int pt, int pl, int Cout) {
  CG g; g.B = B; g.H = H; g.W = W; g.C = C; g.KH = KH; g.KW = KW; g.Ho = Ho; g.Wo = Wo; g.sh = sh; g.sw = sw;
  g.pt = pt; g.pl = pl; g.Cout = Cout;
  return g;
}
int ew_blocks(long work) {
  long b = (work + 255) / 256;
  if (b > 148L * 16) b = 148L * 16;
  return (int)(b < 1 ? 1 : b);
}

}  // namespace

TFDE_API int tfde_im2col_bf16(const float* x, void* out, int B, int H, int W, int C, int KH, int KW, int Ho, int Wo,
                              int sh, int sw, int pt, int pl, int Cout, int ldo, cudaStream_t stream) {
  if (ldo & 7) return TFDE_ERR_INVALID;
  const long M = (long)B * Ho * Wo;
  tfde_launch(im2col_bf16_kernel, dim3(ew_blocks(M * (ldo >> 3))), dim3(256), 0, stream, x, (__nv_bfloat16*)out, mk(B, H, W, C, KH, KW, Ho, Wo, sh, sw, pt, pl, Cout), M, KH * KW * C, ldo);
  TFDE_RETURN_LAUNCH_STATUS();
}
TFDE_API int tfde_im2colT_bf16(const float* x, void* out, int B, int H, int W, int C, int KH, int KW, int Ho, int Wo,
                               int sh, int sw, int pt, int pl, int Cout, long ldo, cudaStream_t stream) {
  const long M = (long)B * Ho * Wo;
  const int Kc = KH * KW * C;
  if ((ldo & 7) || ldo < M) return TFDE_ERR_INVALID;
  dim3 grid((unsigned)((ldo + 31) / 32), (Kc + 31) / 32), block(32, 8);
  tfde_launch(im2colT_bf16_kernel, dim3(grid), dim3(block), 0, stream, x, (__nv_bfloat16*)out, mk(B, H, W, C, KH, KW, Ho, Wo, sh, sw, pt, pl, Cout), M, Kc, ldo);
  TFDE_RETURN_LAUNCH_STATUS();
}
TFDE_API int tfde_dgrad_gather_bf16(const float* dy, const float* mask, void* out, int B, int H, int W, int C, int KH,
                                    int KW, int Ho, int Wo, int sh, int sw, int pt, int pl, int Cout, int ldo,
                                    cudaStream_t stream) {
  if (ldo & 7) return TFDE_ERR_INVALID;
  const long Min = (long)B * H * W;
  tfde_launch(dgrad_gather_bf16_kernel, dim3(ew_blocks(Min * (ldo >> 3))), dim3(256), 0, stream, dy, mask, (__nv_bfloat16*)out, mk(B, H, W, C, KH, KW, Ho, Wo, sh, sw, pt, pl, Cout), Min, KH * KW * Cout, ldo);
  TFDE_RETURN_LAUNCH_STATUS();
}
TFDE_API int tfde_wrot_bf16(const float* w, void* out, int KH, int KW, int Cin, int Cout, int ldo, cudaStream_t stream) {
  if (ldo & 7) return TFDE_ERR_INVALID;
  tfde_launch(wrot_bf16_kernel, dim3(ew_blocks((long)Cin * ldo)), dim3(256), 0, stream, w, (__nv_bfloat16*)out, KH * KW, Cin, Cout, ldo);
  TFDE_RETURN_LAUNCH_STATUS();
}
TFDE_API int tfde_cast_transpose_pad_bf16(const float* x, const float* mask, void* y, long R, int C, long ldy,
                                          cudaStream_t stream) {
  if ((ldy & 7) || ldy < R) return TFDE_ERR_INVALID;
  dim3 grid((C + 31) / 32, (unsigned)((ldy + 31) / 32)), block(32, 8);
  tfde_launch(cast_transpose_pad_bf16_kernel, dim3(grid), dim3(block), 0, stream, x, mask, (__nv_bfloat16*)y, R, C, ldy);
  TFDE_RETURN_LAUNCH_STATUS();
}

// ---------------------------------------------------------------------------------------------
// bf16 matrix transpose y[C, R] (ld = ldy) = x[R, C] (ld = ldx): 64x64 tiles through shared memory, 4-byte
// accesses on both sides. Used to turn the (fast, vectorised) im2col operand into the K-major operand of
// the filter-gradient GEMM instead of gathering it a second time in transposed order.
// ---------------------------------------------------------------------------------------------
namespace {
__global__ void __launch_bounds__(256)
transpose_bf16_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y, long R, int C, long ldx,
                      long ldy) {
  TFDE_PDL_PROLOGUE();
  __shared__ __nv_bfloat16 tile[64][66];
  const long r0 = (long)blockIdx.x * 64;
  const int c0 = blockIdx.y * 64;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;       // 32 x 8
  for (int i = ty; i < 64; i += 8) {                            // row i of the tile, columns 2*tx, 2*tx+1
    const long r = r0 + i;
    const int c = c0 + 2 * tx;
    __nv_bfloat162 v = __floats2bfloat162_rn(0.f, 0.f);
    if (r < R) {
      if (c + 1 < C) v = *reinterpret_cast<const __nv_bfloat162*>(x + r * ldx + c);
      else if (c < C) v.x = x[r * ldx + c];
    }
    tile[i][2 * tx] = v.x; tile[i][2 * tx + 1] = v.y;
  }
  __syncthreads();
  for (int i = ty; i < 64; i += 8) {                            // output row c0+i, columns r0 + 2*tx, +1
    const int c = c0 + i;
    const long r = r0 + 2 * tx;
    if (c < C && r < ldy) {
      __nv_bfloat162 v;
      v.x = tile[2 * tx][i]; v.y = tile[2 * tx + 1][i];
      if (r + 1 < ldy) *reinterpret_cast<__nv_bfloat162*>(y + (long)c * ldy + r) = v;
      else y[(long)c * ldy + r] = v.x;
    }
  }
}
}  // namespace

// y[C, ldy] = transpose(x[R, ldx]); rows of x beyond R read as zero (ldy may exceed R: zero padding).
TFDE_API int tfde_transpose_bf16(const void* x, void* y, long R, int C, long ldx, long ldy, cudaStream_t stream) {
  if ((ldx & 1) || (ldy & 1) || ldy < R) return TFDE_ERR_INVALID;
  dim3 grid((unsigned)((ldy + 63) / 64), (C + 63) / 64);
  tfde_launch(transpose_bf16_kernel, dim3(grid), dim3(256), 0, stream, (const __nv_bfloat16*)x, (__nv_bfloat16*)y, R, C, ldx, ldy);
  TFDE_RETURN_LAUNCH_STATUS();
}
