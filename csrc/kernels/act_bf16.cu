// bf16-activation kernels (channels-last, C % 8 == 0): every tensor between two layers is bf16, every
// thread moves 16 bytes (8 channels) per access, all arithmetic and all reductions are fp32/fp64.
// This is the activation side of the `mixed_bfloat16` policy for convolutional networks (ResNet-50 config
// of BASELINE.json): batch-norm (train forward / backward with the fused-ReLU mask RECOMPUTED from x instead
// of re-reading y), residual add(+ReLU), ReLU backward, max-pool, global average pool, and the bf16->bf16
// operand gathers (im2col, dgrad gather) that feed the tcgen05 GEMMs. Compared with the fp32 kernels of
// nn_ops.cu the batch-norm backward moves 6 bytes per element instead of 16.
// Reference ops replaced: BatchNormalization / Activation / MaxPooling2D / Conv2D of
// mnist_keras_distributed.py:83-104 under a bf16 compute policy (Keras mixed precision semantics:
// bf16 activations, fp32 variables).
#include "common.cuh"
#include <cuda_bf16.h>

namespace {

struct F8 { float v[8]; };

__device__ __forceinline__ F8 unpack8(const uint4 r) {
  F8 o;
  o.v[0] = __uint_as_float(r.x << 16); o.v[1] = __uint_as_float(r.x & 0xffff0000u);
  o.v[2] = __uint_as_float(r.y << 16); o.v[3] = __uint_as_float(r.y & 0xffff0000u);
  o.v[4] = __uint_as_float(r.z << 16); o.v[5] = __uint_as_float(r.z & 0xffff0000u);
  o.v[6] = __uint_as_float(r.w << 16); o.v[7] = __uint_as_float(r.w & 0xffff0000u);
  return o;
}
__device__ __forceinline__ uint32_t pack2(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ uint4 pack8(const F8& f) {
  uint4 o;
  o.x = pack2(f.v[0], f.v[1]); o.y = pack2(f.v[2], f.v[3]); o.z = pack2(f.v[4], f.v[5]); o.w = pack2(f.v[6], f.v[7]);
  return o;
}
__device__ __forceinline__ F8 load8f(const float* p) {
  const float4 a = __ldg(reinterpret_cast<const float4*>(p)), b = __ldg(reinterpret_cast<const float4*>(p) + 1);
  F8 o; o.v[0] = a.x; o.v[1] = a.y; o.v[2] = a.z; o.v[3] = a.w; o.v[4] = b.x; o.v[5] = b.y; o.v[6] = b.z; o.v[7] = b.w;
  return o;
}
__device__ __forceinline__ uint4 ldg16(const __nv_bfloat16* p) { return __ldg(reinterpret_cast<const uint4*>(p)); }

// ---------------------------------------------------------------------------------------------------
// Column statistics over x[M, C]: block = (TPR threads across 8*TPR channels) x (256/TPR rows).
// grid = (C / (8*TPR), row blocks). Per-block partials -> last block of each channel tile finalises in
// fp64 (fixed order => deterministic). MODE 0: (sum x, sum x^2) -> batch-norm forward statistics.
// MODE 1: (sum g, sum g*xhat) with g = dy * [relu mask recomputed from x] -> batch-norm backward.
// MODE 2: (sum x, -) -> plain column sum (bias gradient).
// ---------------------------------------------------------------------------------------------------
struct StatArgs {
  const __nv_bfloat16* x; const __nv_bfloat16* dy;
  long M; int C; int tpr;
  float* partials; unsigned* counters;
  // MODE 0 outputs
  float* mean; float* invstd; float* scale; float* shift; float* moving_mean; float* moving_var;
  const float* gamma; const float* beta; float momentum, eps; int unbiased;
  // MODE 1 inputs/outputs
  const float* fscale; const float* fshift; int relu; float* dbeta; float* dgamma;
  // MODE 2 output
  float* colsum;
};

// One fat block per SM (512 threads): the per-block partials are few (<= 148 per channel tile), so the
// last block's finalise is a couple of load rounds instead of the dominant tail of a 5-microsecond kernel.
constexpr int ST = 512;

template <int MODE>
__global__ void __launch_bounds__(ST, (MODE == 1) ? 1 : 2) colstats_bf16_kernel(const StatArgs a) {
  // MODE 0/2: two blocks per SM (<= 64 registers, no spills); MODE 1 keeps 4 more per-channel vectors live and
  // gets a whole SM (the host sizes the grid accordingly)
  TFDE_PDL_PROLOGUE();
  extern __shared__ float smem[];   // [2][rows][8*tpr]
  __shared__ int s_last;
  const int tpr = a.tpr, rows = ST / tpr, cw = 8 * tpr;
  const int tx = threadIdx.x % tpr, ty = threadIdx.x / tpr;
  const int c0 = blockIdx.x * cw + tx * 8;
  float s[8], q[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) { s[j] = 0.f; q[j] = 0.f; }
  F8 mu, iv, fs, fb;
  if (MODE == 1) {
    mu = load8f(a.mean + c0); iv = load8f(a.invstd + c0);
    if (a.relu) { fs = load8f(a.fscale + c0); fb = load8f(a.fshift + c0); }
  }
  const long stride = (long)gridDim.y * rows;
  long m = (long)blockIdx.y * rows + ty;
  auto accumulate = [&](const uint4 xr, const uint4 dr) {
    const F8 xv = unpack8(xr);
    if (MODE == 1) {
      const F8 g = unpack8(dr);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float gj = g.v[j];
        if (a.relu && !(fmaf(xv.v[j], fs.v[j], fb.v[j]) > 0.f)) gj = 0.f;
        s[j] += gj; q[j] = fmaf(gj, (xv.v[j] - mu.v[j]) * iv.v[j], q[j]);
      }
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) { s[j] += xv.v[j]; if (MODE == 0) q[j] = fmaf(xv.v[j], xv.v[j], q[j]); }
    }
  };
  // U rows (16 bytes each, 32 with dy) in flight per thread: a pure streaming reduction is bound by
  // memory-level parallelism (measured: 4 rows x 148 blocks reached 2.8 of 6.5 TB/s)
  constexpr int U = (MODE == 1) ? 4 : 8;
  for (; m + (U - 1) * stride < a.M; m += U * stride) {
    uint4 xr[U], dr[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      xr[u] = ldg16(a.x + (m + u * stride) * a.C + c0);
      dr[u] = (MODE == 1) ? ldg16(a.dy + (m + u * stride) * a.C + c0) : make_uint4(0u, 0u, 0u, 0u);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) accumulate(xr[u], dr[u]);
  }
  for (; m < a.M; m += stride)
    accumulate(ldg16(a.x + m * a.C + c0), (MODE == 1) ? ldg16(a.dy + m * a.C + c0) : make_uint4(0u, 0u, 0u, 0u));
  float* sa = smem; float* sb = smem + rows * cw;
#pragma unroll
  for (int j = 0; j < 8; ++j) { sa[ty * cw + tx * 8 + j] = s[j]; sb[ty * cw + tx * 8 + j] = q[j]; }
  __syncthreads();
  float* part = a.partials + ((size_t)blockIdx.x * gridDim.y + blockIdx.y) * 2 * cw;
  for (int c = threadIdx.x; c < cw; c += ST) {
    float ts = 0.f, tq = 0.f;
    for (int r = 0; r < rows; ++r) { ts += sa[r * cw + c]; tq += sb[r * cw + c]; }
    __stcg(part + c, ts); __stcg(part + cw + c, tq);
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_last = (atomicAdd(a.counters + blockIdx.x, 1u) == gridDim.y - 1);
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  // Finalise with the whole block: 256/cw thread groups each sum an interleaved subset of the row-block
  // partials (4 loads in flight), then the groups are combined in fixed order through shared memory.
  // Each thread owns 4 channels (one 16-byte load per partial row and statistic); 4*ST/cw thread groups walk
  // interleaved subsets of the row blocks with 4 rows in flight: <= 5 dependent load rounds for any shape.
  const int groups = 4 * ST / cw;
  {
    const int gi = threadIdx.x / (cw >> 2), c4 = (threadIdx.x % (cw >> 2)) << 2;
    double ds[4] = {0.0, 0.0, 0.0, 0.0}, dq[4] = {0.0, 0.0, 0.0, 0.0};
    const float* base = a.partials + (size_t)blockIdx.x * gridDim.y * 2 * cw;
    unsigned rb = gi;
    for (; rb + 3 * groups < gridDim.y; rb += 4 * groups) {
      float4 vs[4], vq[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const float* pp = base + (size_t)(rb + u * groups) * 2 * cw;
        vs[u] = __ldcg(reinterpret_cast<const float4*>(pp + c4)); vq[u] = __ldcg(reinterpret_cast<const float4*>(pp + cw + c4));
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        ds[0] += (double)vs[u].x; ds[1] += (double)vs[u].y; ds[2] += (double)vs[u].z; ds[3] += (double)vs[u].w;
        dq[0] += (double)vq[u].x; dq[1] += (double)vq[u].y; dq[2] += (double)vq[u].z; dq[3] += (double)vq[u].w;
      }
    }
    for (; rb < gridDim.y; rb += groups) {
      const float* pp = base + (size_t)rb * 2 * cw;
      const float4 v = __ldcg(reinterpret_cast<const float4*>(pp + c4)), w = __ldcg(reinterpret_cast<const float4*>(pp + cw + c4));
      ds[0] += (double)v.x; ds[1] += (double)v.y; ds[2] += (double)v.z; ds[3] += (double)v.w;
      dq[0] += (double)w.x; dq[1] += (double)w.y; dq[2] += (double)w.z; dq[3] += (double)w.w;
    }
    double* dsm = reinterpret_cast<double*>(smem);      // [2][groups][cw] doubles = 2 * 4*ST * 8 B = 32 KB
    __syncthreads();
#pragma unroll
    for (int j = 0; j < 4; ++j) { dsm[gi * cw + c4 + j] = ds[j]; dsm[groups * cw + gi * cw + c4 + j] = dq[j]; }
    __syncthreads();
  }
  for (int cl = threadIdx.x; cl < cw; cl += ST) {
    const int c = blockIdx.x * cw + cl;
    double ds = 0.0, dq = 0.0;
    {
      const double* dsm = reinterpret_cast<const double*>(smem);
      for (int gi = 0; gi < groups; ++gi) { ds += dsm[gi * cw + cl]; dq += dsm[groups * cw + gi * cw + cl]; }
    }
    if (MODE == 0) {
      const double mean = ds / (double)a.M;
      double var = dq / (double)a.M - mean * mean;
      if (var < 0.0) var = 0.0;
      const float inv = rsqrtf((float)var + a.eps);
      a.mean[c] = (float)mean; a.invstd[c] = inv;
      const float sc = (a.gamma ? a.gamma[c] : 1.f) * inv;
      a.scale[c] = sc; a.shift[c] = (a.beta ? a.beta[c] : 0.f) - (float)mean * sc;
      const double vm = (a.unbiased && a.M > 1) ? var * ((double)a.M / (double)(a.M - 1)) : var;
      a.moving_mean[c] = a.moving_mean[c] * a.momentum + (float)mean * (1.f - a.momentum);
      a.moving_var[c] = a.moving_var[c] * a.momentum + (float)vm * (1.f - a.momentum);
    } else if (MODE == 1) {
      a.dbeta[c] = (float)ds; a.dgamma[c] = (float)dq;
    } else {
      a.colsum[c] = (float)ds;
    }
  }
  if (threadIdx.x == 0) a.counters[blockIdx.x] = 0u;
}

// ---------------------------------------------------------------------------------------------------
// Batch-norm statistics from the partial column sums a GEMM epilogue left behind (psum / psq: fp32 [G][C],
// one row per 32 output rows; gemm_tcgen05_persistent.cu). Same two-phase deterministic scheme as
// colstats_bf16_kernel, over 1/8 of the bytes of the tensor itself; finalises exactly like MODE 0.
// block = 256 threads = (cw/4 lanes of 4 channels) x (1024/cw rows); grid = (C/cw, row blocks).
// ---------------------------------------------------------------------------------------------------
struct PartArgs {
  const float* psum; const float* psq; int G; long M; int C; int cw;
  float* partials; unsigned* counters;
  float* mean; float* invstd; float* scale; float* shift; float* moving_mean; float* moving_var;
  const float* gamma; const float* beta; float momentum, eps; int unbiased;
};

__global__ void __launch_bounds__(256) bn_finalize_partials_kernel(const PartArgs a) {
  TFDE_PDL_PROLOGUE();
  __shared__ float4 sa[256], sb[256];
  __shared__ double dsa[1024], dsb[1024];
  __shared__ int s_last;
  const int cw = a.cw, lanes = cw >> 2, rows = 256 / lanes;
  const int tx = threadIdx.x % lanes, ty = threadIdx.x / lanes;
  const int c0 = blockIdx.x * cw + tx * 4;
  float4 s = make_float4(0.f, 0.f, 0.f, 0.f), q = s;
  const int stride = gridDim.y * rows;
  int g = blockIdx.y * rows + ty;
  for (; g + 3 * stride < a.G; g += 4 * stride) {
    float4 vs[4], vq[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      vs[u] = __ldg(reinterpret_cast<const float4*>(a.psum + (size_t)(g + u * stride) * a.C + c0));
      vq[u] = __ldg(reinterpret_cast<const float4*>(a.psq + (size_t)(g + u * stride) * a.C + c0));
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      s.x += vs[u].x; s.y += vs[u].y; s.z += vs[u].z; s.w += vs[u].w;
      q.x += vq[u].x; q.y += vq[u].y; q.z += vq[u].z; q.w += vq[u].w;
    }
  }
  for (; g < a.G; g += stride) {
    const float4 v = __ldg(reinterpret_cast<const float4*>(a.psum + (size_t)g * a.C + c0));
    const float4 w = __ldg(reinterpret_cast<const float4*>(a.psq + (size_t)g * a.C + c0));
    s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w; q.x += w.x; q.y += w.y; q.z += w.z; q.w += w.w;
  }
  sa[threadIdx.x] = s; sb[threadIdx.x] = q;
  __syncthreads();
  float* part = a.partials + ((size_t)blockIdx.x * gridDim.y + blockIdx.y) * 2 * cw;
  if (ty == 0) {
    for (int r = 1; r < rows; ++r) {
      const float4 v = sa[r * lanes + tx], w = sb[r * lanes + tx];
      s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w; q.x += w.x; q.y += w.y; q.z += w.z; q.w += w.w;
    }
    __stcg(reinterpret_cast<float4*>(part + tx * 4), s);
    __stcg(reinterpret_cast<float4*>(part + cw + tx * 4), q);
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_last = (atomicAdd(a.counters + blockIdx.x, 1u) == gridDim.y - 1);
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  {
    double ds[4] = {0.0, 0.0, 0.0, 0.0}, dq[4] = {0.0, 0.0, 0.0, 0.0};
    const float* base = a.partials + (size_t)blockIdx.x * gridDim.y * 2 * cw;
    for (unsigned rb = ty; rb < gridDim.y; rb += rows) {
      const float* pp = base + (size_t)rb * 2 * cw;
      const float4 v = __ldcg(reinterpret_cast<const float4*>(pp + tx * 4));
      const float4 w = __ldcg(reinterpret_cast<const float4*>(pp + cw + tx * 4));
      ds[0] += (double)v.x; ds[1] += (double)v.y; ds[2] += (double)v.z; ds[3] += (double)v.w;
      dq[0] += (double)w.x; dq[1] += (double)w.y; dq[2] += (double)w.z; dq[3] += (double)w.w;
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) { dsa[ty * cw + tx * 4 + j] = ds[j]; dsb[ty * cw + tx * 4 + j] = dq[j]; }
    __syncthreads();
  }
  for (int cl = threadIdx.x; cl < cw; cl += 256) {
    const int c = blockIdx.x * cw + cl;
    double ds = 0.0, dq = 0.0;
    for (int r = 0; r < rows; ++r) { ds += dsa[r * cw + cl]; dq += dsb[r * cw + cl]; }
    const double mean = ds / (double)a.M;
    double var = dq / (double)a.M - mean * mean;
    if (var < 0.0) var = 0.0;
    const float inv = rsqrtf((float)var + a.eps);
    a.mean[c] = (float)mean; a.invstd[c] = inv;
    const float sc = (a.gamma ? a.gamma[c] : 1.f) * inv;
    a.scale[c] = sc; a.shift[c] = (a.beta ? a.beta[c] : 0.f) - (float)mean * sc;
    const double vm = (a.unbiased && a.M > 1) ? var * ((double)a.M / (double)(a.M - 1)) : var;
    a.moving_mean[c] = a.moving_mean[c] * a.momentum + (float)mean * (1.f - a.momentum);
    a.moving_var[c] = a.moving_var[c] * a.momentum + (float)vm * (1.f - a.momentum);
  }
  if (threadIdx.x == 0) a.counters[blockIdx.x] = 0u;
}

// y = relu?(x * scale + shift)   (scale = gamma*invstd, shift = beta - mean*scale)
__global__ void bn_apply_bf16_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y, long n8,
                                     int C, const float* __restrict__ scale, const float* __restrict__ shift,
                                     int relu) {
  TFDE_PDL_PROLOGUE();
  const long T = (long)gridDim.x * blockDim.x;
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  const int c8n = C >> 3;
  if (T % c8n == 0) {
    // the grid stride is a multiple of the row length: this thread sees ONE channel group for its whole
    // life, so the per-channel constants are loaded once (no modulo, no parameter loads in the loop)
    const int c = (int)(i % c8n) << 3;
    const F8 sc = load8f(scale + c), sh = load8f(shift + c);
    for (; i + T < n8; i += 2 * T) {          // two independent 16-byte loads in flight
      const uint4 r0 = ldg16(x + (i << 3)), r1 = ldg16(x + ((i + T) << 3));
      const F8 x0 = unpack8(r0), x1 = unpack8(r1);
      F8 o0, o1;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float v0 = fmaf(x0.v[j], sc.v[j], sh.v[j]), v1 = fmaf(x1.v[j], sc.v[j], sh.v[j]);
        o0.v[j] = relu ? fmaxf(v0, 0.f) : v0; o1.v[j] = relu ? fmaxf(v1, 0.f) : v1;
      }
      *reinterpret_cast<uint4*>(y + (i << 3)) = pack8(o0);
      *reinterpret_cast<uint4*>(y + ((i + T) << 3)) = pack8(o1);
    }
    for (; i < n8; i += T) {
      const F8 xv = unpack8(ldg16(x + (i << 3)));
      F8 o;
#pragma unroll
      for (int j = 0; j < 8; ++j) { const float v = fmaf(xv.v[j], sc.v[j], sh.v[j]); o.v[j] = relu ? fmaxf(v, 0.f) : v; }
      *reinterpret_cast<uint4*>(y + (i << 3)) = pack8(o);
    }
    return;
  }
  for (; i < n8; i += T) {
    const int c = (int)((i << 3) % C);
    const F8 xv = unpack8(ldg16(x + (i << 3)));
    const F8 sc = load8f(scale + c), sh = load8f(shift + c);
    F8 o;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float v = fmaf(xv.v[j], sc.v[j], sh.v[j]);
      o.v[j] = relu ? fmaxf(v, 0.f) : v;
    }
    *reinterpret_cast<uint4*>(y + (i << 3)) = pack8(o);
  }
}

// dx = gamma*invstd * (g - dbeta/M - xhat*dgamma/M), g = dy * [x*scale+shift > 0] when the ReLU was fused
__global__ void bn_bwd_apply_bf16_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ dy,
                                         __nv_bfloat16* __restrict__ dx, long n8, long M, int C,
                                         const float* __restrict__ mean, const float* __restrict__ invstd,
                                         const float* __restrict__ gamma, const float* __restrict__ scale,
                                         const float* __restrict__ shift, const float* __restrict__ dbeta,
                                         const float* __restrict__ dgamma, int relu) {
  TFDE_PDL_PROLOGUE();
  const float invM = 1.f / (float)M;
  const long T = (long)gridDim.x * blockDim.x;
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  const int c8n = C >> 3;
  if (T % c8n == 0) {
    // one channel group per thread for its whole life (see bn_apply_bf16_kernel): fold the per-channel
    // terms once:  dx = A*(g - B - (x - mu)*D),  A = gamma*invstd, B = dbeta/M, D = invstd*dgamma/M
    const int c = (int)(i % c8n) << 3;
    const F8 mu = load8f(mean + c), iv = load8f(invstd + c), db = load8f(dbeta + c), dg = load8f(dgamma + c);
    F8 A, Bc, D, sc, sh;
    if (gamma) A = load8f(gamma + c);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      A.v[j] = (gamma ? A.v[j] : 1.f) * iv.v[j]; Bc.v[j] = invM * db.v[j]; D.v[j] = invM * dg.v[j] * iv.v[j];
    }
    if (relu) { sc = load8f(scale + c); sh = load8f(shift + c); }
    auto one = [&](const uint4 xr, const uint4 gr) {
      const F8 xv = unpack8(xr), g = unpack8(gr);
      F8 o;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float gj = g.v[j];
        if (relu && !(fmaf(xv.v[j], sc.v[j], sh.v[j]) > 0.f)) gj = 0.f;
        o.v[j] = A.v[j] * (gj - Bc.v[j] - (xv.v[j] - mu.v[j]) * D.v[j]);
      }
      return pack8(o);
    };
    for (; i + T < n8; i += 2 * T) {          // four independent 16-byte loads in flight
      const uint4 x0 = ldg16(x + (i << 3)), g0 = ldg16(dy + (i << 3));
      const uint4 x1 = ldg16(x + ((i + T) << 3)), g1 = ldg16(dy + ((i + T) << 3));
      *reinterpret_cast<uint4*>(dx + (i << 3)) = one(x0, g0);
      *reinterpret_cast<uint4*>(dx + ((i + T) << 3)) = one(x1, g1);
    }
    for (; i < n8; i += T) *reinterpret_cast<uint4*>(dx + (i << 3)) = one(ldg16(x + (i << 3)), ldg16(dy + (i << 3)));
    return;
  }
  for (; i < n8; i += T) {
    const int c = (int)((i << 3) % C);
    const F8 xv = unpack8(ldg16(x + (i << 3))), g = unpack8(ldg16(dy + (i << 3)));
    const F8 mu = load8f(mean + c), iv = load8f(invstd + c), db = load8f(dbeta + c), dg = load8f(dgamma + c);
    F8 gm, sc, sh, o;
    if (gamma) gm = load8f(gamma + c);
    if (relu) { sc = load8f(scale + c); sh = load8f(shift + c); }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float gj = g.v[j];
      if (relu && !(fmaf(xv.v[j], sc.v[j], sh.v[j]) > 0.f)) gj = 0.f;
      const float xhat = (xv.v[j] - mu.v[j]) * iv.v[j];
      o.v[j] = (gamma ? gm.v[j] : 1.f) * iv.v[j] * (gj - invM * (db.v[j] + xhat * dg.v[j]));
    }
    *reinterpret_cast<uint4*>(dx + (i << 3)) = pack8(o);
  }
}

// y = relu?(a + b)
__global__ void add_bf16_kernel(const __nv_bfloat16* __restrict__ a, const __nv_bfloat16* __restrict__ b,
                                __nv_bfloat16* __restrict__ y, long n8, int relu) {
  TFDE_PDL_PROLOGUE();
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (long)gridDim.x * blockDim.x) {
    const F8 av = unpack8(ldg16(a + (i << 3))), bv = unpack8(ldg16(b + (i << 3)));
    F8 o;
#pragma unroll
    for (int j = 0; j < 8; ++j) { const float v = av.v[j] + bv.v[j]; o.v[j] = relu ? fmaxf(v, 0.f) : v; }
    *reinterpret_cast<uint4*>(y + (i << 3)) = pack8(o);
  }
}

// dx = dy * (y > 0); with `addend` != nullptr: dx = addend + dy * (y > 0)
__global__ void relu_bwd_bf16_kernel(const __nv_bfloat16* __restrict__ y, const __nv_bfloat16* __restrict__ dy,
                                     const __nv_bfloat16* __restrict__ addend, __nv_bfloat16* __restrict__ dx,
                                     long n8) {
  TFDE_PDL_PROLOGUE();
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (long)gridDim.x * blockDim.x) {
    const F8 yv = unpack8(ldg16(y + (i << 3))), g = unpack8(ldg16(dy + (i << 3)));
    F8 o;
    if (addend) o = unpack8(ldg16(addend + (i << 3)));
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float v = yv.v[j] > 0.f ? g.v[j] : 0.f;
      o.v[j] = addend ? o.v[j] + v : v;
    }
    *reinterpret_cast<uint4*>(dx + (i << 3)) = pack8(o);
  }
}

// fp32 -> bf16 and back (network boundary: image in, pooled features out)
__global__ void cast_f32_bf16_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ y, long n) {
  TFDE_PDL_PROLOGUE();
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x)
    y[i] = __float2bfloat16(__ldg(x + i));
}

// ---------------------------------------------------------------------------------------------------
// Max-pool (NHWC), 8 channels per thread. idx = kh*PW + kw of the first maximum.
// ---------------------------------------------------------------------------------------------------
struct PG { int Bn, H, W, C, PH, PW, sh, sw, pt, pl, Ho, Wo; };

__global__ void maxpool_fwd_bf16_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y,
                                        uint8_t* __restrict__ idx, const PG g) {
  TFDE_PDL_PROLOGUE();
  const int c8n = g.C >> 3;
  const long total = (long)g.Bn * g.Ho * g.Wo * c8n;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int c = (int)(i % c8n) * 8; long t = i / c8n;
    const int ow = (int)(t % g.Wo); t /= g.Wo;
    const int oh = (int)(t % g.Ho); const long b = t / g.Ho;
    float best[8]; int bi[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { best[j] = -INFINITY; bi[j] = 0; }
    for (int kh = 0; kh < g.PH; ++kh) {
      const int ih = oh * g.sh - g.pt + kh;
      if ((unsigned)ih >= (unsigned)g.H) continue;
      for (int kw = 0; kw < g.PW; ++kw) {
        const int iw = ow * g.sw - g.pl + kw;
        if ((unsigned)iw >= (unsigned)g.W) continue;
        const F8 v = unpack8(ldg16(x + ((b * g.H + ih) * g.W + iw) * (long)g.C + c));
#pragma unroll
        for (int j = 0; j < 8; ++j) if (v.v[j] > best[j]) { best[j] = v.v[j]; bi[j] = kh * g.PW + kw; }
      }
    }
    F8 o;
#pragma unroll
    for (int j = 0; j < 8; ++j) o.v[j] = best[j];
    const long off = ((b * g.Ho + oh) * g.Wo + ow) * (long)g.C + c;
    *reinterpret_cast<uint4*>(y + off) = pack8(o);
    uint2 ib;
    ib.x = (uint32_t)bi[0] | ((uint32_t)bi[1] << 8) | ((uint32_t)bi[2] << 16) | ((uint32_t)bi[3] << 24);
    ib.y = (uint32_t)bi[4] | ((uint32_t)bi[5] << 8) | ((uint32_t)bi[6] << 16) | ((uint32_t)bi[7] << 24);
    *reinterpret_cast<uint2*>(idx + off) = ib;
  }
}

// Gather formulation (no atomics): each input position sums the windows whose arg-max it is.
__global__ void maxpool_bwd_bf16_kernel(const __nv_bfloat16* __restrict__ dy, const uint8_t* __restrict__ idx,
                                        __nv_bfloat16* __restrict__ dx, const PG g) {
  TFDE_PDL_PROLOGUE();
  const int c8n = g.C >> 3;
  const long total = (long)g.Bn * g.H * g.W * c8n;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int c = (int)(i % c8n) * 8; long t = i / c8n;
    const int w = (int)(t % g.W); t /= g.W;
    const int h = (int)(t % g.H); const long b = t / g.H;
    F8 s;
#pragma unroll
    for (int j = 0; j < 8; ++j) s.v[j] = 0.f;
    const int hh = h + g.pt, ww = w + g.pl;
    int oh_lo = (hh - g.PH + g.sh) / g.sh; if (hh - g.PH + 1 <= 0) oh_lo = 0;
    int ow_lo = (ww - g.PW + g.sw) / g.sw; if (ww - g.PW + 1 <= 0) ow_lo = 0;
    const int oh_hi = min(hh / g.sh, g.Ho - 1), ow_hi = min(ww / g.sw, g.Wo - 1);
    if (g.PH <= 2 * g.sh && g.PW <= 2 * g.sw) {
      // at most 2 x 2 windows cover an input position (the 3x3 / stride-2 pool of the ResNet stem, 2x2 / 2 pools):
      // all eight loads are issued before the first compare instead of one dependent pair per loop trip
      uint2 ib[4]; uint4 dv[4]; uint32_t want[4];
#pragma unroll
      for (int a = 0; a < 4; ++a) {
        const int oh = oh_lo + (a >> 1), ow = ow_lo + (a & 1);
        const int kh = hh - oh * g.sh, kw = ww - ow * g.sw;
        const bool ok = oh <= oh_hi && ow <= ow_hi && kh >= 0 && kh < g.PH && kw >= 0 && kw < g.PW;
        const long o = ((b * g.Ho + oh) * g.Wo + ow) * (long)g.C + c;
        want[a] = ok ? (uint32_t)(kh * g.PW + kw) : 0xffffffffu;      // never equals a stored arg-max byte
        ib[a] = ok ? __ldg(reinterpret_cast<const uint2*>(idx + o)) : make_uint2(0u, 0u);
        dv[a] = ok ? ldg16(dy + o) : make_uint4(0u, 0u, 0u, 0u);
      }
#pragma unroll
      for (int a = 0; a < 4; ++a) {
        const F8 d = unpack8(dv[a]);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const uint32_t got = ((j < 4 ? ib[a].x : ib[a].y) >> (8 * (j & 3))) & 0xffu;
          if (got == want[a]) s.v[j] += d.v[j];
        }
      }
      *reinterpret_cast<uint4*>(dx + ((b * g.H + h) * g.W + w) * (long)g.C + c) = pack8(s);
      continue;
    }
    for (int oh = oh_lo; oh <= oh_hi; ++oh) {
      const int kh = hh - oh * g.sh;
      if (kh < 0 || kh >= g.PH) continue;
      for (int ow = ow_lo; ow <= ow_hi; ++ow) {
        const int kw = ww - ow * g.sw;
        if (kw < 0 || kw >= g.PW) continue;
        const long o = ((b * g.Ho + oh) * g.Wo + ow) * (long)g.C + c;
        const uint2 ib = __ldg(reinterpret_cast<const uint2*>(idx + o));
        const F8 d = unpack8(ldg16(dy + o));
        const uint32_t want = (uint32_t)(kh * g.PW + kw);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const uint32_t got = ((j < 4 ? ib.x : ib.y) >> (8 * (j & 3))) & 0xffu;
          if (got == want) s.v[j] += d.v[j];
        }
      }
    }
    *reinterpret_cast<uint4*>(dx + ((b * g.H + h) * g.W + w) * (long)g.C + c) = pack8(s);
  }
}

// Global average pool: bf16 [B, HW, C] -> fp32 [B, C]; backward fp32 [B, C] -> bf16 [B, HW, C].
__global__ void gap_fwd_bf16_kernel(const __nv_bfloat16* __restrict__ x, float* __restrict__ y, int Bn, int HW, int C) {
  TFDE_PDL_PROLOGUE();
  const int c8n = C >> 3;
  const long total = (long)Bn * c8n;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int c = (int)(i % c8n) * 8; const long b = i / c8n;
    float s[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) s[j] = 0.f;
    for (int t = 0; t < HW; ++t) {
      const F8 v = unpack8(ldg16(x + (b * HW + t) * (long)C + c));
#pragma unroll
      for (int j = 0; j < 8; ++j) s[j] += v.v[j];
    }
    const float inv = 1.f / (float)HW;
#pragma unroll
    for (int j = 0; j < 8; ++j) y[b * C + c + j] = s[j] * inv;
  }
}
__global__ void gap_bwd_bf16_kernel(const float* __restrict__ dy, __nv_bfloat16* __restrict__ dx, int Bn, int HW, int C) {
  TFDE_PDL_PROLOGUE();
  const int c8n = C >> 3;
  const long total = (long)Bn * HW * c8n;
  const float inv = 1.f / (float)HW;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int c = (int)(i % c8n) * 8; const long b = i / ((long)HW * c8n);
    F8 v = load8f(dy + b * C + c);
#pragma unroll
    for (int j = 0; j < 8; ++j) v.v[j] *= inv;
    *reinterpret_cast<uint4*>(dx + (i << 3)) = pack8(v);
  }
}

// ---------------------------------------------------------------------------------------------------
// bf16 -> bf16 operand gathers for the tcgen05 GEMMs (C % 8 == 0: the 8 taps of one thread are one
// 16-byte load). Same index maps as conv_gather.cu.
// ---------------------------------------------------------------------------------------------------
struct CG { int B, H, W, C, KH, KW, Ho, Wo, sh, sw, pt, pl, Cout; };

__global__ void im2col_b2b_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ out, const CG g,
                                  long M, int Kc, int ldo) {
  TFDE_PDL_PROLOGUE();
  // One warp per output row m: the (b, oh, ow) decomposition is warp-uniform and done once in 32-bit
  // arithmetic; lanes walk the row's 16-byte groups (the version with one flat 64-bit index per 16 bytes
  // spent its time in emulated 64-bit divisions: 1.8 TB/s).
  const unsigned k8n = (unsigned)ldo >> 3;
  const unsigned lane = threadIdx.x & 31u;
  const unsigned warps = (gridDim.x * blockDim.x) >> 5;
  const unsigned Mu = (unsigned)M, Wo = (unsigned)g.Wo, Ho = (unsigned)g.Ho, Cu = (unsigned)g.C, KW = (unsigned)g.KW;
  // a lane visits the same 16-byte groups of every row: decode the first NIP of them (kh, kw, ci) once
  constexpr int NIP = 4;
  int pkh[NIP], pkw[NIP]; unsigned pci[NIP];
#pragma unroll
  for (int it = 0; it < NIP; ++it) {
    const unsigned k0 = (lane + 32u * it) << 3;
    const unsigned tap = k0 / Cu;
    pci[it] = k0 - tap * Cu; pkh[it] = (int)(tap / KW); pkw[it] = (int)(tap - (tap / KW) * KW);
  }
  for (unsigned m = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; m < Mu; m += warps) {
    const unsigned t = m / Wo, ow = m - t * Wo;
    const unsigned b = t / Ho, oh = t - b * Ho;
    const int ih0 = (int)(oh * g.sh) - g.pt, iw0 = (int)(ow * g.sw) - g.pl;
    const __nv_bfloat16* xb = x + (size_t)b * g.H * g.W * g.C;
    __nv_bfloat16* orow = out + (size_t)m * ldo;
#pragma unroll
    for (int it = 0; it < NIP; ++it) {
      const unsigned k8 = lane + 32u * it;
      if (k8 < k8n) {
        const unsigned k0 = k8 << 3;
        uint4 o = make_uint4(0u, 0u, 0u, 0u);
        const int ih = ih0 + pkh[it], iw = iw0 + pkw[it];
        if (k0 < (unsigned)Kc && (unsigned)ih < (unsigned)g.H && (unsigned)iw < (unsigned)g.W)
          o = ldg16(xb + ((size_t)ih * g.W + iw) * g.C + pci[it]);
        *reinterpret_cast<uint4*>(orow + k0) = o;
      }
    }
    for (unsigned k8 = lane + 32u * NIP; k8 < k8n; k8 += 32u) {
      const unsigned k0 = k8 << 3;
      uint4 o = make_uint4(0u, 0u, 0u, 0u);
      if (k0 < (unsigned)Kc) {
        const unsigned tap = k0 / Cu, ci = k0 - tap * Cu;
        const unsigned kh = tap / KW, kw = tap - kh * KW;
        const int ih = ih0 + (int)kh, iw = iw0 + (int)kw;
        if ((unsigned)ih < (unsigned)g.H && (unsigned)iw < (unsigned)g.W)
          o = ldg16(xb + ((size_t)ih * g.W + iw) * g.C + ci);
      }
      *reinterpret_cast<uint4*>(orow + k0) = o;
    }
  }
}

__global__ void dgrad_gather_b2b_kernel(const __nv_bfloat16* __restrict__ dy, __nv_bfloat16* __restrict__ out,
                                        const CG g, long Min, int Kd, int ldo) {
  TFDE_PDL_PROLOGUE();
  // one warp per input position m = (b, h, w); see im2col_b2b_kernel
  const unsigned k8n = (unsigned)ldo >> 3;
  const unsigned lane = threadIdx.x & 31u;
  const unsigned warps = (gridDim.x * blockDim.x) >> 5;
  const unsigned Mu = (unsigned)Min, Wd = (unsigned)g.W, Hd = (unsigned)g.H, Co = (unsigned)g.Cout, KW = (unsigned)g.KW;
  const bool unit = g.sh == 1 && g.sw == 1;        // stride 1 (every 3x3 of ResNet): no per-element divisions
  constexpr int NIP = 4;
  int pkh[NIP], pkw[NIP]; unsigned pco[NIP];
#pragma unroll
  for (int it = 0; it < NIP; ++it) {
    const unsigned k0 = (lane + 32u * it) << 3;
    const unsigned tap = k0 / Co;
    pco[it] = k0 - tap * Co; pkh[it] = (int)(tap / KW); pkw[it] = (int)(tap - (tap / KW) * KW);
  }
  for (unsigned m = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; m < Mu; m += warps) {
    const unsigned t = m / Wd, w = m - t * Wd;
    const unsigned b = t / Hd, h = t - b * Hd;
    const __nv_bfloat16* db = dy + (size_t)b * g.Ho * g.Wo * g.Cout;
    __nv_bfloat16* orow = out + (size_t)m * ldo;
    if (unit) {
#pragma unroll
      for (int it = 0; it < NIP; ++it) {
        const unsigned k8 = lane + 32u * it;
        if (k8 < k8n) {
          const unsigned k0 = k8 << 3;
          uint4 o = make_uint4(0u, 0u, 0u, 0u);
          const int oh = (int)h + g.pt - pkh[it], ow = (int)w + g.pl - pkw[it];
          if (k0 < (unsigned)Kd && (unsigned)oh < (unsigned)g.Ho && (unsigned)ow < (unsigned)g.Wo)
            o = ldg16(db + ((size_t)oh * g.Wo + ow) * g.Cout + pco[it]);
          *reinterpret_cast<uint4*>(orow + k0) = o;
        }
      }
    }
    for (unsigned k8 = unit ? lane + 32u * NIP : lane; k8 < k8n; k8 += 32u) {
      const unsigned k0 = k8 << 3;
      uint4 o = make_uint4(0u, 0u, 0u, 0u);
      if (k0 < (unsigned)Kd) {
        const unsigned tap = k0 / Co, co = k0 - tap * Co;
        const unsigned kh = tap / KW, kw = tap - kh * KW;
        const int th = (int)h + g.pt - (int)kh, tw = (int)w + g.pl - (int)kw;
        if (th >= 0 && tw >= 0) {
          const int oh = th / g.sh, ow = tw / g.sw;
          if (oh * g.sh == th && ow * g.sw == tw && oh < g.Ho && ow < g.Wo)
            o = ldg16(db + ((size_t)oh * g.Wo + ow) * g.Cout + co);
        }
      }
      *reinterpret_cast<uint4*>(orow + k0) = o;
    }
  }
}

// Data gradient of a strided 1x1 convolution: the GEMM runs on the compact [B*Ho*Wo, Cin] problem and this
// kernel spreads its rows over the input grid (positions the stride skipped get zero) -- instead of a
// [B*H*W, Cout] gathered operand that is 1 - 1/(sh*sw) zeros and a GEMM sh*sw times too large.
__global__ void strided_scatter_b2b_kernel(const __nv_bfloat16* __restrict__ g, __nv_bfloat16* __restrict__ dx,
                                           int Bn, int H, int W, int C, int Ho, int Wo, int sh, int sw) {
  TFDE_PDL_PROLOGUE();
  const int c8n = C >> 3;
  const long total = (long)Bn * H * W * c8n;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int c = (int)(i % c8n) * 8; long t = i / c8n;
    const int w = (int)(t % W); t /= W;
    const int h = (int)(t % H); const long b = t / H;
    uint4 o = make_uint4(0u, 0u, 0u, 0u);
    const int oh = h / sh, ow = w / sw;
    if (oh * sh == h && ow * sw == w && oh < Ho && ow < Wo) o = ldg16(g + ((b * Ho + oh) * Wo + ow) * (long)C + c);
    *reinterpret_cast<uint4*>(dx + (i << 3)) = o;
  }
}

inline int ew_blocks(long work) {
  long b = (work + 255) / 256;
  if (b > 148L * 16) b = 148L * 16;
  return (int)(b < 1 ? 1 : b);
}

// tile geometry of the statistics kernels
inline int stat_tpr(int C) { int t = C / 8; return t > 32 ? 32 : t; }
inline bool stat_ok(int C) {
  if (C < 8 || (C & 7)) return false;
  const int t = stat_tpr(C);
  return (ST % t) == 0 && (C % (8 * t)) == 0;
}
inline int stat_row_blocks(long M, int C, int blocks_per_sm = 2) {
  const int tpr = stat_tpr(C), rows = ST / tpr, ctiles = C / (8 * tpr);
  long want = (M + (long)rows * 8 - 1) / ((long)rows * 8);     // >= 8 rows per thread
  long cap = (148L * blocks_per_sm + ctiles - 1) / ctiles;     // exactly one resident wave
  if (want > cap) want = cap;
  if (want < 1) want = 1;
  return (int)want;
}

template <int MODE>
int launch_stats(StatArgs& a, float* ws, size_t ws_floats, unsigned* counters, cudaStream_t stream) {
  if (!stat_ok(a.C)) return TFDE_ERR_INVALID;
  a.tpr = stat_tpr(a.C);
  const int cw = 8 * a.tpr, ctiles = a.C / cw, rb = stat_row_blocks(a.M, a.C, MODE == 1 ? 1 : 2);
  if ((size_t)ctiles * rb * 2 * cw > ws_floats || ctiles > TFDE_MAX_COUNTERS) return TFDE_ERR_INVALID;
  a.partials = ws; a.counters = counters;
  const size_t smem = (size_t)2 * (ST / a.tpr) * cw * sizeof(float);      // = 32 KB
  tfde_launch(colstats_bf16_kernel<MODE>, dim3(ctiles, rb), dim3(ST), smem, stream, a);
  TFDE_RETURN_LAUNCH_STATUS();
}

}  // namespace

TFDE_API int tfde_act16_supported_channels(int C) { return stat_ok(C) ? 1 : 0; }

TFDE_API size_t tfde_act16_ws_floats(long M, int C) {
  if (!stat_ok(C)) return 0;
  const int tpr = stat_tpr(C), cw = 8 * tpr;
  return (size_t)(C / cw) * stat_row_blocks(M, C) * 2 * cw;
}

// Batch-norm, training forward: statistics (+ moving averages, + folded scale/shift) then y = x*scale+shift.
TFDE_API int tfde_bn_fwd_train_bf16(const void* x, void* y, long M, int C, const float* gamma, const float* beta,
                                    float* moving_mean, float* moving_var, float momentum, float eps, int relu,
                                    int unbiased_moving, float* save_mean, float* save_invstd, float* scale,
                                    float* shift, float* ws, size_t ws_floats, unsigned* counters,
                                    cudaStream_t stream) {
  StatArgs a{};
  a.x = (const __nv_bfloat16*)x; a.M = M; a.C = C;
  a.mean = save_mean; a.invstd = save_invstd; a.scale = scale; a.shift = shift;
  a.moving_mean = moving_mean; a.moving_var = moving_var; a.gamma = gamma; a.beta = beta;
  a.momentum = momentum; a.eps = eps; a.unbiased = unbiased_moving;
  const int rc = launch_stats<0>(a, ws, ws_floats, counters, stream);
  if (rc != 0) return rc;
  const long n8 = M * C / 8;
  tfde_launch(bn_apply_bf16_kernel, dim3(ew_blocks(n8)), dim3(256), 0, stream, (const __nv_bfloat16*)x, (__nv_bfloat16*)y, n8, C, scale,
                                                          shift, relu);
  TFDE_RETURN_LAUNCH_STATUS();
}

// y = relu?(x*scale + shift) with caller-provided per-channel scale/shift (inference with moving statistics).
// Batch-norm training forward when the producing GEMM already left per-32-row partial statistics
// (tfde_gemm_bf16_nn_stats_persistent): reduce the [G][C] partials, finalise, apply. x is read ONCE.
TFDE_API int tfde_bn_fwd_from_partials_bf16(const void* x, void* y, long M, int C, const float* psum, const float* psq,
                                            int G, const float* gamma, const float* beta, float* moving_mean,
                                            float* moving_var, float momentum, float eps, int relu,
                                            int unbiased_moving, float* save_mean, float* save_invstd, float* scale,
                                            float* shift, float* ws, size_t ws_floats, unsigned* counters,
                                            cudaStream_t stream) {
  if ((C & 63) || G <= 0 || !psum || !psq) return TFDE_ERR_INVALID;
  PartArgs a{};
  a.psum = psum; a.psq = psq; a.G = G; a.M = M; a.C = C; a.cw = (C % 128 == 0) ? 128 : 64;
  a.mean = save_mean; a.invstd = save_invstd; a.scale = scale; a.shift = shift;
  a.moving_mean = moving_mean; a.moving_var = moving_var; a.gamma = gamma; a.beta = beta;
  a.momentum = momentum; a.eps = eps; a.unbiased = unbiased_moving;
  const int ctiles = C / a.cw, rows = 256 / (a.cw / 4);
  int rb = (G + rows * 4 - 1) / (rows * 4);
  if (rb > 64) rb = 64;
  if (rb < 1) rb = 1;
  if ((size_t)ctiles * rb * 2 * a.cw > ws_floats || ctiles > TFDE_MAX_COUNTERS) return TFDE_ERR_INVALID;
  a.partials = ws; a.counters = counters;
  tfde_launch(bn_finalize_partials_kernel, dim3(ctiles, rb), dim3(256), 0, stream, a);
  const long n8 = M * C / 8;
  tfde_launch(bn_apply_bf16_kernel, dim3(ew_blocks(n8)), dim3(256), 0, stream, (const __nv_bfloat16*)x,
              (__nv_bfloat16*)y, n8, C, (const float*)scale, (const float*)shift, relu);
  TFDE_RETURN_LAUNCH_STATUS();
}

TFDE_API int tfde_bn_apply_bf16(const void* x, void* y, long M, int C, const float* scale, const float* shift,
                                int relu, cudaStream_t stream) {
  if (C & 7) return TFDE_ERR_INVALID;
  const long n8 = M * C / 8;
  tfde_launch(bn_apply_bf16_kernel, dim3(ew_blocks(n8)), dim3(256), 0, stream, (const __nv_bfloat16*)x, (__nv_bfloat16*)y, n8, C, scale,
                                                          shift, relu);
  TFDE_RETURN_LAUNCH_STATUS();
}

TFDE_API int tfde_bn_bwd_bf16(const void* x, const void* dy, void* dx, long M, int C, const float* gamma,
                              const float* save_mean, const float* save_invstd, const float* scale,
                              const float* shift, int relu, float* dgamma, float* dbeta, float* ws,
                              size_t ws_floats, unsigned* counters, cudaStream_t stream) {
  StatArgs a{};
  a.x = (const __nv_bfloat16*)x; a.dy = (const __nv_bfloat16*)dy; a.M = M; a.C = C;
  a.mean = const_cast<float*>(save_mean); a.invstd = const_cast<float*>(save_invstd);
  a.fscale = scale; a.fshift = shift; a.relu = relu; a.dbeta = dbeta; a.dgamma = dgamma;
  const int rc = launch_stats<1>(a, ws, ws_floats, counters, stream);
  if (rc != 0) return rc;
  const long n8 = M * C / 8;
  tfde_launch(bn_bwd_apply_bf16_kernel, dim3(ew_blocks(n8)), dim3(256), 0, stream, (const __nv_bfloat16*)x, (const __nv_bfloat16*)dy,
                                                              (__nv_bfloat16*)dx, n8, M, C, save_mean, save_invstd,
                                                              gamma, scale, shift, dbeta, dgamma, relu);
  TFDE_RETURN_LAUNCH_STATUS();
}

TFDE_API int tfde_colsum_bf16(const void* x, long M, int C, float* out, float* ws, size_t ws_floats,
                              unsigned* counters, cudaStream_t stream) {
  StatArgs a{};
  a.x = (const __nv_bfloat16*)x; a.M = M; a.C = C; a.colsum = out;
  return launch_stats<2>(a, ws, ws_floats, counters, stream);
}

TFDE_API int tfde_add_bf16(const void* a, const void* b, void* y, long n, int relu, cudaStream_t stream) {
  if (n & 7) return TFDE_ERR_INVALID;
  tfde_launch(add_bf16_kernel, dim3(ew_blocks(n / 8)), dim3(256), 0, stream, (const __nv_bfloat16*)a, (const __nv_bfloat16*)b,
                                                        (__nv_bfloat16*)y, n / 8, relu);
  TFDE_RETURN_LAUNCH_STATUS();
}

TFDE_API int tfde_relu_bwd_bf16(const void* y, const void* dy, const void* addend, void* dx, long n,
                                cudaStream_t stream) {
  if (n & 7) return TFDE_ERR_INVALID;
  tfde_launch(relu_bwd_bf16_kernel, dim3(ew_blocks(n / 8)), dim3(256), 0, stream, (const __nv_bfloat16*)y, (const __nv_bfloat16*)dy,
                                                             (const __nv_bfloat16*)addend, (__nv_bfloat16*)dx, n / 8);
  TFDE_RETURN_LAUNCH_STATUS();
}

TFDE_API int tfde_cast_f32_bf16(const float* x, void* y, long n, cudaStream_t stream) {
  tfde_launch(cast_f32_bf16_kernel, dim3(ew_blocks(n)), dim3(256), 0, stream, x, (__nv_bfloat16*)y, n);
  TFDE_RETURN_LAUNCH_STATUS();
}

TFDE_API int tfde_maxpool_fwd_bf16(const void* x, void* y, uint8_t* idx, int Bn, int H, int W, int C, int PH, int PW,
                                   int sh, int sw, int pt, int pl, int Ho, int Wo, cudaStream_t stream) {
  if (C & 7) return TFDE_ERR_INVALID;
  PG g{Bn, H, W, C, PH, PW, sh, sw, pt, pl, Ho, Wo};
  tfde_launch(maxpool_fwd_bf16_kernel, dim3(ew_blocks((long)Bn * Ho * Wo * (C / 8))), dim3(256), 0, stream, 
      (const __nv_bfloat16*)x, (__nv_bfloat16*)y, idx, g);
  TFDE_RETURN_LAUNCH_STATUS();
}
TFDE_API int tfde_maxpool_bwd_bf16(const void* dy, const uint8_t* idx, void* dx, int Bn, int H, int W, int C, int PH,
                                   int PW, int sh, int sw, int pt, int pl, int Ho, int Wo, cudaStream_t stream) {
  if (C & 7) return TFDE_ERR_INVALID;
  PG g{Bn, H, W, C, PH, PW, sh, sw, pt, pl, Ho, Wo};
  tfde_launch(maxpool_bwd_bf16_kernel, dim3(ew_blocks((long)Bn * H * W * (C / 8))), dim3(256), 0, stream, 
      (const __nv_bfloat16*)dy, idx, (__nv_bfloat16*)dx, g);
  TFDE_RETURN_LAUNCH_STATUS();
}

TFDE_API int tfde_gap_fwd_bf16(const void* x, float* y, int Bn, int HW, int C, cudaStream_t stream) {
  if (C & 7) return TFDE_ERR_INVALID;
  tfde_launch(gap_fwd_bf16_kernel, dim3(ew_blocks((long)Bn * (C / 8))), dim3(256), 0, stream, (const __nv_bfloat16*)x, y, Bn, HW, C);
  TFDE_RETURN_LAUNCH_STATUS();
}
TFDE_API int tfde_gap_bwd_bf16(const float* dy, void* dx, int Bn, int HW, int C, cudaStream_t stream) {
  if (C & 7) return TFDE_ERR_INVALID;
  tfde_launch(gap_bwd_bf16_kernel, dim3(ew_blocks((long)Bn * HW * (C / 8))), dim3(256), 0, stream, dy, (__nv_bfloat16*)dx, Bn, HW, C);
  TFDE_RETURN_LAUNCH_STATUS();
}

TFDE_API int tfde_strided_scatter_b2b(const void* g, void* dx, int Bn, int H, int W, int C, int Ho, int Wo, int sh,
                                      int sw, cudaStream_t stream) {
  if (C & 7) return TFDE_ERR_INVALID;
  tfde_launch(strided_scatter_b2b_kernel, dim3(ew_blocks((long)Bn * H * W * (C / 8))), dim3(256), 0, stream,
              (const __nv_bfloat16*)g, (__nv_bfloat16*)dx, Bn, H, W, C, Ho, Wo, sh, sw);
  TFDE_RETURN_LAUNCH_STATUS();
}

TFDE_API int tfde_im2col_b2b(const void* x, void* out, int B, int H, int W, int C, int KH, int KW, int Ho, int Wo,
                             int sh, int sw, int pt, int pl, int Cout, int ldo, cudaStream_t stream) {
  if ((ldo & 7) || (C & 7)) return TFDE_ERR_INVALID;
  const long M = (long)B * Ho * Wo;
  CG g{B, H, W, C, KH, KW, Ho, Wo, sh, sw, pt, pl, Cout};
  tfde_launch(im2col_b2b_kernel, dim3(ew_blocks(M * (ldo >> 3))), dim3(256), 0, stream, (const __nv_bfloat16*)x, (__nv_bfloat16*)out, g, M,
                                                                   KH * KW * C, ldo);
  TFDE_RETURN_LAUNCH_STATUS();
}
TFDE_API int tfde_dgrad_gather_b2b(const void* dy, void* out, int B, int H, int W, int C, int KH, int KW, int Ho,
                                   int Wo, int sh, int sw, int pt, int pl, int Cout, int ldo, cudaStream_t stream) {
  if ((ldo & 7) || (Cout & 7)) return TFDE_ERR_INVALID;
  const long Min = (long)B * H * W;
  CG g{B, H, W, C, KH, KW, Ho, Wo, sh, sw, pt, pl, Cout};
  tfde_launch(dgrad_gather_b2b_kernel, dim3(ew_blocks(Min * (ldo >> 3))), dim3(256), 0, stream, (const __nv_bfloat16*)dy,
                                                                           (__nv_bfloat16*)out, g, Min, KH * KW * Cout, ldo);
  TFDE_RETURN_LAUNCH_STATUS();
}
