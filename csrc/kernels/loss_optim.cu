// Loss (+gradient +accuracy) and single-GPU optimizer kernels, sm_100a.
//   G9  sparse categorical cross-entropy from logits (distributed_with_keras.py:41,
//       tf2_mnist_distributed.py:81-83) and from probabilities (mnist_keras_distributed.py:108,114);
//       one warp per row, produces loss sum, correct count and d(logits|probs) in one pass.
//   G11 SGD / momentum / Adam applied to the flat parameter arena (distributed_with_keras.py:42,
//       mnist_keras_distributed.py:111). With more than one replica this kernel is not used: the
//       update is the epilogue of the fused all-reduce in allreduce_p2p.cu.
// The loss scalars are written into the 4-float tail of the gradient arena
// ([loss*scale, correct, rows, 0]) so they ride the gradient all-reduce (G14) for free.
#include "common.cuh"
#include <cuda_bf16.h>

namespace {

constexpr int XW = 8;  // warps per block

// mode 0: logits, mode 1: probabilities
template <int MODE>
__global__ void __launch_bounds__(XW * 32)
xent_kernel(const float* __restrict__ in, const void* __restrict__ labels, int label_is_i64,
            float* __restrict__ dgrad, int rows, int C, float scale, float* __restrict__ partials,
            unsigned* __restrict__ counter, float* __restrict__ out4) {
  TFDE_PDL_PROLOGUE();
  __shared__ float s_l[XW], s_c[XW];
  __shared__ int s_last;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float loss = 0.f, correct = 0.f;
  for (int row = blockIdx.x * XW + warp; row < rows; row += gridDim.x * XW) {
    const float* r = in + (long)row * C;
    const int y = label_is_i64 ? (int)reinterpret_cast<const long long*>(labels)[row]
                               : reinterpret_cast<const int*>(labels)[row];
    if ((unsigned)y >= (unsigned)C) {
      // label outside [0, C): TensorFlow raises on the CPU and returns NaN on the GPU for it. Here the row poisons
      // the loss with NaN and contributes no gradient (and no out-of-bounds read).
      for (int c = lane; c < C; c += 32) dgrad[(long)row * C + c] = 0.f;
      if (lane == 0) loss += __int_as_float(0x7fc00000);
      continue;
    }
    // arg-max (first maximum) and max value
    float mx = -INFINITY; int am = 0x7fffffff;
    for (int c = lane; c < C; c += 32) {
      const float v = __ldg(r + c);
      if (v > mx) { mx = v; am = c; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, mx, o);
      const int oa = __shfl_xor_sync(0xffffffffu, am, o);
      if (ov > mx || (ov == mx && oa < am)) { mx = ov; am = oa; }
    }
    if (MODE == 0) {
      float s = 0.f;
      for (int c = lane; c < C; c += 32) s += expf(__ldg(r + c) - mx);
      s = warp_sum(s);
      const float lse = mx + logf(s);
      const float inv = 1.f / s;
      for (int c = lane; c < C; c += 32) {
        const float p = expf(__ldg(r + c) - mx) * inv;
        dgrad[(long)row * C + c] = (p - (c == y ? 1.f : 0.f)) * scale;
      }
      if (lane == 0) loss += lse - __ldg(r + y);
    } else {
      const float py = __ldg(r + y);
      const float cl = fminf(fmaxf(py, 1e-7f), 1.f - 1e-7f);
      const bool inside = (py >= 1e-7f) && (py <= 1.f - 1e-7f);
      for (int c = lane; c < C; c += 32)
        dgrad[(long)row * C + c] = (c == y && inside) ? (-scale / cl) : 0.f;
      if (lane == 0) loss += -logf(cl);
    }
    if (lane == 0 && am == y) correct += 1.f;
  }
  if (lane == 0) { s_l[warp] = loss; s_c[warp] = correct; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float l = 0.f, c = 0.f;
#pragma unroll
    for (int i = 0; i < XW; ++i) { l += s_l[i]; c += s_c[i]; }
    __stcg(partials + 2 * blockIdx.x, l);
    __stcg(partials + 2 * blockIdx.x + 1, c);
    __threadfence();
    s_last = (atomicAdd(counter, 1u) == gridDim.x - 1);
  }
  __syncthreads();
  if (!s_last) return;
  if (threadIdx.x == 0) {
    __threadfence();
    float l = 0.f, c = 0.f;
    for (unsigned b = 0; b < gridDim.x; ++b) {
      l += __ldcg(partials + 2 * b); c += __ldcg(partials + 2 * b + 1);
    }
    out4[0] = l * scale; out4[1] = c; out4[2] = (float)rows; out4[3] = 0.f;
    *counter = 0u;
  }
}

// ------------------------------------------------------------------------------------------
// Optimizer on the flat arena. opt: 0 SGD, 1 momentum, 2 nesterov momentum, 3 Adam.
// `n` parameters followed by `n_extra` reduced-only scalars (metrics tail) that are accumulated
// into `metrics` (running sums read by fit() once per epoch / log interval).
// ------------------------------------------------------------------------------------------
struct OptArgs {
  float* w; const float* g; float* s1; float* s2; __nv_bfloat16* w_bf16;
  long n; int n_extra; float* metrics; float* last_extra; int opt;
};

__device__ __forceinline__ float apply_one(int opt, float w, float g, float* s1, float* s2, long i,
                                           float lr, float mu, float b1, float b2, float eps,
                                           float lr_t) {
  if (opt == 0) return w - lr * g;
  if (opt == 1 || opt == 2) {
    const float v = mu * s1[i] - lr * g;
    s1[i] = v;
    return w + (opt == 2 ? (mu * v - lr * g) : v);
  }
  const float m = b1 * s1[i] + (1.f - b1) * g;
  const float v = b2 * s2[i] + (1.f - b2) * g * g;
  s1[i] = m; s2[i] = v;
  return w - lr_t * m / (sqrtf(v) + eps);
}

__device__ __forceinline__ void hparams_advance(TfdeHParams* hp) {
  hp->beta1_pow *= hp->beta1;
  hp->beta2_pow *= hp->beta2;
  hp->step += 1u;
}

// The last block to finish advances the device-side step / Adam powers: by then every block has
// consumed the old values, so no second launch and no cooperative grid sync is needed.
__global__ void optimizer_kernel(OptArgs a, TfdeHParams* hp, unsigned* __restrict__ counter) {
  TFDE_PDL_PROLOGUE();
  const float lr = hp->lr, mu = hp->momentum, b1 = hp->beta1, b2 = hp->beta2, eps = hp->eps;
  const float gs = hp->grad_scale;
  // Adam bias correction for step t = hp->step + 1, powers maintained on device.
  const float b1p = hp->beta1_pow * b1, b2p = hp->beta2_pow * b2;
  const float lr_t = (a.opt == 3) ? lr * sqrtf(1.f - b2p) / (1.f - b1p) : lr;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < a.n;
       i += (long)gridDim.x * blockDim.x) {
    const float nw = apply_one(a.opt, a.w[i], __ldg(a.g + i) * gs, a.s1, a.s2, i, lr, mu, b1, b2, eps, lr_t);
    a.w[i] = nw;
    if (a.w_bf16) a.w_bf16[i] = __float2bfloat16(nw);
  }
  if (blockIdx.x == 0 && threadIdx.x < a.n_extra) {
    const float v = __ldg(a.g + a.n + threadIdx.x);
    if (a.metrics) a.metrics[threadIdx.x] += v;
    if (a.last_extra) a.last_extra[threadIdx.x] = v;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    if (atomicAdd(counter, 1u) == gridDim.x - 1) {
      hparams_advance(hp);
      *counter = 0u;
    }
  }
}

__global__ void hparams_advance_kernel(TfdeHParams* hp) { hparams_advance(hp); }

}  // namespace

TFDE_API size_t tfde_xent_ws_floats() { return (size_t)148 * 2 * 2; }

TFDE_API int tfde_xent(const float* in, const void* labels, int label_is_i64, float* dgrad, int rows,
                       int C, float scale, int from_probs, float* partials, unsigned* counter,
                       float* out4, cudaStream_t stream) {
  int grid = (rows + XW - 1) / XW;
  if (grid > 148 * 2) grid = 148 * 2;
  if (grid < 1) grid = 1;
  if (from_probs)
    tfde_launch(xent_kernel<1>, dim3(grid), dim3(XW * 32), 0, stream, in, labels, label_is_i64, dgrad, rows, C, scale, partials, counter, out4);
  else
    tfde_launch(xent_kernel<0>, dim3(grid), dim3(XW * 32), 0, stream, in, labels, label_is_i64, dgrad, rows, C, scale, partials, counter, out4);
  TFDE_RETURN_LAUNCH_STATUS();
}

TFDE_API int tfde_optimizer_apply(float* w, const float* g, float* s1, float* s2, void* w_bf16,
                                  long n, int n_extra, float* metrics, float* last_extra, int opt,
                                  void* hparams, unsigned* counter, cudaStream_t stream) {
  if (opt < 0 || opt > 3 || n_extra > 32) return TFDE_ERR_INVALID;
  if ((opt >= 1 && !s1) || (opt == 3 && !s2)) return TFDE_ERR_INVALID;
  OptArgs a{w, g, s1, s2, (__nv_bfloat16*)w_bf16, n, n_extra, metrics, last_extra, opt};
  long grid = (n + 255) / 256;
  if (grid > 148L * 8) grid = 148L * 8;
  if (grid < 1) grid = 1;
  tfde_launch(optimizer_kernel, dim3((unsigned)grid), dim3(256), 0, stream, a, (TfdeHParams*)hparams, counter);
  TFDE_RETURN_LAUNCH_STATUS();
}

// Advance only (used by paths where another kernel applied the update).
TFDE_API int tfde_hparams_advance(void* hparams, cudaStream_t stream) {
  hparams_advance_kernel<<<1, 1, 0, stream>>>((TfdeHParams*)hparams);
  TFDE_RETURN_LAUNCH_STATUS();
}
