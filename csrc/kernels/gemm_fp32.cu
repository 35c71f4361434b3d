// fp32 CUDA-core GEMM family with gather loaders (dense, transposed, im2col, im2col^T, dgrad
// gather), deterministic split-K and fused epilogues. This is the exact-fp32 path: the reference
// trains in fp32 (distributed_with_keras.py:21, mnist_keras_distributed.py:159) and its layers have
// N = 6..64 output channels and K = 9..5408, i.e. shapes that are bandwidth/latency bound, not
// tensor-core bound. Large bf16 GEMMs go to gemm_tcgen05.cu instead.
//
// One kernel template covers:
//   G1 conv fwd      C[(b,oh,ow), co] = im2col(x) . W[(kh,kw,ci), co]  (+bias, ReLU)
//   G2 conv wgrad    dW[(kh,kw,ci), co] = im2col(x)^T . dy   (+ bias-grad as an extra all-ones row)
//   G3 conv dgrad    dx[(b,h,w), ci] = gather(dy) . Wrot[(kh,kw,co), ci]
//   G4 dense fwd     y = x . W (+bias)(+ReLU)
//   G5 dense wgrad   dW = x^T . dy (+db as ones row), dense dgrad dx = dy . W^T
// ReLU backward never runs as its own kernel: the loaders multiply dy by (y > 0) on the fly (G8).
#include "common.cuh"
#include <cmath>

namespace {

constexpr int BK = 16;
constexpr int NTHREADS = 256;

enum AMode { A_DENSE = 0, A_DENSE_T = 1, A_IM2COL = 2, A_IM2COL_T = 3, A_DGRAD = 4 };
enum BMode { B_DENSE = 0, B_DENSE_T = 1, B_WROT = 2 };

struct ConvGeom {
  int H, W, C;      // input spatial size and channels
  int KH, KW;       // kernel
  int Ho, Wo;       // output spatial size
  int sh, sw;       // strides
  int pt, pl;       // top / left padding
  int Cout;
};

// Optional "push" epilogue: the filter-gradient tile is ALSO stored straight into the peers' all-reduce
// receive slots over NVLink (GEMM -> reduce-scatter in one kernel): the transfer of the 1.38 MB dense
// gradient overlaps the math tile by tile and the all-reduce kernel that follows skips its push phase
// for this range. Layout contract = allreduce_p2p.cu (slot[rank] of parity (epoch+1)&1 in every peer).
struct TfdePush {
  int enabled, world, rank, mode;      // mode 0 one-shot (every peer gets everything), 1 two-shot (owner only)
  long slot_floats;                    // floats per receive slot
  long arena_off;                      // index of C[0] inside the flat gradient arena
  float* recv[8];                      // peer-mapped receive buffers [2][world][slot_floats]
  const uint32_t* epoch;               // all-reduce epoch counter of block 0 (device memory)
  uint32_t* done_flag;                 // set to the epoch this push belongs to
};

struct GemmParams {
  TfdePush push;
  const float* A; const float* Amask; int lda;
  const float* B; const float* Bmask; int ldb;
  float* C; int ldc;
  float* C2;            // destination of the all-ones row (bias gradient), or nullptr
  const float* bias;    // added per output column, or nullptr
  int M, N, K;
  int ones_row;         // A_op row that is identically 1 (== M-1 when used), else -1
  int relu;
  int splits;
  float* ws;            // split-K partials [splits][M][N]
  unsigned* counters;   // one ticket per output tile, zero on entry, left zero on exit
  ConvGeom g;
};

// ------------------------------------------------------------------------------------------
// A-operand loaders. row() precomputes whatever depends only on the A_op row index.
// ------------------------------------------------------------------------------------------
template <int MODE> struct ALoader;

template <> struct ALoader<A_DENSE> {  // A_op[m,k] = A[m*lda + k]
  static constexpr bool kFast = true;
  struct Row { long off; bool ok; };
  static __device__ __forceinline__ Row row(const GemmParams& p, int m) {
    return Row{(long)m * p.lda, m < p.M};
  }
  static __device__ __forceinline__ float load(const GemmParams& p, const Row& r, int m, int k) {
    if (!r.ok || k >= p.K) return 0.f;
    float v = __ldg(p.A + r.off + k);
    if (p.Amask != nullptr && !(__ldg(p.Amask + r.off + k) > 0.f)) v = 0.f;
    return v;
  }
};

template <> struct ALoader<A_DENSE_T> {  // A_op[m,k] = A[k*lda + m]; optional ones row
  static constexpr bool kFast = false;
  struct Row { int m; int kind; };  // kind: 0 = out of range, 1 = data, 2 = ones
  static __device__ __forceinline__ Row row(const GemmParams& p, int m) {
    return Row{m, m >= p.M ? 0 : (m == p.ones_row ? 2 : 1)};
  }
  static __device__ __forceinline__ float load(const GemmParams& p, const Row& r, int m, int k) {
    if (r.kind == 0 || k >= p.K) return 0.f;
    if (r.kind == 2) return 1.f;
    return __ldg(p.A + (long)k * p.lda + r.m);
  }
};

template <> struct ALoader<A_IM2COL> {  // A_op[(b,oh,ow), (kh,kw,ci)] = x[b, oh*sh-pt+kh, ow*sw-pl+kw, ci]
  static constexpr bool kFast = true;
  struct Row { long base; int ih0, iw0; bool ok; };
  static __device__ __forceinline__ Row row(const GemmParams& p, int m) {
    Row r; r.ok = m < p.M;
    const int mm = r.ok ? m : 0;
    const int ow = mm % p.g.Wo; const int t = mm / p.g.Wo;
    const int oh = t % p.g.Ho; const int b = t / p.g.Ho;
    r.ih0 = oh * p.g.sh - p.g.pt; r.iw0 = ow * p.g.sw - p.g.pl;
    r.base = (long)b * p.g.H * p.g.W * p.g.C;
    return r;
  }
  static __device__ __forceinline__ float load(const GemmParams& p, const Row& r, int m, int k) {
    if (!r.ok || k >= p.K) return 0.f;
    const int ci = k % p.g.C; const int t = k / p.g.C;
    const int kw = t % p.g.KW; const int kh = t / p.g.KW;
    const int ih = r.ih0 + kh, iw = r.iw0 + kw;
    if ((unsigned)ih >= (unsigned)p.g.H || (unsigned)iw >= (unsigned)p.g.W) return 0.f;
    return __ldg(p.A + r.base + ((long)ih * p.g.W + iw) * p.g.C + ci);
  }
};

template <> struct ALoader<A_IM2COL_T> {  // A_op[(kh,kw,ci), (b,oh,ow)] ; optional ones row
  static constexpr bool kFast = false;
  struct Row { int kh, kw, ci; int kind; };
  static __device__ __forceinline__ Row row(const GemmParams& p, int m) {
    Row r; r.kind = m >= p.M ? 0 : (m == p.ones_row ? 2 : 1);
    const int mm = r.kind == 1 ? m : 0;
    r.ci = mm % p.g.C; const int t = mm / p.g.C;
    r.kw = t % p.g.KW; r.kh = t / p.g.KW;
    return r;
  }
  static __device__ __forceinline__ float load(const GemmParams& p, const Row& r, int m, int k) {
    if (r.kind == 0 || k >= p.K) return 0.f;
    if (r.kind == 2) return 1.f;
    const int ow = k % p.g.Wo; const int t = k / p.g.Wo;
    const int oh = t % p.g.Ho; const int b = t / p.g.Ho;
    const int ih = oh * p.g.sh - p.g.pt + r.kh, iw = ow * p.g.sw - p.g.pl + r.kw;
    if ((unsigned)ih >= (unsigned)p.g.H || (unsigned)iw >= (unsigned)p.g.W) return 0.f;
    return __ldg(p.A + (((long)b * p.g.H + ih) * p.g.W + iw) * p.g.C + r.ci);
  }
};

template <> struct ALoader<A_DGRAD> {  // A_op[(b,h,w), (kh,kw,co)] = dy[b,(h+pt-kh)/sh,(w+pl-kw)/sw,co]
  static constexpr bool kFast = true;
  struct Row { long base; int h, w; bool ok; };
  static __device__ __forceinline__ Row row(const GemmParams& p, int m) {
    Row r; r.ok = m < p.M;
    const int mm = r.ok ? m : 0;
    r.w = mm % p.g.W; const int t = mm / p.g.W;
    r.h = t % p.g.H; const int b = t / p.g.H;
    r.base = (long)b * p.g.Ho * p.g.Wo * p.g.Cout;
    return r;
  }
  static __device__ __forceinline__ float load(const GemmParams& p, const Row& r, int m, int k) {
    if (!r.ok || k >= p.K) return 0.f;
    const int co = k % p.g.Cout; const int t = k / p.g.Cout;
    const int kw = t % p.g.KW; const int kh = t / p.g.KW;
    const int th = r.h + p.g.pt - kh, tw = r.w + p.g.pl - kw;
    if (th < 0 || tw < 0) return 0.f;
    const int oh = th / p.g.sh, ow = tw / p.g.sw;
    if (oh * p.g.sh != th || ow * p.g.sw != tw || oh >= p.g.Ho || ow >= p.g.Wo) return 0.f;
    const long off = r.base + ((long)oh * p.g.Wo + ow) * p.g.Cout + co;
    float v = __ldg(p.A + off);
    if (p.Amask != nullptr && !(__ldg(p.Amask + off) > 0.f)) v = 0.f;
    return v;
  }
};

// ------------------------------------------------------------------------------------------
// B-operand loaders
// ------------------------------------------------------------------------------------------
template <int MODE> struct BLoader;
template <> struct BLoader<B_DENSE> {  // B_op[k,n] = B[k*ldb + n]
  static constexpr bool kFast = false;
  static __device__ __forceinline__ float load(const GemmParams& p, int k, int n) {
    if (k >= p.K || n >= p.N) return 0.f;
    const long off = (long)k * p.ldb + n;
    float v = __ldg(p.B + off);
    if (p.Bmask != nullptr && !(__ldg(p.Bmask + off) > 0.f)) v = 0.f;
    return v;
  }
};
template <> struct BLoader<B_DENSE_T> {  // B_op[k,n] = B[n*ldb + k]
  static constexpr bool kFast = true;
  static __device__ __forceinline__ float load(const GemmParams& p, int k, int n) {
    if (k >= p.K || n >= p.N) return 0.f;
    return __ldg(p.B + (long)n * p.ldb + k);
  }
};
template <> struct BLoader<B_WROT> {  // B_op[(kh,kw,co), ci] = W[kh,kw,ci,co]
  static constexpr bool kFast = true;
  static __device__ __forceinline__ float load(const GemmParams& p, int k, int n) {
    if (k >= p.K || n >= p.N) return 0.f;
    const int co = k % p.g.Cout; const int t = k / p.g.Cout;  // t = kh*KW + kw
    return __ldg(p.B + ((long)t * p.g.C + n) * p.g.Cout + co);
  }
};

// ------------------------------------------------------------------------------------------
// The kernel. grid = (tiles_n, tiles_m, splits); 256 threads; each thread owns TM x TN outputs.
// ------------------------------------------------------------------------------------------
template <int AM, int BM_, int BMODE, int BN_, int TM, int TN, int BK_ = 16>
__global__ void __launch_bounds__(NTHREADS) gemm_fp32_kernel(const GemmParams p) {
  static_assert((BM_ / TM) * (BN_ / TN) == NTHREADS, "tile/thread mismatch");
  TFDE_PDL_PROLOGUE();
  // k-tile depth of this instantiation (shadows the file-level default). BK_ = 64 is used when the whole
  // reduction fits one tile (K <= 64: the batch-64 wgrad/dgrad GEMMs) so every load is in flight at once.
  constexpr int BK = BK_;
  constexpr int APAD = 4;
  __shared__ __align__(16) float As[BK][BM_ + APAD];
  __shared__ __align__(16) float Bs[BK][BN_ + APAD];
  __shared__ int s_last;

  using AL = ALoader<AM>;
  using BL = BLoader<BMODE>;
  const int t = threadIdx.x;
  const int m0 = blockIdx.y * BM_, n0 = blockIdx.x * BN_;
  const int tx = t % (BN_ / TN), ty = t / (BN_ / TN);

  // K range of this split (whole BK tiles)
  const int ktiles = (p.K + BK - 1) / BK;
  const int per = (ktiles + p.splits - 1) / p.splits;
  const int kt0 = blockIdx.z * per;
  const int kt1 = min(ktiles, kt0 + per);

  constexpr int A_ELEMS = BM_ * BK / NTHREADS;
  constexpr int B_ELEMS = BN_ * BK / NTHREADS;
  static_assert(A_ELEMS >= 1 && B_ELEMS >= 1, "tile too small");

  // Per-thread A rows are the same for every k tile: precompute them.
  constexpr int A_ROWS = AL::kFast ? A_ELEMS : 1;
  typename AL::Row arow[A_ROWS];
  if (AL::kFast) {
#pragma unroll
    for (int i = 0; i < A_ROWS; ++i) arow[i] = AL::row(p, m0 + (t + i * NTHREADS) / BK);
  } else {
    arow[0] = AL::row(p, m0 + t % BM_);
  }

  float acc[TM][TN];
#pragma unroll
  for (int i = 0; i < TM; ++i)
#pragma unroll
    for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;

  // Software pipeline: the global loads of tile kt+1 are in flight while tile kt is multiplied out of
  // shared memory (these GEMMs have only a handful of k tiles per CTA, so exposed load latency -- not
  // FLOPs -- is what the step pays for).
  float areg[A_ELEMS], breg[B_ELEMS];
  auto load_tile = [&](int kt) {
    const int k0 = kt * BK;
#pragma unroll
    for (int i = 0; i < A_ELEMS; ++i) {
      const int e = t + i * NTHREADS;
      if (AL::kFast) areg[i] = AL::load(p, arow[i], m0 + e / BK, k0 + e % BK);
      else           areg[i] = AL::load(p, arow[0], m0 + e % BM_, k0 + e / BM_);
    }
#pragma unroll
    for (int i = 0; i < B_ELEMS; ++i) {
      const int e = t + i * NTHREADS;
      if (BL::kFast) breg[i] = BL::load(p, k0 + e % BK, n0 + e / BK);
      else           breg[i] = BL::load(p, k0 + e / BN_, n0 + e % BN_);
    }
  };
  if (kt0 < kt1) load_tile(kt0);
  for (int kt = kt0; kt < kt1; ++kt) {
    __syncthreads();  // previous tile fully consumed
#pragma unroll
    for (int i = 0; i < A_ELEMS; ++i) {
      const int e = t + i * NTHREADS;
      if (AL::kFast) As[e % BK][e / BK] = areg[i];
      else           As[e / BM_][e % BM_] = areg[i];
    }
#pragma unroll
    for (int i = 0; i < B_ELEMS; ++i) {
      const int e = t + i * NTHREADS;
      if (BL::kFast) Bs[e % BK][e / BK] = breg[i];
      else           Bs[e / BN_][e % BN_] = breg[i];
    }
    __syncthreads();
    if (kt + 1 < kt1) load_tile(kt + 1);
#pragma unroll
    for (int k = 0; k < BK; ++k) {
      float a[TM], b[TN];
#pragma unroll
      for (int i = 0; i < TM; ++i) a[i] = As[k][ty * TM + i];
#pragma unroll
      for (int j = 0; j < TN; ++j) b[j] = Bs[k][tx * TN + j];
#pragma unroll
      for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
  }

  if (p.splits > 1) {
    // Deterministic split-K: every split stores its partial, the last block to finish a tile sums
    // the partials in split order (independent of arrival order) and runs the epilogue. Partials are
    // stored with a leading dimension padded to 4 so both sides use 16-byte accesses, and the
    // reduction keeps 4 splits x TM float4 loads in flight per thread (L2 latency, not bandwidth,
    // bounds it).
    static_assert(TN == 4, "split-K partial layout assumes 4 columns per thread");
    const int ldw = (p.N + 3) & ~3;
    const size_t zstride = (size_t)p.M * ldw;
    float* mine = p.ws + (size_t)blockIdx.z * zstride;
    const int ncol = n0 + tx * TN;
#pragma unroll
    for (int i = 0; i < TM; ++i) {
      const int m = m0 + ty * TM + i;
      if (m < p.M && ncol < ldw)
        __stcg(reinterpret_cast<float4*>(mine + (size_t)m * ldw + ncol),
               make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]));
    }
    __threadfence();
    __syncthreads();
    const unsigned tile = blockIdx.y * gridDim.x + blockIdx.x;
    if (t == 0) {
      const unsigned ticket = atomicAdd(p.counters + tile, 1u);
      s_last = (ticket == (unsigned)p.splits - 1u);
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
#pragma unroll
    for (int i = 0; i < TM; ++i)
#pragma unroll
      for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;
    if (ncol < ldw) {
      constexpr int ZU = 4;
      for (int z0 = 0; z0 < p.splits; z0 += ZU) {
        float4 v[ZU][TM];
#pragma unroll
        for (int u = 0; u < ZU; ++u) {
          const float* part = p.ws + (size_t)(z0 + u) * zstride;
#pragma unroll
          for (int i = 0; i < TM; ++i) {
            const int m = m0 + ty * TM + i;
            v[u][i] = (z0 + u < p.splits && m < p.M)
                          ? __ldcg(reinterpret_cast<const float4*>(part + (size_t)m * ldw + ncol))
                          : make_float4(0.f, 0.f, 0.f, 0.f);
          }
        }
#pragma unroll
        for (int u = 0; u < ZU; ++u)   // fixed z order => bit-reproducible
#pragma unroll
          for (int i = 0; i < TM; ++i) {
            acc[i][0] += v[u][i].x; acc[i][1] += v[u][i].y; acc[i][2] += v[u][i].z; acc[i][3] += v[u][i].w;
          }
      }
    }
    if (t == 0) p.counters[tile] = 0u;  // self-cleaning for the next launch
  }

  // Epilogue: bias, ReLU, ones-row routing (+ optional push of the tile into the peers' receive slots).
  int parity = 0;
  if (p.push.enabled) {
    const uint32_t e = __ldg(p.push.epoch) + 1u;
    parity = (int)(e & 1u);
    if (t == 0 && blockIdx.x == 0 && blockIdx.y == 0) *p.push.done_flag = e;
  }
#pragma unroll
  for (int i = 0; i < TM; ++i) {
    const int m = m0 + ty * TM + i;
    if (m >= p.M) continue;
#pragma unroll
    for (int j = 0; j < TN; ++j) {
      const int n = n0 + tx * TN + j;
      if (n >= p.N) continue;
      float v = acc[i][j];
      if (m == p.ones_row) {
        if (p.C2 != nullptr) p.C2[n] = v;
        continue;
      }
      if (p.bias != nullptr) v += __ldg(p.bias + n);
      if (p.relu) v = fmaxf(v, 0.f);
      p.C[(size_t)m * p.ldc + n] = v;
    }
    if (p.push.enabled && m != p.ones_row) {
      // push this row's 4 finished outputs (no bias / activation on a filter gradient) as one 16-byte
      // NVLink store per destination; ranges are 16-byte aligned by construction (arena offsets % 4 == 0)
      const int nq = n0 + tx * TN;
      if (nq < p.N) {
        const long gi = p.push.arena_off + (long)m * p.ldc + nq;      // index in the gradient arena
        const size_t slot_base = ((size_t)parity * p.push.world + p.push.rank) * p.push.slot_floats;
        const bool vec = (TN == 4) && (nq + 3 < p.N) && ((gi & 3) == 0);
        const float4 v4 = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
        if (p.push.mode == 0) {
          for (int r = 0; r < p.push.world; ++r) {
            if (r == p.push.rank) continue;
            float* dst = p.push.recv[r] + slot_base + gi;
            if (vec) __stcg(reinterpret_cast<float4*>(dst), v4);
            else for (int j = 0; j < TN && nq + j < p.N; ++j) __stcg(dst + j, acc[i][j]);
          }
        } else {
          // a 4-group never straddles two owners: slot_floats % 4 == 0 and gi % 4 == 0 on the vector path
          for (int j0 = 0; j0 < TN; j0 += (vec ? TN : 1)) {
            if (nq + j0 >= p.N) break;
            const long g = gi + j0;
            const int owner = (int)(g / p.push.slot_floats);
            if (owner == p.push.rank) continue;
            float* dst = p.push.recv[owner] + slot_base + (g - (long)owner * p.push.slot_floats);
            if (vec) __stcg(reinterpret_cast<float4*>(dst), v4);
            else __stcg(dst, acc[i][j0]);
          }
        }
      }
    }
  }
}

template <int AM, int BMODE>
int launch_cfg(const GemmParams& p, cudaStream_t stream) {
  // Tile shape by N: the reference's layers have N in {6,10,12,24,32,64,200}.
  if (p.N <= 16) {
    dim3 grid((p.N + 15) / 16, (p.M + 255) / 256, p.splits);
    tfde_launch(gemm_fp32_kernel<AM, 256, BMODE, 16, 4, 4>, grid, dim3(NTHREADS), 0, stream, p);
  } else if (p.N <= 32) {
    dim3 grid((p.N + 31) / 32, (p.M + 127) / 128, p.splits);
    tfde_launch(gemm_fp32_kernel<AM, 128, BMODE, 32, 4, 4>, grid, dim3(NTHREADS), 0, stream, p);
  } else {
    dim3 grid((p.N + 63) / 64, (p.M + 63) / 64, p.splits);
    if (p.K <= 64 && p.splits == 1 && (AM == A_DENSE || AM == A_DENSE_T))
      tfde_launch(gemm_fp32_kernel<AM, 64, BMODE, 64, 4, 4, 64>, grid, dim3(NTHREADS), 0, stream, p);   // one deep k tile
    else
      tfde_launch(gemm_fp32_kernel<AM, 64, BMODE, 64, 4, 4>, grid, dim3(NTHREADS), 0, stream, p);
  }
  TFDE_RETURN_LAUNCH_STATUS();
}

int tile_m_for(int N) { return N <= 16 ? 256 : (N <= 32 ? 128 : 64); }
int tile_n_for(int N) { return N <= 16 ? 16 : (N <= 32 ? 32 : 64); }

// Pick the split-K factor: fill ~2 CTAs per SM, keep >= 4 k-tiles per split, respect workspace.
int choose_splits(int M, int N, int K, size_t ws_bytes, int max_tiles_counters) {
  const int tiles = ((M + tile_m_for(N) - 1) / tile_m_for(N)) * ((N + tile_n_for(N) - 1) / tile_n_for(N));
  if (tiles > max_tiles_counters) return 1;
  if (K <= 64) return 1;   // one deep k tile covers the whole reduction: nothing to split
  const int ktiles = (K + BK - 1) / BK;
  // exposed-latency model: main loop ~ (ktiles/s) dependent tile loads, reduction ~ s/4 dependent L2
  // round trips  =>  s ~ 2*sqrt(ktiles), capped by the SM count and by >= 2 k-tiles per split.
  int s = (int)(2.0 * sqrt((double)ktiles) + 0.5);
  s = min(s, min(ktiles / 2, (2 * 148) / max(tiles, 1)));
  s = max(1, min(s, 128));
  const size_t per = (size_t)M * ((N + 3) & ~3) * sizeof(float);
  if (per == 0) return 1;
  const size_t fit = ws_bytes / per;
  if ((size_t)s > fit) s = (int)max((size_t)1, fit);
  return s;
}

int dispatch(int amode, int bmode, const GemmParams& p, cudaStream_t s) {
  if (p.M <= 0 || p.N <= 0) return 0;
  if (amode == A_DENSE && bmode == B_DENSE) return launch_cfg<A_DENSE, B_DENSE>(p, s);
  if (amode == A_DENSE_T && bmode == B_DENSE) return launch_cfg<A_DENSE_T, B_DENSE>(p, s);
  if (amode == A_DENSE && bmode == B_DENSE_T) return launch_cfg<A_DENSE, B_DENSE_T>(p, s);
  if (amode == A_IM2COL && bmode == B_DENSE) return launch_cfg<A_IM2COL, B_DENSE>(p, s);
  if (amode == A_IM2COL_T && bmode == B_DENSE) return launch_cfg<A_IM2COL_T, B_DENSE>(p, s);
  if (amode == A_DGRAD && bmode == B_WROT) return launch_cfg<A_DGRAD, B_WROT>(p, s);
  return TFDE_ERR_INVALID;
}

constexpr int kMaxCounters = TFDE_MAX_COUNTERS;  // tickets available in the caller-provided counter buffer

}  // namespace

// ---------------------------------------------------------------------------------------------
// C ABI (declarations with reference citations: include/tfde_kernels.h)
// ---------------------------------------------------------------------------------------------

// y[M,N] = act(x[M,K] . w[K,N] + bias)
TFDE_API int tfde_dense_fwd(const float* x, const float* w, const float* bias, float* y, int M,
                            int K, int N, int relu, float* ws, size_t ws_bytes,
                            unsigned* counters, cudaStream_t stream) {
  GemmParams p{};
  p.A = x; p.lda = K; p.B = w; p.ldb = N; p.C = y; p.ldc = N; p.bias = bias;
  p.M = M; p.N = N; p.K = K; p.ones_row = -1; p.relu = relu;
  p.ws = ws; p.counters = counters;
  p.splits = (ws && counters) ? choose_splits(M, N, K, ws_bytes, kMaxCounters) : 1;
  return dispatch(A_DENSE, B_DENSE, p, stream);
}

// dw[K,N] = x[M,K]^T . (dy[M,N] * (ymask > 0));  db[N] = column sums (optional)
TFDE_API int tfde_dense_wgrad(const float* x, const float* dy, const float* ymask, float* dw,
                              float* db, int M, int K, int N, float* ws, size_t ws_bytes,
                              unsigned* counters, cudaStream_t stream) {
  GemmParams p{};
  p.A = x; p.lda = K;               // A_op[k_in, m] = x[m*K + k_in]
  p.B = dy; p.Bmask = ymask; p.ldb = N;
  p.C = dw; p.ldc = N; p.C2 = db;
  p.M = db ? K + 1 : K; p.N = N; p.K = M;
  p.ones_row = db ? K : -1;
  p.ws = ws; p.counters = counters;
  p.splits = (ws && counters) ? choose_splits(p.M, N, M, ws_bytes, kMaxCounters) : 1;
  return dispatch(A_DENSE_T, B_DENSE, p, stream);
}

// tfde_dense_wgrad whose epilogue additionally pushes dw into the peers' all-reduce receive slots
// (`push` is a TfdePush; see allreduce_p2p.cu for the slot layout). GEMM -> reduce-scatter in one kernel.
TFDE_API int tfde_dense_wgrad_push(const float* x, const float* dy, const float* ymask, float* dw,
                                   float* db, int M, int K, int N, float* ws, size_t ws_bytes,
                                   unsigned* counters, const TfdePush* push, cudaStream_t stream) {
  GemmParams p{};
  if (push != nullptr) p.push = *push;
  if (p.push.enabled && (p.push.world < 2 || p.push.world > 8)) return TFDE_ERR_INVALID;
  p.A = x; p.lda = K;
  p.B = dy; p.Bmask = ymask; p.ldb = N;
  p.C = dw; p.ldc = N; p.C2 = db;
  p.M = db ? K + 1 : K; p.N = N; p.K = M;
  p.ones_row = db ? K : -1;
  p.ws = ws; p.counters = counters;
  p.splits = (ws && counters) ? choose_splits(p.M, N, M, ws_bytes, kMaxCounters) : 1;
  return dispatch(A_DENSE_T, B_DENSE, p, stream);
}
TFDE_API size_t tfde_push_desc_bytes() { return sizeof(TfdePush); }

// dx[M,K] = (dy[M,N] * (ymask > 0)) . w[K,N]^T
TFDE_API int tfde_dense_dgrad(const float* dy, const float* ymask, const float* w, float* dx,
                              int M, int K, int N, float* ws, size_t ws_bytes,
                              unsigned* counters, cudaStream_t stream) {
  GemmParams p{};
  p.A = dy; p.Amask = ymask; p.lda = N;
  p.B = w; p.ldb = N;               // B_op[n, k_in] = w[k_in*N + n]
  p.C = dx; p.ldc = K;
  p.M = M; p.N = K; p.K = N; p.ones_row = -1;
  p.ws = ws; p.counters = counters;
  p.splits = (ws && counters) ? choose_splits(M, K, N, ws_bytes, kMaxCounters) : 1;
  return dispatch(A_DENSE, B_DENSE_T, p, stream);
}

static ConvGeom make_geom(int H, int W, int C, int KH, int KW, int Ho, int Wo, int sh, int sw,
                          int pt, int pl, int Cout) {
  ConvGeom g; g.H = H; g.W = W; g.C = C; g.KH = KH; g.KW = KW; g.Ho = Ho; g.Wo = Wo;
  g.sh = sh; g.sw = sw; g.pt = pt; g.pl = pl; g.Cout = Cout;
  return g;
}

// y[B,Ho,Wo,Cout] = act(conv(x[B,H,W,C], w[KH,KW,C,Cout]) + bias)
TFDE_API int tfde_conv2d_fwd(const float* x, const float* w, const float* bias, float* y, int Bn,
                             int H, int W, int C, int KH, int KW, int Ho, int Wo, int sh, int sw,
                             int pt, int pl, int Cout, int relu, float* ws, size_t ws_bytes,
                             unsigned* counters, cudaStream_t stream) {
  GemmParams p{};
  p.g = make_geom(H, W, C, KH, KW, Ho, Wo, sh, sw, pt, pl, Cout);
  p.A = x; p.B = w; p.ldb = Cout; p.C = y; p.ldc = Cout; p.bias = bias;
  p.M = Bn * Ho * Wo; p.N = Cout; p.K = KH * KW * C; p.ones_row = -1; p.relu = relu;
  p.ws = ws; p.counters = counters;
  p.splits = (ws && counters) ? choose_splits(p.M, p.N, p.K, ws_bytes, kMaxCounters) : 1;
  return dispatch(A_IM2COL, B_DENSE, p, stream);
}

// dw[KH,KW,C,Cout] = im2col(x)^T . (dy * (ymask>0)); db optional
TFDE_API int tfde_conv2d_wgrad(const float* x, const float* dy, const float* ymask, float* dw,
                               float* db, int Bn, int H, int W, int C, int KH, int KW, int Ho,
                               int Wo, int sh, int sw, int pt, int pl, int Cout, float* ws,
                               size_t ws_bytes, unsigned* counters, cudaStream_t stream) {
  GemmParams p{};
  p.g = make_geom(H, W, C, KH, KW, Ho, Wo, sh, sw, pt, pl, Cout);
  const int Kc = KH * KW * C;
  p.A = x; p.B = dy; p.Bmask = ymask; p.ldb = Cout; p.C = dw; p.ldc = Cout; p.C2 = db;
  p.M = db ? Kc + 1 : Kc; p.N = Cout; p.K = Bn * Ho * Wo; p.ones_row = db ? Kc : -1;
  p.ws = ws; p.counters = counters;
  p.splits = (ws && counters) ? choose_splits(p.M, p.N, p.K, ws_bytes, kMaxCounters) : 1;
  return dispatch(A_IM2COL_T, B_DENSE, p, stream);
}

// dx[B,H,W,C] = gather(dy * (ymask>0)) . rot(w)
TFDE_API int tfde_conv2d_dgrad(const float* dy, const float* ymask, const float* w, float* dx,
                               int Bn, int H, int W, int C, int KH, int KW, int Ho, int Wo, int sh,
                               int sw, int pt, int pl, int Cout, float* ws, size_t ws_bytes,
                               unsigned* counters, cudaStream_t stream) {
  GemmParams p{};
  p.g = make_geom(H, W, C, KH, KW, Ho, Wo, sh, sw, pt, pl, Cout);
  p.A = dy; p.Amask = ymask; p.B = w; p.C = dx; p.ldc = C;
  p.M = Bn * H * W; p.N = C; p.K = KH * KW * Cout; p.ones_row = -1;
  p.ws = ws; p.counters = counters;
  p.splits = (ws && counters) ? choose_splits(p.M, p.N, p.K, ws_bytes, kMaxCounters) : 1;
  return dispatch(A_DGRAD, B_WROT, p, stream);
}

TFDE_API int tfde_gemm_max_counters() { return kMaxCounters; }

// ---------------------------------------------------------------------------------------------
// Dense forward for callers that are not able to allocate device memory.
//
// Who this is for: code that is handed its input and output buffers and nothing else. Typical cases are
// a plug-in inside a serving process whose memory pool is owned (and fully reserved) by the host
// framework -- TensorFlow, which the reference scripts run on, grabs almost all GPU memory at start-up
// unless told otherwise (see the per_process_gpu_memory_fraction note at
// mnist_keras_distributed.py:170-174), so a cudaMalloc from a co-resident library simply fails -- and
// latency-critical loops in which cudaMalloc/cudaFree are off limits because they synchronise the device.
// tfde_dense_fwd needs split-K scratch (partial sums + tickets) from its caller; this variant brings its
// own, a fixed array that lives in the library image.
//
// What you must know before using it:
//  * The scratch is ONE file-scope __device__ array (kStaticScratchBytes, about 2 MiB). Every GPU that
//    loads this library keeps that much memory reserved until the process exits.
//  * Problems whose split-K schedule needs more than kStaticPartialBytes of partial sums are refused
//    (return value 1 == cudaErrorInvalidValue, reason printed on stderr); nothing is launched. Use
//    tfde_dense_fwd with your own workspace for those.
//  * All calls on a device share the array. The tickets are cleared at the start of every call, the
//    partial sums are overwritten before they are read, so back-to-back calls on one stream are fine.
//    Two calls that can run at the same time on the GPU (different streams, or graph branches with no
//    dependency between them) race on the array and produce garbage.
//  * The very first call on each device looks up the device address of the array (a host-side symbol is
//    not a device pointer). Make that first call outside of any stream capture, with the device you
//    intend to use current.
//  * No cudaMalloc, cudaFree or other allocating routine is called, here or in the kernel.
// ---------------------------------------------------------------------------------------------
#include <cstdio>
#include <mutex>

namespace {
constexpr size_t kStaticTicketBytes = (size_t)TFDE_MAX_COUNTERS * sizeof(unsigned);  // 16 KiB
constexpr size_t kStaticPartialBytes = (size_t)2 << 20;                              // 2 MiB
constexpr size_t kStaticScratchBytes = kStaticTicketBytes + kStaticPartialBytes;
static __device__ __align__(256) unsigned char s_static_scratch[kStaticScratchBytes];

constexpr int kMaxDevicesTracked = 64;
unsigned char* g_static_scratch_dev[kMaxDevicesTracked];  // device address per GPU, filled on first use
std::mutex g_static_scratch_mu;

// Device address of s_static_scratch on the current device (naming the array in host code would give the
// host-side shadow symbol, which a kernel must never be handed).
int static_scratch_address(unsigned char** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  if (dev < 0 || dev >= kMaxDevicesTracked) return TFDE_ERR_INVALID;
  std::lock_guard<std::mutex> lock(g_static_scratch_mu);
  if (g_static_scratch_dev[dev] == nullptr) {
    void* ptr = nullptr;
    e = cudaGetSymbolAddress(&ptr, s_static_scratch);
    if (e != cudaSuccess) return (int)e;
    g_static_scratch_dev[dev] = static_cast<unsigned char*>(ptr);
  }
  *out = g_static_scratch_dev[dev];
  return 0;
}
}  // namespace

// y[M,N] = act(x[M,K] . w[K,N] + bias), same math and kernel as tfde_dense_fwd, scratch taken from the
// library's own static device array. See the block comment above for limits and hazards.
TFDE_API int tfde_dense_fwd_static_scratch(const float* x, const float* w, const float* bias, float* y,
                                           int M, int K, int N, int relu, cudaStream_t stream) {
  if (M <= 0 || N <= 0 || K <= 0 || !x || !w || !y) return TFDE_ERR_INVALID;
  // The split-K schedule this shape would get with unlimited workspace decides how much scratch it needs.
  const int splits = choose_splits(M, N, K, (size_t)-1, kMaxCounters);
  const size_t need = splits > 1 ? (size_t)splits * M * ((N + 3) & ~3) * sizeof(float) : 0;
  if (need > kStaticPartialBytes) {
    fprintf(stderr,
            "tfde_dense_fwd_static_scratch: M=%d K=%d N=%d needs %zu bytes of split-K scratch but the "
            "built-in array holds %zu; call tfde_dense_fwd with a caller-provided workspace instead\n",
            M, K, N, need, kStaticPartialBytes);
    return TFDE_ERR_INVALID;
  }
  unsigned char* base = nullptr;
  const int rc = static_scratch_address(&base);
  if (rc != 0) return rc;
  unsigned* tickets = reinterpret_cast<unsigned*>(base);
  float* partials = reinterpret_cast<float*>(base + kStaticTicketBytes);
  if (splits > 1) {
    // tickets must read zero when the kernel starts; the array keeps whatever the previous call
    // (possibly an aborted one) left behind.
    const cudaError_t e = cudaMemsetAsync(tickets, 0, kStaticTicketBytes, stream);
    if (e != cudaSuccess) return (int)e;
  }
  GemmParams p{};
  p.A = x; p.lda = K; p.B = w; p.ldb = N; p.C = y; p.ldc = N; p.bias = bias;
  p.M = M; p.N = N; p.K = K; p.ones_row = -1; p.relu = relu;
  p.ws = partials; p.counters = tickets; p.splits = splits;
  return dispatch(A_DENSE, B_DENSE, p, stream);
}
