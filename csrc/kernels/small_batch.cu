// Latency-optimised kernels for the reference's own operating point: per-replica batch 64
// (distributed_with_keras.py:13), where a training step is ~180 MFLOP and every kernel is bound by
// exposed memory latency and by the number of dependent round trips, not by FLOPs or bandwidth.
//
//  1. dense_skinny_fwd  -- y[M<=64, N<=64 tile] = act(x . W + b) with a LONG reduction (K = 5408 for the
//     Dense(64) of Model A, distributed_with_keras.py:37). Split-K over ~all SMs; every CTA issues all of
//     its loads at once with cp.async; the partial tiles of the 8 CTAs of a thread-block CLUSTER are
//     reduced through distributed shared memory (no global round trip), and only one partial per cluster
//     reaches L2, where the last-arriving CTA of each 8-row band sums them in cluster order.
//  2. head_fwd_bwd      -- the classifier head in ONE kernel: logits = h . W2 + b2 (Dense(10),
//     distributed_with_keras.py:38 / Dense(10, softmax), mnist_keras_distributed.py:108), softmax
//     cross-entropy + accuracy (from logits or from probabilities), d(logits), dW2 = h^T . dz, db2 and
//     dh = (dz . W2^T) * (h > 0). Replaces four launches (G4 + G9 + G5 of SURVEY.md section 2.7).
//
// All reductions have a fixed order (bit-reproducible, eager == graph replay).
#include "common.cuh"
#include <cooperative_groups.h>

namespace cg = cooperative_groups;

namespace {

// ================================================================================================
// 1. skinny dense forward with cluster / DSMEM split-K
// ================================================================================================
constexpr int SK_CLUSTER = 8;
constexpr int SK_TM = 64, SK_TN = 64;
constexpr int SK_KMAX = 64;                 // max K slice per CTA
constexpr int SK_XLD = SK_KMAX + 4;         // padded row stride of the x slice (multiple of 4 floats)
constexpr int SK_THREADS = 256;
constexpr int SK_BAND = SK_TM / SK_CLUSTER; // rows of the tile each CTA of a cluster finalises (8)

struct SkinnyParams {
  const float* x; const float* w; const float* bias; float* y;
  int M, N, K;
  int kper;          // K elements per CTA (multiple of 4, <= SK_KMAX)
  int nclusters;     // clusters along K
  int relu;
  float* ws;         // [tiles][nclusters][SK_TM*SK_TN] cluster partials
  unsigned* counters;  // [tiles][SK_CLUSTER]
};

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
  const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() {
  asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
}

__global__ void __cluster_dims__(SK_CLUSTER, 1, 1) __launch_bounds__(SK_THREADS)
dense_skinny_fwd_kernel(const SkinnyParams p) {
  TFDE_PDL_PROLOGUE();
  __shared__ __align__(16) float xs[SK_TM][SK_XLD];      // x slice  [64 rows][kper]
  __shared__ __align__(16) float wsm[SK_KMAX][SK_TN];    // W slice  [kper][64 cols]
  // this CTA's partial tile (read by cluster peers) re-uses the x slice once the products are done
  // (static shared memory is capped at 48 KB; xs holds 64*68 >= 64*64 floats)
  float (*ps)[SK_TN] = reinterpret_cast<float (*)[SK_TN]>(&xs[0][0]);
  __shared__ int s_last;

  cg::cluster_group cluster = cg::this_cluster();
  const int t = threadIdx.x;
  const unsigned crank = cluster.block_rank();
  const int cluster_id = blockIdx.x / SK_CLUSTER;
  const int tile = blockIdx.y;                           // tile index over (m tiles x n tiles)
  const int tiles_n = (p.N + SK_TN - 1) / SK_TN;
  const int m0 = (tile / tiles_n) * SK_TM, n0 = (tile % tiles_n) * SK_TN;
  const int k0 = (int)blockIdx.x * p.kper;
  const int klen = max(0, min(p.kper, p.K - k0));        // multiple of 4 because K % 4 == 0

  // ---- issue every load of this CTA at once (16-byte cp.async), zero-fill what is out of range ----
  const int kq = p.kper >> 2;                            // float4 per x row slice
  for (int e = t; e < SK_TM * kq; e += SK_THREADS) {
    const int r = e / kq, c4 = e % kq;
    float* dst = &xs[r][c4 * 4];
    if (m0 + r < p.M && c4 * 4 < klen) cp_async16(dst, p.x + (size_t)(m0 + r) * p.K + k0 + c4 * 4);
    else *reinterpret_cast<float4*>(dst) = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  for (int e = t; e < p.kper * (SK_TN / 4); e += SK_THREADS) {
    const int k = e / (SK_TN / 4), c4 = e % (SK_TN / 4);
    float* dst = &wsm[k][c4 * 4];
    if (k < klen && n0 + c4 * 4 < p.N) cp_async16(dst, p.w + (size_t)(k0 + k) * p.N + n0 + c4 * 4);
    else *reinterpret_cast<float4*>(dst) = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  cp_async_wait_all();
  __syncthreads();

  // ---- partial 64x64 tile: 4x4 outputs per thread ---------------------------------------------
  const int tx = t & 15, ty = t >> 4;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  for (int k = 0; k < p.kper; ++k) {
    const float4 b = *reinterpret_cast<const float4*>(&wsm[k][tx * 4]);
    float a[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) a[i] = xs[ty * 4 + i][k];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      acc[i][0] = fmaf(a[i], b.x, acc[i][0]); acc[i][1] = fmaf(a[i], b.y, acc[i][1]);
      acc[i][2] = fmaf(a[i], b.z, acc[i][2]); acc[i][3] = fmaf(a[i], b.w, acc[i][3]);
    }
  }
  __syncthreads();                                       // everyone is done reading xs before it becomes ps
#pragma unroll
  for (int i = 0; i < 4; ++i)
    *reinterpret_cast<float4*>(&ps[ty * 4 + i][tx * 4]) = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);

  // ---- cluster reduction through distributed shared memory -----------------------------------------
  cluster.sync();                                        // every partial tile of the cluster is in smem
  // this CTA finalises rows [crank*8, crank*8+8) of the tile: 512 floats, 2 per thread
  const int brow = (int)crank * SK_BAND + (t >> 5);      // 8 rows, 32 threads x float2 per row
  const int bcol = (t & 31) * 2;
  float2 s = make_float2(0.f, 0.f);
#pragma unroll
  for (unsigned peer = 0; peer < SK_CLUSTER; ++peer) {   // fixed peer order
    const float* remote = cluster.map_shared_rank(&ps[0][0], peer);
    const float2 v = *reinterpret_cast<const float2*>(remote + brow * SK_TN + bcol);
    s.x += v.x; s.y += v.y;
  }
  cluster.sync();                                        // nobody leaves while its smem is being read

  const int m = m0 + brow, n = n0 + bcol;
  if (p.nclusters > 1) {
    float* part = p.ws + ((size_t)tile * p.nclusters + cluster_id) * (SK_TM * SK_TN);
    __stcg(reinterpret_cast<float2*>(part + brow * SK_TN + bcol), s);
    __threadfence();
    __syncthreads();
    unsigned* ctr = p.counters + tile * SK_CLUSTER + crank;
    if (t == 0) s_last = (atomicAdd(ctr, 1u) == (unsigned)p.nclusters - 1u);
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    const float* base = p.ws + (size_t)tile * p.nclusters * (SK_TM * SK_TN) + brow * SK_TN + bcol;
    s = make_float2(0.f, 0.f);
    for (int c0 = 0; c0 < p.nclusters; c0 += 8) {        // 8 independent L2 loads in flight
      float2 v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u)
        v[u] = (c0 + u < p.nclusters) ? __ldcg(reinterpret_cast<const float2*>(base + (size_t)(c0 + u) * (SK_TM * SK_TN)))
                                      : make_float2(0.f, 0.f);
#pragma unroll
      for (int u = 0; u < 8; ++u) { s.x += v[u].x; s.y += v[u].y; }   // cluster order
    }
    if (t == 0) *ctr = 0u;
  }
  if (m < p.M) {
    float v0 = s.x, v1 = s.y;
    if (p.bias) { if (n < p.N) v0 += __ldg(p.bias + n); if (n + 1 < p.N) v1 += __ldg(p.bias + n + 1); }
    if (p.relu) { v0 = fmaxf(v0, 0.f); v1 = fmaxf(v1, 0.f); }
    if (n < p.N) p.y[(size_t)m * p.N + n] = v0;
    if (n + 1 < p.N) p.y[(size_t)m * p.N + n + 1] = v1;
  }
}

// ================================================================================================
// 2. fused classifier head: Dense(N<=32) + softmax cross-entropy + backward of that Dense
// ================================================================================================
constexpr int HD_ROWS_MAX = 64;  // max rows of h per block (small batches use 16 for more CTAs)
constexpr int HD_THREADS = 256;
constexpr int HD_NMAX = 32;

struct HeadParams {
  const float* h; const float* w; const float* bias; const void* labels; int label_is_i64;
  float* out;        // logits (from_probs == 0) or probabilities (from_probs == 1), [M,N], may be NULL
  float* dh;         // [M,K], may be NULL (no upstream gradient needed)
  float* dw; float* db;   // [K,N], [N] (db may be NULL)
  float* out4;       // [loss*scale, correct, rows, 0]
  int M, K, N;
  int relu_mask;     // multiply dh by (h > 0) (the producing layer had a fused ReLU)
  int from_probs;
  int rpb;           // rows per block (16 or 64)
  float scale;
  float* partials;   // [gridDim.x][K*N + N + 2]
  unsigned* counter;
};

// dynamic shared memory: hs[rpb][K+1] | wsm[K][N] | zs[rpb][N] (logits -> probs -> dz) | bias | red
__global__ void __launch_bounds__(HD_THREADS)
head_fwd_bwd_kernel(const HeadParams p) {
  TFDE_PDL_PROLOGUE();
  extern __shared__ __align__(16) float hsm[];
  const int K = p.K, N = p.N, KP = K + 1;
  const int HD_ROWS = p.rpb;
  float* hs = hsm;
  float* wsm = hs + HD_ROWS * KP;
  float* zs = wsm + K * N;
  float* bsm = zs + HD_ROWS * N;      // bias [N]
  float* red = bsm + N;               // [2 * 8] loss / correct per warp
  __shared__ int s_last;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int r0 = blockIdx.x * HD_ROWS;
  const int rows = min(HD_ROWS, p.M - r0);

  for (int e = t; e < HD_ROWS * K; e += HD_THREADS) {
    const int r = e / K, k = e % K;
    hs[r * KP + k] = (r < rows) ? __ldg(p.h + (size_t)(r0 + r) * K + k) : 0.f;
  }
  for (int e = t; e < K * N; e += HD_THREADS) wsm[e] = __ldg(p.w + e);
  if (t < N) bsm[t] = p.bias ? __ldg(p.bias + t) : 0.f;
  __syncthreads();

  // logits: thread -> (row = e % 64, col = e / 64)
  for (int e = t; e < HD_ROWS * N; e += HD_THREADS) {
    const int r = e % HD_ROWS, c = e / HD_ROWS;
    float a = bsm[c];
    const float* hr = hs + r * KP;
    for (int k = 0; k < K; ++k) a = fmaf(hr[k], wsm[k * N + c], a);
    zs[r * N + c] = a;
  }
  __syncthreads();

  // softmax / loss / accuracy / dz: one warp per row, lane = class (N <= 32)
  float loss = 0.f, correct = 0.f;
  for (int r = warp; r < rows; r += HD_THREADS / 32) {
    const int yraw = p.label_is_i64 ? (int)reinterpret_cast<const long long*>(p.labels)[r0 + r]
                                    : reinterpret_cast<const int*>(p.labels)[r0 + r];
    // a label outside [0, N) poisons the loss with NaN (TensorFlow's GPU kernel does the same) and gets no gradient
    const bool yok = (unsigned)yraw < (unsigned)N;
    const int y = yok ? yraw : 0;
    const float z = lane < N ? zs[r * N + lane] : -INFINITY;
    float mx = z; int am = lane < N ? lane : 0x7fffffff;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, mx, o);
      const int oa = __shfl_xor_sync(0xffffffffu, am, o);
      if (ov > mx || (ov == mx && oa < am)) { mx = ov; am = oa; }
    }
    const float ex = lane < N ? expf(z - mx) : 0.f;
    const float sum = warp_sum(ex);
    const float prob = ex / sum;
    const float py = __shfl_sync(0xffffffffu, prob, y);
    const float zy = __shfl_sync(0xffffffffu, z, y);
    float dz;
    if (!p.from_probs) {
      dz = (prob - (lane == y ? 1.f : 0.f)) * p.scale;
      if (lane == 0) loss += (mx + logf(sum)) - zy;
    } else {
      const float cl = fminf(fmaxf(py, 1e-7f), 1.f - 1e-7f);
      const bool inside = (py >= 1e-7f) && (py <= 1.f - 1e-7f);
      dz = inside ? (prob - (lane == y ? 1.f : 0.f)) * (p.scale * py / cl) : 0.f;
      if (lane == 0) loss += -logf(cl);
    }
    if (!yok) { dz = 0.f; if (lane == 0) loss = __int_as_float(0x7fc00000); }
    if (lane == 0 && yok && am == y) correct += 1.f;
    if (lane < N) {
      if (p.out) p.out[(size_t)(r0 + r) * N + lane] = p.from_probs ? prob : z;
      zs[r * N + lane] = dz;
    }
  }
  // rows beyond `rows` (tail block) contribute nothing
  for (int e = t; e < (HD_ROWS - rows) * N; e += HD_THREADS) zs[rows * N + e] = 0.f;
  if (lane == 0) { red[warp] = loss; red[8 + warp] = correct; }
  __syncthreads();

  // dh = (dz . W^T) * (h > 0)
  if (p.dh != nullptr) {
    for (int e = t; e < rows * K; e += HD_THREADS) {
      const int r = e / K, k = e % K;
      float a = 0.f;
#pragma unroll 4
      for (int c = 0; c < N; ++c) a = fmaf(zs[r * N + c], wsm[k * N + c], a);
      if (p.relu_mask && !(hs[r * KP + k] > 0.f)) a = 0.f;
      p.dh[(size_t)(r0 + r) * K + k] = a;
    }
  }
  // dW partial [K,N], db partial [N], loss/correct partial
  float* part = p.partials + (size_t)blockIdx.x * (K * N + N + 2);
  const bool single = gridDim.x == 1;
  for (int e = t; e < K * N + N; e += HD_THREADS) {
    float a = 0.f;
    if (e < K * N) {
      const int k = e / N, c = e % N;
      for (int r = 0; r < HD_ROWS; ++r) a = fmaf(hs[r * KP + k], zs[r * N + c], a);
      if (single) p.dw[e] = a; else __stcg(part + e, a);
    } else {
      const int c = e - K * N;
      for (int r = 0; r < HD_ROWS; ++r) a += zs[r * N + c];
      if (single) { if (p.db) p.db[c] = a; } else __stcg(part + e, a);
    }
  }
  if (t == 0) {
    float l = 0.f, c = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) { l += red[i]; c += red[8 + i]; }
    if (single) { p.out4[0] = l * p.scale; p.out4[1] = c; p.out4[2] = (float)p.M; p.out4[3] = 0.f; }
    else { __stcg(part + K * N + N, l); __stcg(part + K * N + N + 1, c); }
  }
  if (single) return;
  __threadfence();
  __syncthreads();
  if (t == 0) s_last = (atomicAdd(p.counter, 1u) == gridDim.x - 1);
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  const int per = K * N + N + 2;
  for (int e = t; e < per; e += HD_THREADS) {
    float a = 0.f;
    for (unsigned b0 = 0; b0 < gridDim.x; b0 += 8) {
      float v[8];
#pragma unroll
      for (unsigned u = 0; u < 8; ++u) v[u] = (b0 + u < gridDim.x) ? __ldcg(p.partials + (size_t)(b0 + u) * per + e) : 0.f;
#pragma unroll
      for (unsigned u = 0; u < 8; ++u) a += v[u];
    }
    if (e < K * N) p.dw[e] = a;
    else if (e < K * N + N) { if (p.db) p.db[e - K * N] = a; }
    else if (e == K * N + N) p.out4[0] = a * p.scale;
    else p.out4[1] = a;
  }
  if (t == 0) { p.out4[2] = (float)p.M; p.out4[3] = 0.f; *p.counter = 0u; }
}

size_t head_smem_bytes(int K, int N, int rpb = HD_ROWS_MAX) {
  return ((size_t)rpb * (K + 1) + (size_t)K * N + (size_t)rpb * N + N + 16) * sizeof(float);
}
int head_rows_per_block(int M) {
  static int forced = -1;      // TFDE_HEAD_RPB=16|64 overrides the heuristic (experiments)
  if (forced < 0) { const char* e = getenv("TFDE_HEAD_RPB"); forced = e ? atoi(e) : 0; }
  if (forced == 4 || forced == 8 || forced == 16 || forced == 32 || forced == 64) return forced;
  return M <= 256 ? 16 : HD_ROWS_MAX;
}

// ================================================================================================
// 3. skinny dense backward: dx, dW and db of Dense(K -> 64) at batch <= 64 in ONE launch
// ================================================================================================
// The batch-64 operating point of the reference (distributed_with_keras.py:37: Dense(64) on 5408 features).
// The generic GEMM path runs the filter gradient and the data gradient as two 85-block kernels of ~17 us each
// (forked); both are one 64x64x64 tile per block and pure latency. Here block b owns input features
// [64b, 64b+64): it stages g = dy * (y > 0) [64 x 64] once, then
//   dx[m, k] = sum_n g[m, n] * W[k, n]   (W rows k are contiguous: the Keras (in, out) layout)
//   dW[k, n] = sum_m x[m, k] * g[m, n]
// with the second operand tile re-using the same shared memory, and block 0 also writes db = sum_m g.
// Thread (ty, tx) of 16x16 owns the 4x4 outputs {ty+16i} x {tx+16j}: conflict-free shared reads, coalesced
// 64-byte global segments. Deterministic (no split, fixed order).
constexpr int SB_T = 64;

__global__ void __launch_bounds__(256)
dense_bwd_skinny_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ dy,
                        const float* __restrict__ ymask, float* __restrict__ dx, float* __restrict__ dw,
                        float* __restrict__ db, int M, int K) {
  TFDE_PDL_PROLOGUE();
  __shared__ float gs[SB_T][SB_T + 1];      // g[m][n]
  __shared__ float ts[SB_T][SB_T + 1];      // W tile [k][n], then x tile [m][k]
  const int t = threadIdx.x, tx = t & 15, ty = t >> 4;
  const int k0 = blockIdx.x * SB_T;
  // stage g and the W tile (float4 global loads, scalar shared stores because of the +1 padding)
  for (int e = t; e < SB_T * SB_T / 4; e += 256) {
    const int r = e >> 4, c4 = (e & 15) << 2;
    float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
    if (r < M) {
      g = __ldg(reinterpret_cast<const float4*>(dy + (size_t)r * SB_T + c4));
      if (ymask != nullptr) {
        const float4 y = __ldg(reinterpret_cast<const float4*>(ymask + (size_t)r * SB_T + c4));
        if (!(y.x > 0.f)) g.x = 0.f; if (!(y.y > 0.f)) g.y = 0.f; if (!(y.z > 0.f)) g.z = 0.f; if (!(y.w > 0.f)) g.w = 0.f;
      }
    }
    gs[r][c4] = g.x; gs[r][c4 + 1] = g.y; gs[r][c4 + 2] = g.z; gs[r][c4 + 3] = g.w;
    float4 wv = make_float4(0.f, 0.f, 0.f, 0.f);
    if (k0 + r < K) wv = __ldg(reinterpret_cast<const float4*>(w + (size_t)(k0 + r) * SB_T + c4));
    ts[r][c4] = wv.x; ts[r][c4 + 1] = wv.y; ts[r][c4 + 2] = wv.z; ts[r][c4 + 3] = wv.w;
  }
  // prefetch this thread's part of the x tile into registers while the data gradient is computed
  float4 xr[4];
#pragma unroll
  for (int u = 0; u < 4; ++u) {
    const int e = t + u * 256, r = e >> 4, c4 = (e & 15) << 2;
    xr[u] = (r < M && k0 + c4 < K) ? __ldg(reinterpret_cast<const float4*>(x + (size_t)r * K + k0 + c4))
                                   : make_float4(0.f, 0.f, 0.f, 0.f);
  }
  __syncthreads();
  if (blockIdx.x == 0 && db != nullptr && t < SB_T) {
    float s = 0.f;
    for (int m = 0; m < SB_T; ++m) s += gs[m][t];
    db[t] = s;
  }
  float acc[4][4];
  if (dx != nullptr) {
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
#pragma unroll 8
    for (int n = 0; n < SB_T; ++n) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = gs[ty + 16 * i][n];
#pragma unroll
      for (int j = 0; j < 4; ++j) b[j] = ts[tx + 16 * j][n];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int m = ty + 16 * i;
      if (m < M) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int k = k0 + tx + 16 * j;
          if (k < K) dx[(size_t)m * K + k] = acc[i][j];
        }
      }
    }
  }
  __syncthreads();                           // every thread is done with the W tile
#pragma unroll
  for (int u = 0; u < 4; ++u) {
    const int e = t + u * 256, r = e >> 4, c4 = (e & 15) << 2;
    ts[r][c4] = xr[u].x; ts[r][c4 + 1] = xr[u].y; ts[r][c4 + 2] = xr[u].z; ts[r][c4 + 3] = xr[u].w;
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
#pragma unroll 8
  for (int m = 0; m < SB_T; ++m) {
    float a[4], b[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) a[i] = ts[m][ty + 16 * i];      // x[m][k], k = ty + 16 i
#pragma unroll
    for (int j = 0; j < 4; ++j) b[j] = gs[m][tx + 16 * j];      // g[m][n], n = tx + 16 j
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int k = k0 + ty + 16 * i;
    if (k < K) {
#pragma unroll
      for (int j = 0; j < 4; ++j) dw[(size_t)k * SB_T + tx + 16 * j] = acc[i][j];
    }
  }
}

}  // namespace

// ---------------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------------
// 1 if tfde_dense_skinny_fwd handles this shape (otherwise use tfde_dense_fwd).
TFDE_API int tfde_dense_skinny_supported(int M, int K, int N) {
  const int tiles = ((M + SK_TM - 1) / SK_TM) * ((N + SK_TN - 1) / SK_TN);
  return (K % 4 == 0 && N % 4 == 0 && K >= 512 && tiles >= 1 && tiles <= 8) ? 1 : 0;
}
TFDE_API size_t tfde_dense_skinny_ws_floats(int M, int K, int N) {
  const int tiles = ((M + SK_TM - 1) / SK_TM) * ((N + SK_TN - 1) / SK_TN);
  return (size_t)tiles * 32 * SK_TM * SK_TN;
}

// y[M,N] = act(x[M,K] . w[K,N] + bias) for small M*N and long K (cluster split-K through DSMEM).
// ws: tfde_dense_skinny_ws_floats() floats; counters: >= tiles*8 zeroed tickets (self-cleaning).
TFDE_API int tfde_dense_skinny_fwd(const float* x, const float* w, const float* bias, float* y, int M, int K,
                                   int N, int relu, float* ws, unsigned* counters, cudaStream_t stream) {
  if (!tfde_dense_skinny_supported(M, K, N)) return TFDE_ERR_INVALID;
  const int tiles = ((M + SK_TM - 1) / SK_TM) * ((N + SK_TN - 1) / SK_TN);
  // K slices: as many CTAs as fit ~one wave (148 SMs / tiles), whole clusters of 8, slice <= 64, multiple of 4
  int max_ctas = (148 / tiles) / SK_CLUSTER * SK_CLUSTER;
  if (max_ctas < SK_CLUSTER) max_ctas = SK_CLUSTER;
  int kper = (K + max_ctas - 1) / max_ctas;
  kper = (kper + 3) & ~3;
  if (kper < 16) kper = 16;
  int nclusters;
  if (kper > SK_KMAX) {               // K too long for one wave at 64 per CTA: more clusters (up to 32)
    kper = SK_KMAX;
  }
  int ctas = (K + kper - 1) / kper;
  nclusters = (ctas + SK_CLUSTER - 1) / SK_CLUSTER;
  if (nclusters > 32) return TFDE_ERR_INVALID;   // K > 16384: not a job for this kernel
  SkinnyParams p{x, w, bias, y, M, N, K, kper, nclusters, relu, ws, counters};
  dim3 grid(nclusters * SK_CLUSTER, tiles);
  tfde_launch(dense_skinny_fwd_kernel, grid, dim3(SK_THREADS), 0, stream, p);
  TFDE_RETURN_LAUNCH_STATUS();
}

TFDE_API int tfde_head_supported(int K, int N) {
  return (N >= 1 && N <= HD_NMAX && K >= 1 && head_smem_bytes(K, N) <= 200 * 1024) ? 1 : 0;
}
TFDE_API size_t tfde_head_ws_floats(int M, int K, int N) {
  const int rpb = head_rows_per_block(M);
  return (size_t)((M + rpb - 1) / rpb) * ((size_t)K * N + N + 2);
}

// Fused Dense(N) + sparse categorical cross-entropy (+accuracy) + backward of the Dense.
// The first call with a given (K, N) may raise the kernel's dynamic shared-memory limit: make it outside
// of stream capture.
TFDE_API int tfde_head_fwd_bwd(const float* h, const float* w, const float* bias, const void* labels,
                               int label_is_i64, float* out, float* dh, float* dw, float* db, float* out4,
                               int M, int K, int N, int relu_mask, int from_probs, float scale,
                               float* partials, unsigned* counter, cudaStream_t stream) {
  if (!tfde_head_supported(K, N) || M <= 0) return TFDE_ERR_INVALID;
  const int rpb = head_rows_per_block(M);
  const size_t smem = head_smem_bytes(K, N, rpb);
  static size_t configured[TFDE_MAX_DEVICES] = {};
  { const cudaError_t e = tfde_ensure_dyn_smem(head_fwd_bwd_kernel, smem, configured); if (e != cudaSuccess) return (int)e; }
  HeadParams p{h, w, bias, labels, label_is_i64, out, dh, dw, db, out4, M, K, N, relu_mask, from_probs, rpb, scale,
               partials, counter};
  tfde_launch(head_fwd_bwd_kernel, dim3((M + rpb - 1) / rpb), dim3(HD_THREADS), smem, stream, p);
  TFDE_RETURN_LAUNCH_STATUS();
}

// 1 if tfde_dense_bwd_skinny handles this shape: batch <= 64 rows, exactly 64 units, K % 4 == 0.
TFDE_API int tfde_dense_bwd_skinny_supported(int M, int K, int N) {
  return (M >= 1 && M <= SB_T && N == SB_T && K >= 4 && (K & 3) == 0) ? 1 : 0;
}

// Backward of Dense(K -> 64) at batch <= 64 in one launch: dx [M,K] (may be NULL), dw [K,64], db [64] (may be
// NULL); ymask = the layer's own output when it had a fused ReLU (gradient masked by y > 0), else NULL.
// ref: the Dense(64, relu) of distributed_with_keras.py:37 (cuBLAS GEMMs + Eigen reduce in TF).
TFDE_API int tfde_dense_bwd_skinny(const float* x, const float* w, const float* dy, const float* ymask, float* dx,
                                   float* dw, float* db, int M, int K, int N, cudaStream_t stream) {
  if (!tfde_dense_bwd_skinny_supported(M, K, N) || !x || !w || !dy || !dw) return TFDE_ERR_INVALID;
  tfde_launch(dense_bwd_skinny_kernel, dim3((K + SB_T - 1) / SB_T), dim3(256), 0, stream, x, w, dy, ymask, dx,
              dw, db, M, K);
  TFDE_RETURN_LAUNCH_STATUS();
}

