// Whole training step of "Model A" in ONE persistent cooperative kernel (optionally K steps per launch).
//
// Model (distributed_with_keras.py:33-43): Conv2D(32,3,relu) -> MaxPooling2D() -> Flatten -> Dense(64,relu)
// -> Dense(ncls) ; SparseCategoricalCrossentropy(from_logits=True) ; SGD (momentum / Adam as extras);
// per-replica batch <= 64 (distributed_with_keras.py:13). The reference runs this step as ~25 TensorFlow ops
// plus an NCCL all-reduce (distributed_with_keras.py:16,63); round 1 of this repository ran it as 6 dependent
// kernels (68 us, every one of them latency-bound). Here the step is one launch:
//
//   grid = 169 CTAs = one per pooled pixel (13x13); CTA p owns input features [32p, 32p+32) of the Dense(64)
//   layer, i.e. rows [32p, 32p+32) of its (5408, 64) kernel -- 8 KB of weights that only this CTA ever reads
//   or writes. The 1.38 MB activation tensor and its gradient never exist in global memory.
//
//   P1  conv3x3+bias+ReLU+maxpool for pixel p, all rows of the batch (kept in shared memory for the backward
//       pass), then the split-K partial  h_p[64,64] = act_p[64,32] . W_p[32,64]        -> global partials
//   B1  grid barrier
//   P2  CTA b < 64 reduces row b of h over the 169 partials (fixed order), + bias, ReLU, and runs the classifier
//       head for that row: logits, softmax-CE, d(logits), dh[b] = d(logits).Wh^T * (h>0)    -> global (20 KB)
//   B2  grid barrier
//   P3  every CTA: dW_p = act_p^T . dh (+ optimizer / reduce-scatter push), dx_p = dh . W_p^T, pool routing + ReLU
//       mask + conv filter-gradient partial for pixel p. CTAs 0..15 also produce the head-kernel gradient rows,
//       CTA 16 the bias gradients and the loss / accuracy scalars.
//   T   two-level ticket (13 groups of 13 CTAs): the last CTA of a group reduces its group's conv partials, the
//       last group winner reduces the 13 group partials, applies the optimizer to the small tensors (conv
//       kernel+bias, dense bias, head kernel+bias), writes the metrics tail, advances the step counter and
//       releases the next step.
//
// Multi-GPU (MirroredStrategy / MultiWorkerMirroredStrategy, G12): the gradient all-reduce is part of the SAME
// kernel, slice by slice, over peer memory (NVLink 5 / NVSwitch) -- GEMM -> reduce-scatter -> all-gather with
// no NCCL call and no separate launch:
//   * slice p is owned by rank p % W. Right after computing dW_p a non-owner stores it into the owner's
//     receive buffer; it then carries on with dx / the conv backward while the bytes fly.
//   * after its ticket the owner sums the W contributions in rank order (bit-identical replicas) and stores the
//     reduced slice into every peer's `gsum` buffer; every rank applies the optimizer to its own copy of slice p
//     (optimizer state stays replicated: checkpoints need no gather). A non-owner picks the reduced slice up
//     lazily, under the NEXT step's convolution, right before its GEMM needs the new weights.
//   * the small tensors + the 4 metric scalars (G14) are exchanged one-shot by the ticket winner
//     (double-buffered by step parity).
//   * transport is flag-in-data: every 8 bytes on the wire carry 4 bytes of payload and the step's epoch, so there
//     is no separate flag, no st.release.sys fence (an NVLink round trip) -- one-way latency only.
// All waits are bounded (status word + abort), flags carry a monotonically increasing epoch kept in device
// memory, so the kernel can be replayed from a CUDA graph.
//
// Numerics: exact fp32 FFMA everywhere (the reference trains in fp32), fixed summation orders.
#include "common.cuh"
#include "umma_tf32.cuh"
#include <cstdio>
#include <cstdlib>
#include <mutex>

namespace {

constexpr int FS_PIX = 169;       // pooled pixels = CTAs
constexpr int FS_PW = 13;
constexpr int FS_HW = 28;         // input height/width
constexpr int FS_THREADS = 256;
constexpr int FS_B = 64;          // max rows of the per-replica batch
constexpr int FS_C = 32;          // conv channels
constexpr int FS_U = 64;          // dense units
constexpr int FS_NC = 16;         // max classes (padded head width)
constexpr int FS_SLICE = FS_C * FS_U;        // 2048 floats of the dense kernel per CTA
constexpr int FS_ND = FS_PIX * FS_SLICE;     // 346112 floats
constexpr int FS_MAX_RANKS = 8;
constexpr int FS_SMALL_MAX = 2048;           // floats of the packed small-tensor block (>= 288+32+64+64*16+16+4)
constexpr int FS_FLAG_STRIDE = 176;          // >= FS_PIX, multiple of 16

// state words (each counter on its own 128-byte line)
constexpr int ST_BAR = 0, ST_TICKET = 32, ST_DONE = 64, ST_EPOCH0 = 96, ST_GROUP = 128;   // ST_GROUP + 32*g, g < 13
constexpr int FS_GROUPS = 13;                // two-level ticket: 13 groups of 13 CTAs
constexpr int FS_STATE_WORDS = ST_GROUP + 32 * FS_GROUPS;
constexpr int FS_PROF_STAMPS = 16;
// global scratch: hbuf = h [64][64] | dh [64][64] | dz [64][16] | per-row loss, correct [64][2]
constexpr int HB_H = 0, HB_DH = FS_B * FS_U, HB_DZ = 2 * FS_B * FS_U, HB_STAT = HB_DZ + FS_B * FS_NC;
constexpr int FS_HBUF_FLOATS = HB_STAT + 2 * FS_B;
// cpart = per-CTA conv partials [169][320] | per-group partials [13][320]
constexpr int FS_CPART_FLOATS = (FS_PIX + FS_GROUPS) * 320;

// shared memory layout (floats)
constexpr int PATCH_LD = 17, ACT_LD = 68, WS_LD = 68, WH_LD = 17, DH_LD = 68, DX_LD = 33, DZ_LD = 17;
constexpr int SM_PATCH = 0;
constexpr int SM_ACT = SM_PATCH + FS_B * PATCH_LD;      // actT [32][68]
constexpr int SM_ARG = SM_ACT + FS_C * ACT_LD;          // argw [64][4] u32
constexpr int SM_WS = SM_ARG + FS_B * 4;                // ws   [32][68]
constexpr int SM_DH = SM_WS + FS_C * WS_LD;             // dh   [64][68]  (tail: staging of the reduced small block)
constexpr int SM_DX = SM_DH + FS_B * DH_LD;             // dxs  [64][33]
constexpr int SM_CW = SM_DX + FS_B * DX_LD;             // cw   [10][32]
constexpr int SM_RED = SM_CW + 10 * FS_C;               // red  [8][10][32] (P2: [16][16] float4 ; tail: [80] float4)
constexpr int SM_WH = SM_RED + 8 * 10 * FS_C;           // whs  [64][17] + bh [16]      (row CTAs only)
constexpr int SM_ROW = SM_WH + FS_U * WH_LD + FS_NC;    // hrow [64]
constexpr int SM_HG = SM_ROW + FS_U;                    // hcol [64][4] + dzs [64][17]  (head-gradient CTAs only)
constexpr int SM_TOTAL = SM_HG + FS_B * 4 + FS_B * DZ_LD;
constexpr int FS_SMALL_F4 = (288 + 32 + 64 + 64 * FS_NC + FS_NC + 4 + 3) / 4;    // float4 groups of the largest small block
static_assert(SM_DH + FS_MAX_RANKS * FS_SMALL_F4 * 4 <= SM_TOTAL, "tail exchange staging must fit in dh..end");
constexpr int FS_STAGE_FLOATS = FS_MAX_RANKS * FS_SMALL_F4 * 4;
static_assert(SM_ACT % 4 == 0 && SM_WS % 4 == 0 && SM_DH % 4 == 0 && SM_RED % 4 == 0 && SM_HG % 4 == 0, "align");

// Shared memory layout of the TENSOR-CORE variant (TC): the three per-CTA GEMMs run as kind::tf32 tcgen05.mma with the
// accumulators in TMEM. Operands are the no-swizzle K-major core-block tiles of umma_tf32.cuh; a tile that is needed in
// both orientations is written twice (activations, weight slice, gradient tile). The M = 128 instruction shape reads
// 64 rows past each A tile: the tiles are ordered so that those reads stay inside the allocation.
constexpr int T_PATCH = 0;
constexpr int T_ACTK = T_PATCH + FS_B * PATCH_LD;        // act  [b 64][c 32]  RB 256   A of the forward GEMM
constexpr int T_WTK = T_ACTK + 2048;                     // W^T  [n 64][c 32]  RB 256   B of the forward GEMM
constexpr int T_ACTB = T_WTK + 2048;                     // act^T[c 32][b 64]  RB 512   B of the filter gradient
constexpr int T_WSB = T_ACTB + 2048;                     // W    [c 32][n 64]  RB 512   B of the data gradient
constexpr int T_DHB = T_WSB + 2048;                      // dh   [b 64][n 64]  RB 512   A of the data gradient
constexpr int T_DHT = T_DHB + 4096;                      // dh^T [n 64][b 64]  RB 512   A of the filter gradient
constexpr int T_ARG = T_DHT + 4096;
constexpr int T_DX = T_ARG + FS_B * 4;
constexpr int T_CW = T_DX + FS_B * DX_LD;
constexpr int T_RED = T_CW + 10 * FS_C;
constexpr int T_WH = T_RED + 8 * 10 * FS_C;
constexpr int T_ROW = T_WH + FS_U * WH_LD + FS_NC;
constexpr int T_HG = T_ROW + FS_U;
constexpr int T_WSX = T_HG + FS_B * 4 + FS_B * DZ_LD;    // exact fp32 master copy of the weight slice [c 32][68] (the MMA tiles hold TF32-rounded copies)
constexpr int T_GB = T_RED;                              // dW staging [32][68] (TMEM rows are units, threads want 2x4 tiles); aliases `red`
constexpr int T_TOTAL = T_WSX + FS_C * WS_LD;
static_assert(FS_C * 68 <= 8 * 10 * FS_C, "dW staging must fit in red");
static_assert(T_ACTK % 4 == 0 && T_DHB % 4 == 0 && T_RED % 4 == 0 && T_HG % 4 == 0 && T_GB % 4 == 0, "align");
static_assert(T_DHT + 2 * 4096 <= T_TOTAL, "M=128 over-read of the dh^T tile must stay inside the allocation");
static_assert(T_DHB + 8 * ((288 + 32 + 64 + 64 * FS_NC + FS_NC + 4 + 3) / 4) * 4 <= T_TOTAL, "tail exchange staging (TC)");
constexpr int FS_TMEM_COLS = 128;                        // D1 [0,64) | dW^T [64,96) | dx [96,128)

struct Hyper { float lr, mu, b1, b2, eps, gs, lr_t; };

// The CTA's copy of its dense-kernel slice in shared memory: plain [c][68] (FFMA variant) or the two blocked tiles.
template <bool TC>
struct WTile {
  float* base;      // FFMA: ws ; TC: sm (tile offsets are added here)
  __device__ __forceinline__ float4 load(int c, int n) const {
    if (TC) return *reinterpret_cast<const float4*>(base + T_WSX + c * WS_LD + n);
    return *reinterpret_cast<const float4*>(base + c * WS_LD + n);
  }
  __device__ __forceinline__ void store(int c, int n, const float4 v) const {
    if (TC) {
      *reinterpret_cast<float4*>(base + T_WSX + c * WS_LD + n) = v;
      const float4 r = make_float4(umma::tf32r(v.x), umma::tf32r(v.y), umma::tf32r(v.z), umma::tf32r(v.w));
      *reinterpret_cast<float4*>(base + T_WSB + umma::blk(c, n, 512)) = r;
      float* wt = base + T_WTK;
      wt[umma::blk(n, c, 256)] = r.x; wt[umma::blk(n + 1, c, 256)] = r.y;
      wt[umma::blk(n + 2, c, 256)] = r.z; wt[umma::blk(n + 3, c, 256)] = r.w;
    } else {
      *reinterpret_cast<float4*>(base + c * WS_LD + n) = v;
    }
  }
};

struct FsParams {
  const void* x; const long long* y;
  int x_is_u8; float in_scale; long x_step_bytes;
  int B, nsteps, ncls, opt;
  float* w; float* g; float* s1; float* s2;
  long off_ck, off_cb, off_dk, off_db, off_hk, off_hb, n;
  float loss_scale;
  TfdeHParams* hp; float* metrics; float* tails;
  float* part; float* hbuf; float* cpart; float* small;
  uint32_t* state; int* status;
  unsigned long long timeout_ns;
  int rank, world;
  float* rs[FS_MAX_RANKS];        // rs[o]: owner o's receive buffer [world][FS_ND]
  float* gsum[FS_MAX_RANKS];      // gsum[r]: rank r's reduced dense-kernel gradient [FS_ND]
  float* srecv[FS_MAX_RANKS];     // srecv[r]: rank r's small-block receive slots [2][world][FS_SMALL_MAX]
  uint32_t* flags[FS_MAX_RANKS];  // flags[r]: RS [world][176] | AG [176] | S [world]
  unsigned long long* prof;       // optional [FS_PIX][FS_PROF_STAMPS] globaltimer stamps of the LAST step of a launch
  // streamed inputs (end-to-end path): ready[s] becomes >= ready_tag once step s's images and labels have landed in
  // x / y -- written by a 4-byte copy queued behind that step's copies on the copy stream, while this kernel runs
  const uint32_t* ready; uint32_t ready_tag;
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_gpu(uint32_t* p, uint32_t v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_gpu(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// Bounded spin until *flag >= target (wrap-safe). SYS: the flag is written by a peer GPU. Returns false on
// timeout or when another CTA already recorded a failure (the whole grid then drains quickly).
#define FS_STAMP(i) do { if (p.prof && threadIdx.x == 0 && s == (p.nsteps > 1 ? p.nsteps - 2 : 0)) \
    p.prof[(long)blockIdx.x * FS_PROF_STAMPS + (i)] = globaltimer_ns(); } while (0)

template <bool SYS>
__device__ __forceinline__ bool wait_flag(const FsParams& p, const uint32_t* flag, uint32_t target, int code,
                                          int who, int where) {
  const unsigned long long t0 = globaltimer_ns();
  unsigned spins = 0;
  while ((int)((SYS ? ld_acquire_sys(flag) : ld_acquire_gpu(flag)) - target) < 0) {
    if ((++spins & 0xff) == 0) {
      if (*reinterpret_cast<volatile int*>(p.status) != 0) return false;
      if (globaltimer_ns() - t0 > p.timeout_ns) {
        if (atomicCAS(p.status, 0, code) == 0) { p.status[1] = (int)target; p.status[2] = who; p.status[3] = where; }
        return false;
      }
    }
  }
  return true;
}

// Grid-wide barrier on a monotonically increasing arrival counter (all CTAs are co-resident: cooperative launch).
__device__ __forceinline__ void grid_barrier(const FsParams& p, uint32_t target, int where) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    atomicAdd(p.state + ST_BAR, 1u);
    wait_flag<false>(p, p.state + ST_BAR, target, 10, (int)blockIdx.x, where);
  }
  __syncthreads();
}

// ---- flag-in-data ("LL") transport over peer memory ------------------------------------------------------------
// A float4 travels as two 16-byte stores {d0, e, d1, e} {d2, e, d3, e}: every 8-byte half carries the epoch it belongs
// to, so the receiver needs no separate flag and the sender no fence (a st.release.sys costs an NVLink round trip
// before the flag even leaves) -- one-way latency only, at twice the bytes, the right trade for 8 KB slices.
// `slot` is the LL image of a float array: float4 index i lives at uint4 index 2i.
__device__ __forceinline__ void ll_store4(float* slot, long i4, const float4 v, uint32_t e) {
  uint4* dst = reinterpret_cast<uint4*>(slot) + 2 * i4;
  // plain (weak) 16-byte stores: each one is a single aligned transaction carrying its own epoch words, so neither
  // ordering between them nor a system-scope qualifier is needed (strong .sys stores to peer memory issue far more
  // slowly: 28 of them per thread cost 4.6 us in the 8-GPU tail)
  asm volatile("st.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(dst), "r"(__float_as_uint(v.x)), "r"(e),
               "r"(__float_as_uint(v.y)), "r"(e) : "memory");
  asm volatile("st.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(dst + 1), "r"(__float_as_uint(v.z)), "r"(e),
               "r"(__float_as_uint(v.w)), "r"(e) : "memory");
}
__device__ __forceinline__ float4 ll_load4(const FsParams& p, const float* slot, long i4, uint32_t e, int code, int who,
                                           int where) {
  const uint4* src = reinterpret_cast<const uint4*>(slot) + 2 * i4;
  uint4 a, b;
  const unsigned long long t0 = globaltimer_ns();
  unsigned spins = 0;
  for (;;) {
    asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(a.x), "=r"(a.y), "=r"(a.z), "=r"(a.w) : "l"(src) : "memory");
    asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(b.x), "=r"(b.y), "=r"(b.z), "=r"(b.w) : "l"(src + 1) : "memory");
    if (a.y == e && a.w == e && b.y == e && b.w == e) break;
    if ((++spins & 0xff) == 0) {
      if (*reinterpret_cast<volatile int*>(p.status) != 0) break;
      if (globaltimer_ns() - t0 > p.timeout_ns) {
        if (atomicCAS(p.status, 0, code) == 0) { p.status[1] = (int)e; p.status[2] = who; p.status[3] = where; }
        break;
      }
    }
  }
  return make_float4(__uint_as_float(a.x), __uint_as_float(a.z), __uint_as_float(b.x), __uint_as_float(b.z));
}

// Optimizer arithmetic with every rounding spelled out (no FMA-contraction freedom): the same update is inlined at
// several call sites (owner / non-owner, immediate / deferred) and replicas must stay bit-identical, so the compiler
// must not be free to fuse `a*b + c*d` one way here and the other way there.
__device__ __forceinline__ float update_one(int opt, float w, float g, float* s1, float* s2, long i, const Hyper& h) {
  g = __fmul_rn(g, h.gs);
  if (opt == 0) return __fmaf_rn(-h.lr, g, w);
  if (opt == 1 || opt == 2) {
    const float lg = __fmul_rn(h.lr, g);
    const float v = __fmaf_rn(h.mu, s1[i], -lg);
    s1[i] = v;
    return __fadd_rn(w, opt == 2 ? __fmaf_rn(h.mu, v, -lg) : v);
  }
  const float m = __fmaf_rn(h.b1, s1[i], __fmul_rn(1.f - h.b1, g));
  const float v = __fmaf_rn(h.b2, s2[i], __fmul_rn(__fmul_rn(1.f - h.b2, g), g));
  s1[i] = m; s2[i] = v;
  return __fsub_rn(w, __fdiv_rn(__fmul_rn(h.lr_t, m), __fadd_rn(__fsqrt_rn(v), h.eps)));
}

__device__ __forceinline__ float4 update_four(int opt, float4 w, float4 g, float* s1, float* s2, long i, const Hyper& h) {
  float4 o;
  o.x = update_one(opt, w.x, g.x, s1, s2, i, h);     o.y = update_one(opt, w.y, g.y, s1, s2, i + 1, h);
  o.z = update_one(opt, w.z, g.z, s1, s2, i + 2, h); o.w = update_one(opt, w.w, g.w, s1, s2, i + 3, h);
  return o;
}

// packed small block: [ck 288][cb 32][db 64][hk 64*ncls][hb ncls][tail 4]
__device__ __forceinline__ int small_len(int ncls) { return 288 + 32 + 64 + 64 * ncls + ncls + 4; }
__device__ __forceinline__ long small_to_arena(const FsParams& p, int i) {
  if (i < 288) return p.off_ck + i;
  i -= 288; if (i < 32) return p.off_cb + i;
  i -= 32;  if (i < 64) return p.off_db + i;
  i -= 64;  if (i < 64 * p.ncls) return p.off_hk + i;
  i -= 64 * p.ncls; if (i < p.ncls) return p.off_hb + i;
  return p.n + (i - p.ncls);
}

// Completes the cross-replica reduction of this CTA's dense-kernel slice for epoch `e` and applies the optimizer:
// the owner sums the W contributions in rank order and stores the reduced slice into every peer's `gsum`; a
// non-owner picks the reduced slice up from its own `gsum`. Every rank then updates its copy (global memory and the
// shared-memory copy `ws` that the next step's GEMMs read).
template <bool TC>
__device__ __forceinline__ void finish_slice(const FsParams& p, uint32_t e, int pix, int owner, const float (&gacc)[2][4],
                                             const WTile<TC> ws, float* gslice, float* wslice, long aslice, const Hyper& h) {
  const int t = threadIdx.x, tc = t >> 4, tn = t & 15;
  const int W = p.world, R = p.rank;
  float4 sum[2];
  if (owner == R) {
    const float* base = p.rs[R] + (long)pix * (2 * FS_SLICE);
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const int i4 = (tc * 2 + i) * (FS_U / 4) + tn;
      float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int r0 = 0; r0 < W; r0 += 4) {          // rank order, four peers' loads in flight at a time
        float4 v[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int r = r0 + j;
          if (r < W)
            v[j] = (r == R) ? make_float4(gacc[i][0], gacc[i][1], gacc[i][2], gacc[i][3])
                            : ll_load4(p, base + (long)r * (2 * FS_ND), i4, e, 22, r, pix);
        }
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (r0 + j < W) { a.x += v[j].x; a.y += v[j].y; a.z += v[j].z; a.w += v[j].w; }
      }
      sum[i] = a;
      for (int k = 1; k < W; ++k) {
        const int peer = (R + k) % W;
        ll_store4(p.gsum[peer] + (long)pix * (2 * FS_SLICE), i4, a, e);
      }
    }
  } else {
#pragma unroll
    for (int i = 0; i < 2; ++i)
      sum[i] = ll_load4(p, p.gsum[R] + (long)pix * (2 * FS_SLICE), (tc * 2 + i) * (FS_U / 4) + tn, e, 23, owner, pix);
  }
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const int c = tc * 2 + i, off = c * FS_U + tn * 4;
    const float4 nw = update_four(p.opt, ws.load(c, tn * 4), sum[i], p.s1, p.s2, aslice + off, h);
    ws.store(c, tn * 4, nw);
    *reinterpret_cast<float4*>(gslice + off) = sum[i];
    *reinterpret_cast<float4*>(wslice + off) = nw;
  }
}

template <bool TC>
__global__ void __launch_bounds__(FS_THREADS, 2)
fused_step_kernel(const FsParams p) {
  extern __shared__ __align__(128) float sm[];
  float* patch = sm + (TC ? T_PATCH : SM_PATCH);
  float* actT = sm + (TC ? T_ACTB : SM_ACT);             // FFMA: [c][68] ; TC: blocked [c][b] tile
  uint32_t* argw = reinterpret_cast<uint32_t*>(sm + (TC ? T_ARG : SM_ARG));
  float* dh = sm + (TC ? T_DHB : SM_DH);                 // FFMA: [b][68] ; TC: blocked [b][n] tile (tail: staging)
  float* dxs = sm + (TC ? T_DX : SM_DX);
  float* cw = sm + (TC ? T_CW : SM_CW);
  float* red = sm + (TC ? T_RED : SM_RED);
  float* whs = sm + (TC ? T_WH : SM_WH);
  float* bh = whs + FS_U * WH_LD;
  float* hrow = sm + (TC ? T_ROW : SM_ROW);
  float* hcol = sm + (TC ? T_HG : SM_HG);
  float* dzs = hcol + FS_B * 4;
  const WTile<TC> ws{TC ? sm : sm + SM_WS};
  __shared__ int s_last;
  __shared__ uint64_t s_mbar;                            // TC: completion of the committed tcgen05.mma batch
  __shared__ uint32_t s_tmem;
  uint32_t mma_parity = 0;
  if (TC) {
    if (threadIdx.x < 32) umma::tmem_alloc(&s_tmem, FS_TMEM_COLS);
    if (threadIdx.x == 0) umma::mbar_init(&s_mbar, 1);
    umma::fence_before_sync();
    __syncthreads();
    umma::fence_after_sync();
  }
  const uint32_t tmem = TC ? s_tmem : 0u;

  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int pix = blockIdx.x, py = pix / FS_PW, px = pix - py * FS_PW;
  const int grid = gridDim.x;
  const int W = p.world, R = p.rank;
  const int owner = pix % W;
  const int N = p.ncls;
  const uint32_t epoch0 = p.state[ST_EPOCH0];
  float* gslice = p.g + p.off_dk + (long)pix * FS_SLICE;
  float* wslice = p.w + p.off_dk + (long)pix * FS_SLICE;
  const long aslice = p.off_dk + (long)pix * FS_SLICE;        // arena index of the slice (optimizer slots)
  const int nsmall = small_len(N);
  const bool row_cta = pix < FS_B;          // reduces row `pix` of h and runs the classifier head for it
  const bool hg_cta = pix < 16;             // computes 4 rows of the head kernel gradient
  const int grp = pix / FS_GROUPS;          // two-level ticket group
  __shared__ Hyper s_hprev;                 // hyper-parameters of the step whose slice exchange is still pending
  __shared__ int s_in_ready;                // streamed inputs: the next step's batch was on the device at this step's start
  bool pending = false;                     // W > 1: the previous step's slice is reduced/updated lazily, right
  float gacc[2][4];                         //        before this step's GEMM needs it (off the critical path)
  const int tc = t >> 4, tn = t & 15;
  float pv[4];                              // one row of this thread's 4x4 input patch (thread = (batch row, patch row))
  auto load_patch = [&](int step, float (&v)[4]) {
    const int b = t >> 2, row = t & 3;
#pragma unroll
    for (int j = 0; j < 4; ++j) v[j] = 0.f;
    if (b < p.B) {
      const long off = ((long)b * FS_HW + (2 * py + row)) * FS_HW + 2 * px;
      if (p.x_is_u8) {
        const uint8_t* xb = reinterpret_cast<const uint8_t*>(p.x) + (long)step * p.x_step_bytes + off;
#pragma unroll
        for (int j = 0; j < 4; ++j) v[j] = (float)__ldg(xb + j) * p.in_scale;
      } else {
        const float* xf = reinterpret_cast<const float*>(reinterpret_cast<const char*>(p.x) + (long)step * p.x_step_bytes) + off;
#pragma unroll
        for (int j = 0; j < 4; ++j) v[j] = __ldg(xf + j) * p.in_scale;
      }
    }
  };

  for (int s = 0; s < p.nsteps; ++s) {
    const uint32_t e = epoch0 + (uint32_t)s + 1u;
    const uint32_t bar_base = (e - 1u) * 2u * (uint32_t)grid;
    // hyper-parameters of this step (the ticket winner of the previous step advanced them before releasing us)
    Hyper h;
    float b1p, b2p; uint32_t hp_step;
    {
      const volatile TfdeHParams* hp = p.hp;
      h.lr = hp->lr; h.mu = hp->momentum; h.b1 = hp->beta1; h.b2 = hp->beta2; h.eps = hp->eps; h.gs = hp->grad_scale;
      b1p = hp->beta1_pow * h.b1; b2p = hp->beta2_pow * h.b2; hp_step = hp->step;
      h.lr_t = (p.opt == 3) ? h.lr * sqrtf(1.f - b2p) / (1.f - b1p) : h.lr;
    }
    FS_STAMP(0);
    if (p.ready != nullptr) {
      // streamed inputs: the first step of a launch waits for its batch; every step looks ahead (non-blocking) whether
      // the NEXT batch is already here, so the prefetch at its end does not have to ask (steady state: the copy
      // engine is a whole step ahead)
      if (t == 0) {
        if (s == 0) wait_flag<true>(p, p.ready, p.ready_tag, 12, pix, 0);
        s_in_ready = (s + 1 < p.nsteps) && (int)(ld_acquire_sys(p.ready + s + 1) - p.ready_tag) >= 0;
      }
      if (s == 0) __syncthreads();
    }
    // ================================ P1: conv + pool + split-K dense partial =========================
    int ylab = -1;
    {
      // dense kernel slice (8 KB) -> registers now, shared memory after the conv weights are staged
      // (first step of a launch only: afterwards the shared-memory copy is kept current by this CTA's own updates)
      float4 wreg[2];
      if (s == 0) {
#pragma unroll
        for (int i = 0; i < 2; ++i) wreg[i] = __ldcg(reinterpret_cast<const float4*>(wslice) + t + i * FS_THREADS);
      }
      for (int i = t; i < 320; i += FS_THREADS)
        cw[i] = __ldcg(i < 288 ? p.w + p.off_ck + i : p.w + p.off_cb + (i - 288));
      if (row_cta) {
        if (t == 0 && pix < p.B) ylab = (int)p.y[(long)s * p.B + pix];
        for (int i = t; i < FS_U * FS_NC + FS_NC; i += FS_THREADS) {   // head kernel [64][ncls] -> [64][17], bias
          float v = 0.f;
          if (i < FS_U * FS_NC) { const int k = i >> 4, c = i & 15; if (c < N) v = __ldcg(p.w + p.off_hk + k * N + c); whs[k * WH_LD + c] = v; }
          else { const int c = i - FS_U * FS_NC; if (c < N) v = __ldcg(p.w + p.off_hb + c); bh[c] = v; }
        }
      }
      {
        // this step's 4x4 input patches: loaded here for the first step of a launch, otherwise already in registers
        // (requested before the wait for the previous step's completion: the HBM latency hides behind the tail)
        const int b = t >> 2, row = t & 3;
        if (s == 0) load_patch(0, pv);
#pragma unroll
        for (int j = 0; j < 4; ++j) patch[b * PATCH_LD + row * 4 + j] = pv[j];
      }
      if (s == 0) {
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          const int idx4 = t + i * FS_THREADS, c = idx4 >> 4, n4 = idx4 & 15;
          ws.store(c, n4 * 4, wreg[i]);
        }
      }
    }
    __syncthreads();
    {
      // conv: thread = (row b, 8 channels); weights broadcast from shared memory
      const int b = t & 63, cg = t >> 6;
      float pr[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) pr[i] = patch[b * PATCH_LD + i];
      uint32_t bits = 0;
#pragma unroll
      for (int cc = 0; cc < 8; ++cc) {
        const int c = cg * 8 + cc;
        float wr[9];
#pragma unroll
        for (int i = 0; i < 9; ++i) wr[i] = cw[i * FS_C + c];
        const float bias = cw[9 * FS_C + c];
        float best = -1.f; int besti = 0;
#pragma unroll
        for (int dy = 0; dy < 2; ++dy)
#pragma unroll
          for (int dx = 0; dx < 2; ++dx) {
            float o = bias;
#pragma unroll
            for (int kh = 0; kh < 3; ++kh)
#pragma unroll
              for (int kw = 0; kw < 3; ++kw) o = fmaf(wr[kh * 3 + kw], pr[(dy + kh) * 4 + dx + kw], o);
            o = fmaxf(o, 0.f);
            if (o > best) { best = o; besti = dy * 2 + dx; }
          }
        if (TC) {
          const float br = umma::tf32r(best);
          actT[umma::blk(c, b, 512)] = br;
          sm[T_ACTK + umma::blk(b, c, 256)] = br;
        } else {
          actT[c * ACT_LD + b] = best;
        }
        bits |= (uint32_t)besti << (2 * cc);
      }
      argw[b * 4 + cg] = bits;
    }
    if (pending) {
      // the previous step's slice: its reduce-scatter contributions / reduced values arrived while the tail and this
      // step's convolution ran
      finish_slice<TC>(p, e - 1u, pix, owner, gacc, ws, gslice, wslice, aslice, s_hprev);
      pending = false;
    }
    if (TC) {
      // split-K partial on the tensor core: D1[b, n] = sum_c act[b, c] * W_p[c, n]  (4 x UMMA 128x64x8, TF32)
      umma::fence_operands();
      __syncthreads();
      if (t == 0) {
        umma::fence_after_sync();
        const uint32_t a0 = umma::smem_u32(sm + T_ACTK), b0 = umma::smem_u32(sm + T_WTK);
        const uint32_t id = umma::idesc_tf32(128, 64, 0, 0);
#pragma unroll
        for (int k = 0; k < 4; ++k)
          umma::mma_tf32(tmem, umma::desc(a0 + 2 * k * 128, 128, 1024), umma::desc(b0 + 2 * k * 128, 128, 1024), id, k > 0);
        umma::commit(&s_mbar);
      }
      if ((warp & 3) < 2) {
        // warps 0,1 read columns [0,32) of rows 0..63, warps 4,5 columns [32,64): one accumulator row per thread
        umma::mbar_wait(&s_mbar, mma_parity);
        umma::fence_after_sync();
        const int b = (warp & 3) * 32 + lane, col0 = (warp >> 2) * 32;
        float r[32];
        umma::tmem_ld32(tmem + ((uint32_t)((warp & 3) * 32) << 16) + col0, r);
        float4* dst = reinterpret_cast<float4*>(p.part + (long)pix * (FS_B * FS_U) + b * FS_U + col0);
#pragma unroll
        for (int j = 0; j < 8; ++j) __stcg(dst + j, make_float4(r[4 * j], r[4 * j + 1], r[4 * j + 2], r[4 * j + 3]));
        umma::fence_before_sync();
      }
      mma_parity ^= 1u;
    } else {
    __syncthreads();
    {
      // split-K partial: h_p[b, n] = sum_c act[b, c] * W_p[c, n]; thread tile 4 rows x 4 units
      const int tb = t >> 4, tn = t & 15;
      float acc[4][4];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
#pragma unroll 8
      for (int c = 0; c < FS_C; ++c) {
        const float4 a = *reinterpret_cast<const float4*>(actT + c * ACT_LD + tb * 4);
        const float4 wv = ws.load(c, tn * 4);
        const float av[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          acc[i][0] = fmaf(av[i], wv.x, acc[i][0]); acc[i][1] = fmaf(av[i], wv.y, acc[i][1]);
          acc[i][2] = fmaf(av[i], wv.z, acc[i][2]); acc[i][3] = fmaf(av[i], wv.w, acc[i][3]);
        }
      }
      float4* dst = reinterpret_cast<float4*>(p.part + (long)pix * (FS_B * FS_U));
#pragma unroll
      for (int i = 0; i < 4; ++i)
        __stcg(dst + (tb * 4 + i) * (FS_U / 4) + tn, make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]));
    }
    }   // !TC
    FS_STAMP(1);
    grid_barrier(p, bar_base + (uint32_t)grid, 1);
    FS_STAMP(2);

    // ================================ P2: CTA b < 64 reduces row b of h and runs the head for it ==========
    if (row_cta) {
      const int b = pix;
      {
        const int f4 = t & 15, g16 = t >> 4;
        float4 v[11];
#pragma unroll
        for (int j = 0; j < 11; ++j) {
          const int q = g16 + 16 * j;
          v[j] = q < FS_PIX ? __ldcg(reinterpret_cast<const float4*>(p.part) + (long)q * 1024 + b * 16 + f4)
                            : make_float4(0.f, 0.f, 0.f, 0.f);
        }
        float4 a = v[0];
#pragma unroll
        for (int j = 1; j < 11; ++j) { a.x += v[j].x; a.y += v[j].y; a.z += v[j].z; a.w += v[j].w; }
        reinterpret_cast<float4*>(red)[g16 * 16 + f4] = a;
      }
      __syncthreads();
      if (t < 16) {
        float4 a = reinterpret_cast<const float4*>(red)[t];
#pragma unroll
        for (int g16 = 1; g16 < 16; ++g16) {
          const float4 v = reinterpret_cast<const float4*>(red)[g16 * 16 + t];
          a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w;
        }
        const float4 bv = __ldcg(reinterpret_cast<const float4*>(p.w + p.off_db) + t);
        a.x = fmaxf(a.x + bv.x, 0.f); a.y = fmaxf(a.y + bv.y, 0.f);
        a.z = fmaxf(a.z + bv.z, 0.f); a.w = fmaxf(a.w + bv.w, 0.f);
        *reinterpret_cast<float4*>(hrow + t * 4) = a;
        __stcg(reinterpret_cast<float4*>(p.hbuf + HB_H) + b * 16 + t, a);
      }
      __syncthreads();
      if (warp == 0) {
        // logits (lane = class), softmax cross-entropy, d(logits), dh row
        float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll 4
        for (int k = 0; k < FS_U; k += 4) {
          const float4 hv = *reinterpret_cast<const float4*>(hrow + k);
          a0 = fmaf(hv.x, whs[(k + 0) * WH_LD + lane], a0); a1 = fmaf(hv.y, whs[(k + 1) * WH_LD + lane], a1);
          a2 = fmaf(hv.z, whs[(k + 2) * WH_LD + lane], a2); a3 = fmaf(hv.w, whs[(k + 3) * WH_LD + lane], a3);
        }
        const int y = __shfl_sync(0xffffffffu, ylab, 0);
        const float z = lane < N ? ((a0 + a1) + (a2 + a3)) + bh[lane & 15] : -INFINITY;
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
        const bool live = b < p.B;
        const bool valid = live && (unsigned)y < (unsigned)N;
        const float zy = __shfl_sync(0xffffffffu, z, valid ? y : 0);
        const float dz = valid ? (prob - (lane == y ? 1.f : 0.f)) * p.loss_scale : 0.f;
        if (lane == 0) {
          // a label outside [0, ncls) poisons the loss (TensorFlow's GPU kernel returns NaN for it) and
          // contributes no gradient
          const float l = !live ? 0.f : (valid ? (mx + logf(sum)) - zy : __int_as_float(0x7fc00000));
          __stcg(p.hbuf + HB_STAT + 2 * b, l);
          __stcg(p.hbuf + HB_STAT + 2 * b + 1, (valid && am == y) ? 1.f : 0.f);
        }
        if (lane < FS_NC) __stcg(p.hbuf + HB_DZ + b * FS_NC + lane, dz);
        // dh[b, n] = (sum_c dz[c] * Wh[n, c]) * (h[b, n] > 0), n = lane and lane + 32
        float d0 = 0.f, d1 = 0.f;
#pragma unroll
        for (int c = 0; c < FS_NC; ++c) {
          const float dzc = __shfl_sync(0xffffffffu, dz, c);
          d0 = fmaf(dzc, whs[lane * WH_LD + c], d0);
          d1 = fmaf(dzc, whs[(lane + 32) * WH_LD + c], d1);
        }
        __stcg(p.hbuf + HB_DH + b * FS_U + lane, hrow[lane] > 0.f ? d0 : 0.f);
        __stcg(p.hbuf + HB_DH + b * FS_U + lane + 32, hrow[lane + 32] > 0.f ? d1 : 0.f);
      }
    }
    FS_STAMP(3);
    grid_barrier(p, bar_base + 2u * (uint32_t)grid, 2);
    FS_STAMP(4);

    // ================================ P3: backward of this CTA's slice ================================
    {
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int f4 = t + i * FS_THREADS, b = f4 >> 4, k4 = f4 & 15;
        float4 v = __ldcg(reinterpret_cast<const float4*>(p.hbuf + HB_DH) + f4);
        if (TC) {
          v = make_float4(umma::tf32r(v.x), umma::tf32r(v.y), umma::tf32r(v.z), umma::tf32r(v.w));
          *reinterpret_cast<float4*>(dh + umma::blk(b, k4 * 4, 512)) = v;
          float* dt = sm + T_DHT;
          dt[umma::blk(k4 * 4, b, 512)] = v.x; dt[umma::blk(k4 * 4 + 1, b, 512)] = v.y;
          dt[umma::blk(k4 * 4 + 2, b, 512)] = v.z; dt[umma::blk(k4 * 4 + 3, b, 512)] = v.w;
        } else {
          *reinterpret_cast<float4*>(dh + b * DH_LD + k4 * 4) = v;
        }
      }
      if (pix <= 16) {
        if (hg_cta && t < FS_B) *reinterpret_cast<float4*>(hcol + t * 4) = __ldcg(reinterpret_cast<const float4*>(p.hbuf + HB_H) + t * 16 + pix);
        const float4 v = __ldcg(reinterpret_cast<const float4*>(p.hbuf + HB_DZ) + t);
        float* d = dzs + (t >> 2) * DZ_LD + (t & 3) * 4;
        d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w;
      }
    }
    if (TC) umma::fence_operands();
    __syncthreads();
    FS_STAMP(5);
    if (TC && t == 0) {
      // filter gradient dW_p^T[n, c] = sum_b dh[b, n] * act[b, c] and data gradient dx_p[b, c] = sum_n dh[b, n] * W_p[c, n]
      // on the tensor core (2 x 8 UMMA 128x32x8, TF32); they run while CTAs 0..15 do the head gradients below
      umma::fence_after_sync();
      const uint32_t id = umma::idesc_tf32(128, 32, 0, 0);
      const uint32_t aT = umma::smem_u32(sm + T_DHT), bA = umma::smem_u32(sm + T_ACTB);
      const uint32_t aD = umma::smem_u32(sm + T_DHB), bW = umma::smem_u32(sm + T_WSB);
#pragma unroll
      for (int k = 0; k < 8; ++k)
        umma::mma_tf32(tmem + 64, umma::desc(aT + 2 * k * 128, 128, 2048), umma::desc(bA + 2 * k * 128, 128, 2048), id, k > 0);
#pragma unroll
      for (int k = 0; k < 8; ++k)
        umma::mma_tf32(tmem + 96, umma::desc(aD + 2 * k * 128, 128, 2048), umma::desc(bW + 2 * k * 128, 128, 2048), id, k > 0);
      umma::commit(&s_mbar);
    }
    if (pix <= 16) {
      // CTAs 0..15: head kernel gradient rows [4*pix, 4*pix+4): dWh[k, c] = sum_b h[b, k] * dz[b, c]; CTA 16: the bias
      // gradients and the loss / accuracy scalars. All 256 threads: the batch is split 4 ways, the four partial sums
      // are combined in a fixed order through shared memory (`red` is free until the conv-backward reduction).
      float* sp = p.small;
      const int q4 = t >> 6, u = t & 63;
      if (hg_cta) {
        const int kk = u >> 4, c = u & 15;
        float a = 0.f;
#pragma unroll
        for (int b = q4 * 16; b < q4 * 16 + 16; ++b) a = fmaf(hcol[b * 4 + kk], dzs[b * DZ_LD + c], a);
        red[q4 * 64 + u] = a;
      } else {
        float a = 0.f, z = 0.f;                                 // db[n] = sum_b dh[b, n] ; dbh[c] = sum_b dz[b, c]
#pragma unroll
        for (int b = q4 * 16; b < q4 * 16 + 16; ++b) {
          a += TC ? dh[umma::blk(b, u, 512)] : dh[b * DH_LD + u];
          if (u < FS_NC) z += dzs[b * DZ_LD + u];
        }
        red[q4 * 64 + u] = a;
        if (u < FS_NC) red[256 + q4 * FS_NC + u] = z;
      }
      __syncthreads();
      if (hg_cta) {
        if (t < 64 && (t & 15) < N)
          __stcg(sp + 384 + (4 * pix + (t >> 4)) * N + (t & 15), ((red[t] + red[64 + t]) + red[128 + t]) + red[192 + t]);
      } else {
        if (t < 64) __stcg(sp + 320 + t, ((red[t] + red[64 + t]) + red[128 + t]) + red[192 + t]);
        else if (t < 64 + N) {
          const int c = t - 64;
          __stcg(sp + 384 + FS_U * N + c, ((red[256 + c] + red[256 + FS_NC + c]) + red[256 + 2 * FS_NC + c]) + red[256 + 3 * FS_NC + c]);
        } else if (t >= 96 && t < 128) {
          // loss / correct: fixed-order tree over the 64 per-row values (warp 3)
          float l = __ldcg(p.hbuf + HB_STAT + 2 * lane) + __ldcg(p.hbuf + HB_STAT + 2 * (lane + 32));
          float c = __ldcg(p.hbuf + HB_STAT + 2 * lane + 1) + __ldcg(p.hbuf + HB_STAT + 2 * (lane + 32) + 1);
          l = warp_sum(l); c = warp_sum(c);
          if (lane == 0) {
            float* tl = sp + 384 + FS_U * N + N;
            __stcg(tl + 0, l * p.loss_scale); __stcg(tl + 1, c); __stcg(tl + 2, (float)p.B); __stcg(tl + 3, 0.f);
          }
        }
      }
      __syncthreads();     // `red` is reused by the TC staging tile / the conv-backward reduction
    }
    FS_STAMP(6);
    // dW_p[c, n] = sum_b act[b, c] * dh[b, n] ; thread tile 2 channels x 4 units (kept in registers until the
    // optimizer / the reduction consumes it)
    if (TC) {
      // accumulators come back one row per thread: dW^T rows (n) go through a small staging tile so that every thread
      // ends up with the same 2 x 4 tile of dW_p as in the FFMA variant; dx rows get the ReLU mask on the way out
      float* gb = sm + T_GB;
      if ((warp & 3) < 2) {
        umma::mbar_wait(&s_mbar, mma_parity);
        umma::fence_after_sync();
        const int row = (warp & 3) * 32 + lane;
        float r[32];
        if (warp < 4) {
          umma::tmem_ld32(tmem + ((uint32_t)((warp & 3) * 32) << 16) + 64, r);       // row = unit n, columns = channel c
#pragma unroll
          for (int c = 0; c < 32; ++c) gb[c * 68 + row] = r[c];
        } else {
          umma::tmem_ld32(tmem + ((uint32_t)((warp & 3) * 32) << 16) + 96, r);       // row = batch row b, columns = c
#pragma unroll
          for (int c = 0; c < 32; ++c) dxs[row * DX_LD + c] = actT[umma::blk(c, row, 512)] > 0.f ? r[c] : 0.f;
        }
        umma::fence_before_sync();
      }
      mma_parity ^= 1u;
      __syncthreads();
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        const float4 v = *reinterpret_cast<const float4*>(gb + (tc * 2 + i) * 68 + tn * 4);
        gacc[i][0] = v.x; gacc[i][1] = v.y; gacc[i][2] = v.z; gacc[i][3] = v.w;
      }
    } else
    {
#pragma unroll
      for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) gacc[i][j] = 0.f;
#pragma unroll 4
      for (int b = 0; b < FS_B; b += 4) {
        const float4 a0 = *reinterpret_cast<const float4*>(actT + (tc * 2) * ACT_LD + b);
        const float4 a1 = *reinterpret_cast<const float4*>(actT + (tc * 2 + 1) * ACT_LD + b);
        const float av0[4] = {a0.x, a0.y, a0.z, a0.w}, av1[4] = {a1.x, a1.y, a1.z, a1.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float4 d = *reinterpret_cast<const float4*>(dh + (b + i) * DH_LD + tn * 4);
          gacc[0][0] = fmaf(av0[i], d.x, gacc[0][0]); gacc[0][1] = fmaf(av0[i], d.y, gacc[0][1]);
          gacc[0][2] = fmaf(av0[i], d.z, gacc[0][2]); gacc[0][3] = fmaf(av0[i], d.w, gacc[0][3]);
          gacc[1][0] = fmaf(av1[i], d.x, gacc[1][0]); gacc[1][1] = fmaf(av1[i], d.y, gacc[1][1]);
          gacc[1][2] = fmaf(av1[i], d.z, gacc[1][2]); gacc[1][3] = fmaf(av1[i], d.w, gacc[1][3]);
        }
      }
    }
    if (W > 1 && owner != R) {
      // GEMM -> reduce-scatter: the tile goes straight into the owner's receive buffer over NVLink (flag-in-data)
      float* dst = p.rs[owner] + (long)R * (2 * FS_ND) + (long)pix * (2 * FS_SLICE);
#pragma unroll
      for (int i = 0; i < 2; ++i)
        ll_store4(dst, (tc * 2 + i) * (FS_U / 4) + tn, make_float4(gacc[i][0], gacc[i][1], gacc[i][2], gacc[i][3]), e);
    }
    if (!TC) {
      // dx_p[b, c] = sum_n dh[b, n] * W_p[c, n], masked by ReLU; thread tile 4 rows x 2 channels
      const int tb = t >> 4, tq = t & 15;
      float acc[4][2];
#pragma unroll
      for (int i = 0; i < 4; ++i) acc[i][0] = acc[i][1] = 0.f;
#pragma unroll 4
      for (int n4 = 0; n4 < FS_U / 4; ++n4) {
        const float4 w0 = ws.load(tq * 2, n4 * 4);
        const float4 w1 = ws.load(tq * 2 + 1, n4 * 4);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float4 d = *reinterpret_cast<const float4*>(dh + (tb * 4 + i) * DH_LD + n4 * 4);
          acc[i][0] = fmaf(d.x, w0.x, acc[i][0]); acc[i][0] = fmaf(d.y, w0.y, acc[i][0]);
          acc[i][0] = fmaf(d.z, w0.z, acc[i][0]); acc[i][0] = fmaf(d.w, w0.w, acc[i][0]);
          acc[i][1] = fmaf(d.x, w1.x, acc[i][1]); acc[i][1] = fmaf(d.y, w1.y, acc[i][1]);
          acc[i][1] = fmaf(d.z, w1.z, acc[i][1]); acc[i][1] = fmaf(d.w, w1.w, acc[i][1]);
        }
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int b = tb * 4 + i;
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const int c = tq * 2 + j;
          dxs[b * DX_LD + c] = actT[c * ACT_LD + b] > 0.f ? acc[i][j] : 0.f;
        }
      }
    }
    __syncthreads();
    if (W == 1) {
      // single replica: the optimizer is the epilogue of the filter-gradient tile (dx above consumed the old weights)
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        const int c = tc * 2 + i, off = c * FS_U + tn * 4;
        const float4 gv = make_float4(gacc[i][0], gacc[i][1], gacc[i][2], gacc[i][3]);
        const float4 nw = update_four(p.opt, ws.load(c, tn * 4), gv, p.s1, p.s2, aslice + off, h);
        ws.store(c, tn * 4, nw);
        *reinterpret_cast<float4*>(gslice + off) = gv;
        *reinterpret_cast<float4*>(wslice + off) = nw;
      }
    }
    {
      // conv filter/bias gradient partial of this pixel: lane = channel, warp strides over the rows
      float acc[10];
#pragma unroll
      for (int i = 0; i < 10; ++i) acc[i] = 0.f;
#pragma unroll 2
      for (int b = warp; b < FS_B; b += 8) {
        const float gv = dxs[b * DX_LD + lane];
        const int sel = (argw[b * 4 + (lane >> 3)] >> (2 * (lane & 7))) & 3;
        const int dy = sel >> 1, dx = sel & 1;
        const float* pb = patch + b * PATCH_LD + dy * 4 + dx;
#pragma unroll
        for (int kh = 0; kh < 3; ++kh)
#pragma unroll
          for (int kw = 0; kw < 3; ++kw) acc[kh * 3 + kw] = fmaf(gv, pb[kh * 4 + kw], acc[kh * 3 + kw]);
        acc[9] += gv;
      }
#pragma unroll
      for (int i = 0; i < 10; ++i) red[(warp * 10 + i) * FS_C + lane] = acc[i];
    }
    __syncthreads();
    for (int i = t; i < 320; i += FS_THREADS) {
      float a = 0.f;
#pragma unroll
      for (int wv = 0; wv < 8; ++wv) a += red[wv * 320 + i];
      __stcg(p.cpart + (long)pix * 320 + i, a);
    }
    // ================================ T: two-level ticket; the last CTA finishes the small tensors ===========
    FS_STAMP(7);
    __threadfence();
    __syncthreads();
    if (t == 0) s_last = (atomicAdd(p.state + ST_GROUP + 32 * grp, 1u) + 1u == e * (uint32_t)FS_GROUPS);
    __syncthreads();
    if (s_last) {
      // last CTA of its group of 13: reduce the group's conv partials (fixed order)
      __threadfence();
      if (t < 80) {
        float4 v[FS_GROUPS];
#pragma unroll
        for (int j = 0; j < FS_GROUPS; ++j)
          v[j] = __ldcg(reinterpret_cast<const float4*>(p.cpart + (long)(grp * FS_GROUPS + j) * 320) + t);
        float4 a = v[0];
#pragma unroll
        for (int j = 1; j < FS_GROUPS; ++j) { a.x += v[j].x; a.y += v[j].y; a.z += v[j].z; a.w += v[j].w; }
        __stcg(reinterpret_cast<float4*>(p.cpart + (long)(FS_PIX + grp) * 320) + t, a);
      }
      __threadfence();
      __syncthreads();
      if (t == 0) s_last = (atomicAdd(p.state + ST_TICKET, 1u) + 1u == e * (uint32_t)FS_GROUPS);
      __syncthreads();
    }
    FS_STAMP(8);
    if (s_last) {
      __threadfence();
      float* sp = p.small;
      float* stage = dh;                                   // packed small block in shared memory (dh is dead by now)
      float4* stage4 = reinterpret_cast<float4*>(stage);
      const int n4 = (nsmall + 3) >> 2;
      // everything the tail will need from L2 is requested up front, so the round trips overlap: the head / bias
      // gradients written by CTAs 0..16, the current values of the small parameters, the metric accumulators
      float4 early = make_float4(0.f, 0.f, 0.f, 0.f);
      const int ie = 80 + t;
      if (ie < n4) {
        early.x = __ldcg(sp + 4 * ie);
        early.y = 4 * ie + 1 < nsmall ? __ldcg(sp + 4 * ie + 1) : 0.f;
        early.z = 4 * ie + 2 < nsmall ? __ldcg(sp + 4 * ie + 2) : 0.f;
        early.w = 4 * ie + 3 < nsmall ? __ldcg(sp + 4 * ie + 3) : 0.f;
      }
      const int ntrain = nsmall - 4;
      long ai[8]; float wv[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int i = t + j * FS_THREADS;
        ai[j] = i < ntrain ? small_to_arena(p, i) : -1;
        wv[j] = ai[j] >= 0 ? __ldcg(p.w + ai[j]) : 0.f;
      }
      const float metric_prev = (t < 4 && p.metrics) ? __ldcg(p.metrics + t) : 0.f;
      if (t < 80) {
        float4 v[FS_GROUPS];
#pragma unroll
        for (int j = 0; j < FS_GROUPS; ++j)
          v[j] = __ldcg(reinterpret_cast<const float4*>(p.cpart + (long)(FS_PIX + j) * 320) + t);
        float4 a = v[0];
#pragma unroll
        for (int j = 1; j < FS_GROUPS; ++j) { a.x += v[j].x; a.y += v[j].y; a.z += v[j].z; a.w += v[j].w; }
        stage4[t] = a;
      }
      if (ie < n4) stage4[ie] = early;
      if (FS_THREADS + 80 < n4 && t + FS_THREADS + 80 < n4) {     // ncls > 13: a few more groups than threads
        const int i2 = t + FS_THREADS + 80;
        float4 e2;
        e2.x = __ldcg(sp + 4 * i2); e2.y = 4 * i2 + 1 < nsmall ? __ldcg(sp + 4 * i2 + 1) : 0.f;
        e2.z = 4 * i2 + 2 < nsmall ? __ldcg(sp + 4 * i2 + 2) : 0.f; e2.w = 4 * i2 + 3 < nsmall ? __ldcg(sp + 4 * i2 + 3) : 0.f;
        stage4[i2] = e2;
      }
      __syncthreads();
      FS_STAMP(12);
      // `stage` now holds this replica's packed small gradients + [loss*scale, correct, rows, 0]
      if (W > 1) {
        // one-shot exchange of the packed small block (flag-in-data, double-buffered by step parity): push the own
        // block to every peer, gather the peers' blocks into shared memory with all loads in flight at once (not one
        // spin per peer), then sum in rank order
        const int parity = (int)(e & 1u);
        float4* rbuf4 = stage4 + FS_SMALL_F4;                // [(W-1)][n4], spans dh .. end of shared memory (all dead here)
        for (int i = t; i < n4; i += FS_THREADS) {
          const float4 v = stage4[i];
          for (int k = 1; k < W; ++k) {
            const int peer = (R + k) % W;
            ll_store4(p.srecv[peer] + ((long)parity * W + R) * (2 * FS_SMALL_MAX), i, v, e);
          }
        }
        FS_STAMP(13);
        const int tot = (W - 1) * n4;
        const float* rbase = p.srecv[R] + (long)parity * W * (2 * FS_SMALL_MAX);
        for (int base = t; base < tot; base += 4 * FS_THREADS) {
          uint4 ra[4], rb[4];
          const uint4* src[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int idx = base + j * FS_THREADS;
            const int k1 = idx < tot ? idx / n4 : 0, i4 = idx < tot ? idx - k1 * n4 : 0;
            const int peer = (R + k1 + 1) % W;
            src[j] = reinterpret_cast<const uint4*>(rbase + (long)peer * (2 * FS_SMALL_MAX)) + 2 * i4;
            asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(ra[j].x), "=r"(ra[j].y), "=r"(ra[j].z), "=r"(ra[j].w) : "l"(src[j]) : "memory");
            asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(rb[j].x), "=r"(rb[j].y), "=r"(rb[j].z), "=r"(rb[j].w) : "l"(src[j] + 1) : "memory");
          }
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int idx = base + j * FS_THREADS;
            if (idx < tot) {
              float4 v = make_float4(__uint_as_float(ra[j].x), __uint_as_float(ra[j].z), __uint_as_float(rb[j].x), __uint_as_float(rb[j].z));
              if (!(ra[j].y == e && ra[j].w == e && rb[j].y == e && rb[j].w == e)) {
                const int k1 = idx / n4, i4 = idx - k1 * n4, peer = (R + k1 + 1) % W;
                v = ll_load4(p, rbase + (long)peer * (2 * FS_SMALL_MAX), i4, e, 21, peer, pix);     // not there yet: spin
              }
              rbuf4[idx] = v;
            }
          }
        }
        __syncthreads();
        FS_STAMP(14);
        for (int i = t; i < n4; i += FS_THREADS) {
          float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
          for (int r = 0; r < W; ++r) {
            const float4 u = (r == R) ? stage4[i] : rbuf4[((r - R - 1 + W) % W) * n4 + i];
            a.x += u.x; a.y += u.y; a.z += u.z; a.w += u.w;
          }
          stage4[i] = a;
        }
      }
      __syncthreads();
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        if (ai[j] >= 0) {
          const float gv = stage[t + j * FS_THREADS];
          p.g[ai[j]] = gv;
          p.w[ai[j]] = update_one(p.opt, wv[j], gv, p.s1, p.s2, ai[j], h);
        }
      }
      if (t < 4) {
        const float v = stage[ntrain + t];
        p.g[p.n + t] = v;
        p.w[p.n + t] = v;
        if (p.metrics) p.metrics[t] = metric_prev + v;
        if (p.tails) p.tails[(long)s * 4 + t] = v;
      }
      __syncthreads();
      if (t == 0) {
        TfdeHParams* hp = p.hp;                 // advance the device-side step state (values read at step start)
        hp->beta1_pow = b1p; hp->beta2_pow = b2p; hp->step = hp_step + 1u;
        if (s == p.nsteps - 1) {
          p.state[ST_EPOCH0] = epoch0 + (uint32_t)p.nsteps;
          if (p.tails) __threadfence_system();
        }
        st_release_gpu(p.state + ST_DONE, e);     // release: cumulative over the CTA's writes ordered by the barrier above
      }
    }
    FS_STAMP(9);
    // ================================ dense-kernel slice: reduce across replicas, update ================
    if (W > 1) {
      if (owner == R || s + 1 == p.nsteps) {
        // owner: its peers sent their contributions in the middle of P3, they are here by now -> reduce, publish the
        // reduced slice, update. (Everyone on the last step of a launch: drain.)
        finish_slice<TC>(p, e, pix, owner, gacc, ws, gslice, wslice, aslice, h);
      } else {
        pending = true;                       // non-owner: picks the reduced slice up under the next step's convolution
        if (t == 0) s_hprev = h;
      }
    }
    // ================================ wait for the step to be complete ==================================
    FS_STAMP(10);
    if (s + 1 < p.nsteps) {
      if (p.ready != nullptr && !s_in_ready) {   // (uniform per CTA: written at the top of this step) not seen yet: wait
        if (t == 0) wait_flag<true>(p, p.ready + s + 1, p.ready_tag, 12, pix, 4);
        __syncthreads();
      }
      load_patch(s + 1, pv);               // next step's inputs: in flight while this step's tail finishes
      if (t == 0) wait_flag<false>(p, p.state + ST_DONE, e, 11, pix, 3);
      __syncthreads();
    }
    FS_STAMP(11);
    if (*reinterpret_cast<volatile int*>(p.status) != 0) break;    // a wait timed out somewhere: drain
  }
  if (TC) {
    umma::fence_before_sync();
    __syncthreads();
    if (threadIdx.x < 32) { umma::fence_after_sync(); umma::tmem_dealloc(tmem, FS_TMEM_COLS); }
  }
}

struct DeviceInit { bool done = false; int max_blocks = 0; };
DeviceInit g_init[2][64];      // [variant: 0 exact-fp32 FFMA, 1 TF32 tensor core][device]
std::mutex g_init_mu;

}  // namespace

// Launch descriptor (mirrored by ops/fused_step.py::_FusedStepDesc).
struct TfdeFusedStepDesc {
  const void* x; const void* y;
  int x_is_u8; float in_scale; long x_step_bytes;
  int B, nsteps, ncls, opt;
  float* w; float* g; float* s1; float* s2;
  long off_ck, off_cb, off_dk, off_db, off_hk, off_hb, n;
  float loss_scale; int rank; int world; int use_tc;     // use_tc: 1 = TF32 tcgen05 GEMMs, 0 = exact fp32 FFMA
  void* hp; float* metrics; float* tails;
  float* part; float* hbuf; float* cpart; float* small_block;
  void* state; void* status;
  unsigned long long timeout_ns;
  void* rs[FS_MAX_RANKS]; void* gsum[FS_MAX_RANKS]; void* srecv[FS_MAX_RANKS]; void* flags[FS_MAX_RANKS];
  void* prof;
  void* ready; unsigned ready_tag;          // streamed inputs (tfde_fused_exec_streamed); NULL: inputs are resident
};

TFDE_API size_t tfde_fused_step_desc_bytes() { return sizeof(TfdeFusedStepDesc); }
// scratch sizes (floats / words) so that the Python side never hard-codes them
TFDE_API size_t tfde_fused_step_part_floats() { return (size_t)FS_PIX * FS_B * FS_U; }
TFDE_API size_t tfde_fused_step_hbuf_floats() { return (size_t)FS_HBUF_FLOATS; }
TFDE_API size_t tfde_fused_step_cpart_floats() { return (size_t)FS_CPART_FLOATS; }
TFDE_API size_t tfde_fused_step_small_floats() { return (size_t)FS_SMALL_MAX; }
TFDE_API size_t tfde_fused_step_state_words() { return (size_t)FS_STATE_WORDS; }
TFDE_API size_t tfde_fused_step_rs_floats(int world) { return (size_t)world * 2 * FS_ND; }       // LL image: 2x
TFDE_API size_t tfde_fused_step_gsum_floats() { return (size_t)2 * FS_ND; }
TFDE_API size_t tfde_fused_step_srecv_floats(int world) { return (size_t)2 * world * 2 * FS_SMALL_MAX; }
TFDE_API size_t tfde_fused_step_flag_words(int world) { return (size_t)(world + 1) * FS_FLAG_STRIDE + 16; }
TFDE_API int tfde_fused_step_blocks() { return FS_PIX; }
TFDE_API int tfde_fused_step_prof_stamps() { return FS_PROF_STAMPS; }

// One training step x nsteps of Model A on `stream`. Returns 0, a cudaError_t, or TFDE_ERR_INVALID for shapes the
// kernel does not cover (the caller then uses the layer-by-layer path).
TFDE_API int tfde_fused_step(const TfdeFusedStepDesc* d, cudaStream_t stream) {
  if (d->B < 1 || d->B > FS_B || d->nsteps < 1 || d->ncls < 1 || d->ncls > FS_NC) return TFDE_ERR_INVALID;
  if (d->world < 1 || d->world > FS_MAX_RANKS || d->rank < 0 || d->rank >= d->world) return TFDE_ERR_INVALID;
  if (d->opt < 0 || d->opt > 3 || (d->opt >= 1 && !d->s1) || (d->opt == 3 && !d->s2)) return TFDE_ERR_INVALID;
  if ((d->off_dk & 3) || (d->off_db & 3) || ((uintptr_t)d->w & 15) || ((uintptr_t)d->g & 15)) return TFDE_ERR_INVALID;
  int dev = 0;
  cudaError_t err = cudaGetDevice(&dev);
  if (err != cudaSuccess) return (int)err;
  const int tcv = d->use_tc ? 1 : 0;
  const size_t smem = (size_t)(tcv ? T_TOTAL : SM_TOTAL) * sizeof(float);
  {
    std::lock_guard<std::mutex> lk(g_init_mu);
    DeviceInit& di = g_init[tcv][dev & 63];
    if (!di.done) {
      const void* fn = tcv ? (const void*)fused_step_kernel<true> : (const void*)fused_step_kernel<false>;
      err = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (err != cudaSuccess) return (int)err;
      // two CTAs per SM need the large shared-memory carve-out (the default is sized for ONE block of this kernel)
      err = cudaFuncSetAttribute(fn, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared);
      if (err != cudaSuccess) return (int)err;
      int per_sm = 0, sms = 0;
      err = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, FS_THREADS, smem);
      if (err != cudaSuccess) return (int)err;
      err = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
      if (err != cudaSuccess) return (int)err;
      di.max_blocks = per_sm * sms;
      di.done = true;
    }
    if (di.max_blocks < FS_PIX && !(tcv && getenv("TFDE_FUSED_TC_FORCE"))) {   // the grid could not be co-resident on this GPU
      cudaFuncAttributes fa{};
      const void* fn2 = tcv ? (const void*)fused_step_kernel<true> : (const void*)fused_step_kernel<false>;
      cudaFuncGetAttributes(&fa, fn2);
      size_t avail2 = 0;
      cudaOccupancyAvailableDynamicSMemPerBlock(&avail2, fn2, 2, FS_THREADS);
      int smem_sm = 0;
      cudaDeviceGetAttribute(&smem_sm, cudaDevAttrMaxSharedMemoryPerMultiprocessor, dev);
      fprintf(stderr, "tfde_fused_step: only %d CTAs of %zu bytes shared memory can be co-resident (need %d); regs %d, "
              "static smem %zu, max dyn %d, carveout %d, dyn smem available at 2 CTAs/SM %zu, smem per SM %d\n",
              di.max_blocks, smem, FS_PIX, fa.numRegs, fa.sharedSizeBytes, fa.maxDynamicSharedSizeBytes,
              fa.preferredShmemCarveout, avail2, smem_sm);
      return TFDE_ERR_INVALID;
    }
  }
  FsParams p{};
  p.x = d->x; p.y = (const long long*)d->y; p.x_is_u8 = d->x_is_u8; p.in_scale = d->in_scale;
  p.x_step_bytes = d->x_step_bytes; p.B = d->B; p.nsteps = d->nsteps; p.ncls = d->ncls; p.opt = d->opt;
  p.w = d->w; p.g = d->g; p.s1 = d->s1; p.s2 = d->s2;
  p.off_ck = d->off_ck; p.off_cb = d->off_cb; p.off_dk = d->off_dk; p.off_db = d->off_db;
  p.off_hk = d->off_hk; p.off_hb = d->off_hb; p.n = d->n;
  p.loss_scale = d->loss_scale; p.hp = (TfdeHParams*)d->hp; p.metrics = d->metrics; p.tails = d->tails;
  p.part = d->part; p.hbuf = d->hbuf; p.cpart = d->cpart; p.small = d->small_block;
  p.state = (uint32_t*)d->state; p.status = (int*)d->status;
  p.timeout_ns = d->timeout_ns ? d->timeout_ns : 5000000000ULL;
  p.rank = d->rank; p.world = d->world;
  for (int i = 0; i < FS_MAX_RANKS; ++i) {
    p.rs[i] = (float*)d->rs[i]; p.gsum[i] = (float*)d->gsum[i];
    p.srecv[i] = (float*)d->srecv[i]; p.flags[i] = (uint32_t*)d->flags[i];
  }
  p.prof = (unsigned long long*)d->prof;
  p.ready = (const uint32_t*)d->ready; p.ready_tag = d->ready_tag;
  if (d->world > 1)
    for (int i = 0; i < d->world; ++i)
      if (!p.rs[i] || !p.gsum[i] || !p.srecv[i] || !p.flags[i]) return TFDE_ERR_INVALID;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(FS_PIX); cfg.blockDim = dim3(FS_THREADS); cfg.dynamicSmemBytes = smem; cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeCooperative;
  attr[0].val.cooperative = 1;
  cfg.attrs = attr; cfg.numAttrs = 1;
  if (tcv && g_init[1][dev & 63].max_blocks < FS_PIX) {
    // TFDE_FUSED_TC_FORCE (bring-up experiment): the runtime's occupancy calculator grants kernels that allocate tensor
    // memory ONE CTA per SM, so a cooperative launch of 169 CTAs is refused; launch plainly and rely on the bounded
    // waits to report (status word) rather than hang if the CTAs turn out not to be co-resident
    cfg.numAttrs = 0;
  }
  err = tcv ? cudaLaunchKernelEx(&cfg, fused_step_kernel<true>, p) : cudaLaunchKernelEx(&cfg, fused_step_kernel<false>, p);
  if (err != cudaSuccess) return (int)err;
  TFDE_RETURN_LAUNCH_STATUS();
}

// ---------------------------------------------------------------------------------------------
// Host launcher for the end-to-end path (Keras `steps_per_execution`): ONE call stages the inputs of `nsteps`
// steps and launches them. Per step it issues that step's own host->device copies (images, labels) from the
// caller's pinned memory on `copy_stream`, so the copies of execution i+1 run under the kernel of execution i;
// the kernel is ordered after the copies by an event and is followed by `ev_done` (the host waits on it before
// it reuses this staging slot or reads the losses the kernel wrote into mapped pinned memory).
// ---------------------------------------------------------------------------------------------
TFDE_API int tfde_event_create(void** out) {
  cudaEvent_t e;
  cudaError_t err = cudaEventCreateWithFlags(&e, cudaEventDisableTiming);
  *out = (void*)e;
  return (int)err;
}
TFDE_API int tfde_event_destroy(void* e) { return (int)cudaEventDestroy((cudaEvent_t)e); }
TFDE_API int tfde_event_synchronize(void* e) { return (int)cudaEventSynchronize((cudaEvent_t)e); }
TFDE_API int tfde_event_query(void* e) { return (int)cudaEventQuery((cudaEvent_t)e); }

TFDE_API int tfde_fused_exec(const TfdeFusedStepDesc* d, const void* host_x, const void* host_y,
                             size_t x_bytes_per_step, size_t y_bytes_per_step, cudaStream_t copy_stream,
                             void* ev_copied, void* ev_done, cudaStream_t stream) {
  cudaError_t err;
  char* dx = (char*)d->x; char* dy = (char*)d->y;
  for (int s = 0; s < d->nsteps; ++s) {
    err = cudaMemcpyAsync(dx + (size_t)s * x_bytes_per_step, (const char*)host_x + (size_t)s * x_bytes_per_step,
                          x_bytes_per_step, cudaMemcpyHostToDevice, copy_stream);
    if (err != cudaSuccess) return (int)err;
    err = cudaMemcpyAsync(dy + (size_t)s * y_bytes_per_step, (const char*)host_y + (size_t)s * y_bytes_per_step,
                          y_bytes_per_step, cudaMemcpyHostToDevice, copy_stream);
    if (err != cudaSuccess) return (int)err;
  }
  err = cudaEventRecord((cudaEvent_t)ev_copied, copy_stream);
  if (err != cudaSuccess) return (int)err;
  err = cudaStreamWaitEvent(stream, (cudaEvent_t)ev_copied, 0);
  if (err != cudaSuccess) return (int)err;
  int rc = tfde_fused_step(d, stream);
  if (rc != 0) return rc;
  return (int)cudaEventRecord((cudaEvent_t)ev_done, stream);
}

// Streamed form of tfde_fused_exec: the kernel is launched FIRST-COME -- it does not wait for the copies. Per step the
// copy stream carries that step's images, its labels and then a 4-byte tag into d->ready[s]; the kernel (d->ready,
// d->ready_tag) consumes step s once ready[s] >= tag (bounded wait, status code 12). One launch then covers any
// number of steps with every copy but the first step's hidden under compute, instead of one launch per staging chunk.
// host_tags: pinned uint32[nsteps], all equal to d->ready_tag (tags grow with every use of the same ready array).
// The copies are plain cudaMemcpyAsync calls (copy engine); nothing on the copy stream depends on the kernel.
TFDE_API int tfde_fused_exec_streamed(const TfdeFusedStepDesc* d, const void* host_x, const void* host_y,
                                      const void* host_tags, size_t x_bytes_per_step, size_t y_bytes_per_step,
                                      cudaStream_t copy_stream, void* ev_copied, void* ev_done, cudaStream_t stream) {
  if (d->ready == nullptr || host_tags == nullptr) return TFDE_ERR_INVALID;
  cudaError_t err;
  char* dx = (char*)d->x; char* dy = (char*)d->y;
  auto copy_step = [&](int s) -> cudaError_t {
    cudaError_t e = cudaMemcpyAsync(dx + (size_t)s * x_bytes_per_step, (const char*)host_x + (size_t)s * x_bytes_per_step,
                                    x_bytes_per_step, cudaMemcpyHostToDevice, copy_stream);
    if (e != cudaSuccess) return e;
    e = cudaMemcpyAsync(dy + (size_t)s * y_bytes_per_step, (const char*)host_y + (size_t)s * y_bytes_per_step,
                        y_bytes_per_step, cudaMemcpyHostToDevice, copy_stream);
    if (e != cudaSuccess) return e;
    return cudaMemcpyAsync((char*)d->ready + (size_t)s * 4, (const char*)host_tags + (size_t)s * 4, 4,
                           cudaMemcpyHostToDevice, copy_stream);
  };
  // the first step's batch is queued before the launch, the rest right behind it
  err = copy_step(0);
  if (err != cudaSuccess) return (int)err;
  int rc = tfde_fused_step(d, stream);
  if (rc != 0) return rc;
  for (int s = 1; s < d->nsteps; ++s) {
    err = copy_step(s);
    if (err != cudaSuccess) return (int)err;
  }
  err = cudaEventRecord((cudaEvent_t)ev_copied, copy_stream);
  if (err != cudaSuccess) return (int)err;
  return (int)cudaEventRecord((cudaEvent_t)ev_done, stream);
}

// Diagnostics: resident CTAs per SM the runtime grants either variant at `dyn_smem` bytes of dynamic shared memory.
TFDE_API int tfde_fused_step_occupancy(int use_tc, size_t dyn_smem) {
  const void* fn = use_tc ? (const void*)fused_step_kernel<true> : (const void*)fused_step_kernel<false>;
  cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(dyn_smem > 49152 ? dyn_smem : 49152));
  cudaFuncSetAttribute(fn, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared);
  int per_sm = -1;
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, FS_THREADS, dyn_smem);
  return per_sm;
}
