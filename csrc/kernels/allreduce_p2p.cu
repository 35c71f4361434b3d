// G12: gradient all-reduce over NVLink 5 / NVSwitch peer memory, FUSED with the 1/N scale, the optional
// bf16 wire cast and the SGD / momentum / Adam update -- the per-step collective that TensorFlow runs
// through NCCL cross_device_ops under MultiWorkerMirroredStrategy / MirroredStrategy
// (distributed_with_keras.py:16,63; mnist_keras_distributed.py:243). No NCCL call and no separate
// optimizer kernel on this path (G11 is folded in). Also serves G13 (initial broadcast) and G14 (metric
// aggregation: the 4-float tail of the gradient arena is reduced with the gradients).
//
// Design (B200-first): every GPU reaches every peer at full bandwidth through NVSwitch, so there is
// no ring. All NVLink traffic is *posted writes* (stores into the peer's receive slots); every load
// is local. Two algorithms:
//
//   one-shot  (small arenas / few ranks): each rank pushes its whole gradient into slot[rank] of every
//             peer, one flag barrier, every rank sums the N slots in rank order (bit-identical
//             replicas) and applies the optimizer to its full local weights.  bytes out: (N-1)*S.
//   two-shot  (large arenas): rank r owns slice r. push slice j to rank j (reduce-scatter), barrier,
//             owner reduces + applies the optimizer on its slice only (optimizer state sharded by
//             owner = ZeRO-1 for free), pushes the updated weights to every peer (all-gather),
//             barrier.  bytes out: 2*(N-1)/N*S.
//
// Synchronisation: block b of every rank handles the same sub-range, so a per-(rank, block) flag is
// enough -- no grid-wide sync, no cooperative launch. Flags carry a monotonically increasing epoch
// kept in device memory, which makes the kernel replayable inside a CUDA graph. Receive slots are
// double-buffered by epoch parity, which removes the exit barrier of the one-shot path: a slot
// written at epoch e is rewritten at e+2, and a rank can only reach e+2 after every peer passed the
// barrier of e+1, i.e. after it finished reading e. Waits are bounded: on timeout the kernel records
// (code, epoch, peer, block) in a status word and stops spinning instead of hanging the GPU.
//
// All blocks of one launch must be co-resident (grid <= 148 * occupancy); the host wrapper enforces it.
#include "common.cuh"
#include <cuda_bf16.h>
#include <cstring>

namespace {

constexpr int MAX_RANKS = 8;
constexpr int MAX_BLOCKS = 256;
constexpr int AR_THREADS = 512;

struct ARParams {
  int rank, world;
  int mode;            // 0 one-shot, 1 two-shot
  int opt;             // -1 reduce only, 0 SGD, 1 momentum, 2 nesterov, 3 Adam
  int wire_bf16;       // push gradients as bf16
  long n;              // trainable floats
  long n_total;        // n + extras (reduced, not updated), multiple of 4
  long slot_floats;    // floats per receive slot (one-shot: n_total, two-shot: slice size)
  const float* g;      // local gradient arena [n_total]
  float* w;            // local parameter arena [n_total] (tail receives the reduced extras)
  float* s1; float* s2;            // optimizer slots (two-shot: slice-sized, one-shot: full)
  __nv_bfloat16* w_bf16;           // optional local bf16 shadow of the weights [n_total]
  float* recv[MAX_RANKS];          // peer-mapped receive buffers: [2 parity][world][slot_floats]
  float* wpeer[MAX_RANKS];         // peer-mapped parameter arenas (two-shot all-gather targets)
  __nv_bfloat16* wbf_peer[MAX_RANKS];
  uint32_t* flags[MAX_RANKS];      // peer-mapped signal pads: [2 phase][world][MAX_BLOCKS]
  uint32_t* epoch;                 // local [MAX_BLOCKS]
  float* metrics;                  // local running sums of the extras, or nullptr
  int n_extra;
  TfdeHParams* hp;
  unsigned* done_counter;          // local ticket for the hparams advance
  int* status;                     // local [4]: code, epoch, peer, block
  unsigned long long timeout_ns;
  long pp_lo, pp_hi;               // arena range a producer kernel may have pushed already (GEMM epilogue)
  const uint32_t* pp_done;         // == this epoch when it did; nullptr when the feature is off
  // NVLS (mode 2): multicast mappings of the gradient arena, the parameter arena and a signal pad
  // [2 phase][MAX_BLOCKS]; nv_flags is this rank's ordinary mapping of the same signal pad.
  const float* mc_g; float* mc_w; uint32_t* mc_flags; const uint32_t* nv_flags;
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// Signal "my block `b` finished its pushes of phase `phase`" to every peer, then wait for theirs.
__device__ __forceinline__ void flag_barrier(const ARParams& p, int phase, int b, uint32_t e) {
  __syncthreads();  // all pushes of this block issued
  const int t = threadIdx.x;
  if (t < p.world && t != p.rank) {
    // st.release.sys orders every write that happened-before it (the block's pushes, via the
    // bar.sync above) before the flag becomes visible; no separate membar.sys round trip needed.
    st_release_sys(p.flags[t] + ((size_t)phase * p.world + p.rank) * MAX_BLOCKS + b, e);
    const uint32_t* mine = p.flags[p.rank] + ((size_t)phase * p.world + t) * MAX_BLOCKS + b;
    const unsigned long long t0 = globaltimer_ns();
    unsigned spins = 0;
    while ((int)(ld_acquire_sys(mine) - e) < 0) {
      if ((++spins & 0x3ff) == 0 && globaltimer_ns() - t0 > p.timeout_ns) {
        if (atomicCAS(p.status, 0, 1) == 0) { p.status[1] = (int)e; p.status[2] = t; p.status[3] = b; }
        break;
      }
    }
  }
  __syncthreads();
}

__device__ __forceinline__ float bf16_round(float v) { return __bfloat162float(__float2bfloat16(v)); }

struct Hyper { float lr, mu, b1, b2, eps, gs, lr_t; };

__device__ __forceinline__ float update_one(int opt, float w, float g, float* s1, float* s2, long si,
                                            const Hyper& h) {
  if (opt < 0) return g;
  g *= h.gs;
  if (opt == 0) return w - h.lr * g;
  if (opt == 1 || opt == 2) {
    const float v = h.mu * s1[si] - h.lr * g;
    s1[si] = v;
    return w + (opt == 2 ? (h.mu * v - h.lr * g) : v);
  }
  const float m = h.b1 * s1[si] + (1.f - h.b1) * g;
  const float v = h.b2 * s2[si] + (1.f - h.b2) * g * g;
  s1[si] = m; s2[si] = v;
  return w - h.lr_t * m / (sqrtf(v) + h.eps);
}

// ---- vectorised data movement -------------------------------------------------------------------
// Everything moves as 16-byte groups of 4 floats; loops keep PUSH_U independent groups in flight per
// thread (posted NVLink stores need memory-level parallelism, not occupancy).
constexpr int PUSH_U = 4;

__device__ __forceinline__ uint2 pack_bf16x4(const float4 v) {
  __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y), b = __floats2bfloat162_rn(v.z, v.w);
  uint2 o; o.x = *reinterpret_cast<uint32_t*>(&a); o.y = *reinterpret_cast<uint32_t*>(&b);
  return o;
}
__device__ __forceinline__ float4 unpack_bf16x4(const uint2 r) {
  float4 v;
  v.x = __uint_as_float(r.x << 16); v.y = __uint_as_float(r.x & 0xffff0000u);
  v.z = __uint_as_float(r.y << 16); v.w = __uint_as_float(r.y & 0xffff0000u);
  return v;
}
__device__ __forceinline__ float4 round_bf16x4(const float4 v) { return unpack_bf16x4(pack_bf16x4(v)); }

// Store group `i4` (index in 4-float groups relative to the slot start) of a slot.
__device__ __forceinline__ void slot_store4(float* slot, long i4, const float4 v, int wire_bf16) {
  if (wire_bf16) __stcg(reinterpret_cast<uint2*>(slot) + i4, pack_bf16x4(v));
  else           __stcg(reinterpret_cast<float4*>(slot) + i4, v);
}
__device__ __forceinline__ float4 slot_load4(const float* slot, long i4, int wire_bf16) {
  if (wire_bf16) return unpack_bf16x4(__ldcg(reinterpret_cast<const uint2*>(slot) + i4));
  return __ldcg(reinterpret_cast<const float4*>(slot) + i4);
}

struct Vec4 { float v[4]; };
__device__ __forceinline__ Vec4 to_vec(const float4 f) { Vec4 o; o.v[0] = f.x; o.v[1] = f.y; o.v[2] = f.z; o.v[3] = f.w; return o; }
__device__ __forceinline__ float4 to_f4(const Vec4& o) { return make_float4(o.v[0], o.v[1], o.v[2], o.v[3]); }

// Sum the W contributions of group (global float index gi, slot-relative group index si4) in RANK ORDER
// (identical order on every rank => bit-identical replicas), own contribution read from g.
__device__ __forceinline__ float4 reduce_group(const ARParams& p, const float* base, long slot, long gi, long si4) {
  float4 v[MAX_RANKS];
#pragma unroll
  for (int r = 0; r < MAX_RANKS; ++r) {
    if (r < p.world) {
      if (r == p.rank) {
        v[r] = __ldg(reinterpret_cast<const float4*>(p.g + gi));
        if (p.wire_bf16) v[r] = round_bf16x4(v[r]);
      } else {
        v[r] = slot_load4(base + (size_t)r * slot, si4, p.wire_bf16);
      }
    }
  }
  float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
  for (int r = 0; r < MAX_RANKS; ++r)
    if (r < p.world) { s.x += v[r].x; s.y += v[r].y; s.z += v[r].z; s.w += v[r].w; }
  return s;
}

// Optimizer on one group. gi = global float index of the group, si = index into the optimizer slots.
__device__ __forceinline__ float4 update_group(const ARParams& p, const Hyper& h, long gi, long si, const float4 sum,
                                               bool accumulate_metrics) {
  Vec4 s = to_vec(sum), nw;
  if (gi + 3 < p.n && p.opt >= 0) {
    const Vec4 w = to_vec(*reinterpret_cast<const float4*>(p.w + gi));
    if (p.opt == 0) {
#pragma unroll
      for (int c = 0; c < 4; ++c) nw.v[c] = w.v[c] - h.lr * (s.v[c] * h.gs);
    } else if (p.opt == 1 || p.opt == 2) {
      Vec4 m = to_vec(*reinterpret_cast<const float4*>(p.s1 + si));
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const float g = s.v[c] * h.gs;
        const float v = h.mu * m.v[c] - h.lr * g;
        m.v[c] = v;
        nw.v[c] = w.v[c] + (p.opt == 2 ? (h.mu * v - h.lr * g) : v);
      }
      *reinterpret_cast<float4*>(p.s1 + si) = to_f4(m);
    } else {
      Vec4 m = to_vec(*reinterpret_cast<const float4*>(p.s1 + si));
      Vec4 q = to_vec(*reinterpret_cast<const float4*>(p.s2 + si));
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const float g = s.v[c] * h.gs;
        m.v[c] = h.b1 * m.v[c] + (1.f - h.b1) * g;
        q.v[c] = h.b2 * q.v[c] + (1.f - h.b2) * g * g;
        nw.v[c] = w.v[c] - h.lr_t * m.v[c] / (sqrtf(q.v[c]) + h.eps);
      }
      *reinterpret_cast<float4*>(p.s1 + si) = to_f4(m);
      *reinterpret_cast<float4*>(p.s2 + si) = to_f4(q);
    }
  } else {
    // boundary group (parameters end inside it) or reduce-only: per component
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const long i = gi + c;
      if (i < p.n) nw.v[c] = update_one(p.opt, p.w[i], s.v[c], p.s1, p.s2, si + c, h);
      else {
        nw.v[c] = s.v[c];
        if (accumulate_metrics && p.metrics && i - p.n < p.n_extra) p.metrics[i - p.n] += s.v[c];
      }
    }
  }
  return to_f4(nw);
}

__device__ __forceinline__ void store_bf16_shadow(__nv_bfloat16* dst, long gi, const float4 nw, long n) {
  if (gi + 3 < n) *reinterpret_cast<uint2*>(dst + gi) = pack_bf16x4(nw);
  else {
    const Vec4 o = to_vec(nw);
    for (int c = 0; c < 4; ++c) if (gi + c < n) dst[gi + c] = __float2bfloat16(o.v[c]);
  }
}

__global__ void __launch_bounds__(AR_THREADS)
allreduce_fused_kernel(const ARParams p) {
  TFDE_PDL_PROLOGUE();
  const int b = blockIdx.x, nb = gridDim.x, t = threadIdx.x;
  __shared__ uint32_t s_e;
  if (t == 0) s_e = p.epoch[b] + 1u;
  __syncthreads();
  const uint32_t e = s_e;
  const int parity = (int)(e & 1u);
  const int W = p.world, R = p.rank;

  Hyper h;
  if (p.opt >= 0) {
    h.lr = p.hp->lr; h.mu = p.hp->momentum; h.b1 = p.hp->beta1; h.b2 = p.hp->beta2; h.eps = p.hp->eps;
    h.gs = p.hp->grad_scale;
    const float b1p = p.hp->beta1_pow * h.b1, b2p = p.hp->beta2_pow * h.b2;
    h.lr_t = (p.opt == 3) ? h.lr * sqrtf(1.f - b2p) / (1.f - b1p) : h.lr;
  } else {
    h.lr = h.mu = h.b1 = h.b2 = h.eps = h.lr_t = 0.f; h.gs = 1.f;
  }

  const long slot = p.slot_floats;
  const float* base = p.recv[R] + (size_t)parity * W * slot;   // my receive slots of this parity
  // [pp_lo, pp_hi) was already pushed by the epilogue of the GEMM that produced it (if it ran this step)
  const bool prepushed = p.pp_done != nullptr && *p.pp_done == e;
  const long pp_lo = prepushed ? p.pp_lo : 0, pp_hi = prepushed ? p.pp_hi : 0;
  if (p.mode == 0) {
    // ------------------------------- one-shot -------------------------------------------
    long chunk = (p.n_total + nb - 1) / nb; chunk = (chunk + 3) & ~3L;
    const long lo = min((long)b * chunk, p.n_total), hi = min(lo + chunk, p.n_total);
    const long lo4 = lo >> 2, cnt4 = (hi - lo) >> 2;
    const float4* g4 = reinterpret_cast<const float4*>(p.g) + lo4;
    // push: read each group once, store it into slot[R] of every peer
    for (long i0 = t; i0 < cnt4; i0 += (long)PUSH_U * blockDim.x) {
      float4 v[PUSH_U];
#pragma unroll
      bool need[PUSH_U];
#pragma unroll
      for (int u = 0; u < PUSH_U; ++u) {
        const long i = i0 + (long)u * blockDim.x;
        const long gi = lo + (i << 2);
        need[u] = i < cnt4 && !(gi >= pp_lo && gi < pp_hi);
        if (need[u]) v[u] = __ldg(g4 + i);
      }
      for (int k = 1; k < W; ++k) {
        const int peer = (R + k) % W;
        float* dst = p.recv[peer] + ((size_t)parity * W + R) * slot;
#pragma unroll
        for (int u = 0; u < PUSH_U; ++u) {
          const long i = i0 + (long)u * blockDim.x;
          if (need[u]) slot_store4(dst, lo4 + i, v[u], p.wire_bf16);
        }
      }
    }
    flag_barrier(p, 0, b, e);
    for (long i = t; i < cnt4; i += blockDim.x) {
      const long gi = lo + (i << 2);
      const float4 s = reduce_group(p, base, slot, gi, lo4 + i);
      const float4 nw = update_group(p, h, gi, gi, s, true);
      *reinterpret_cast<float4*>(p.w + gi) = nw;
      if (p.w_bf16) store_bf16_shadow(p.w_bf16, gi, nw, p.n);
    }
  } else {
    // ------------------------------- two-shot -------------------------------------------
    // slice j = [j*slot, (j+1)*slot) ; this block handles sub-range [lo, hi) of every slice.
    long chunk = (slot + nb - 1) / nb; chunk = (chunk + 3) & ~3L;
    const long lo = min((long)b * chunk, slot), hi = min(lo + chunk, slot);
    const long lo4 = lo >> 2;
    for (int k = 1; k < W; ++k) {          // reduce-scatter by push: slice `peer` of my gradient -> peer
      const int peer = (R + k) % W;
      const long gbeg = (long)peer * slot + lo;
      const long cnt4 = (min((long)peer * slot + hi, p.n_total) - gbeg) >> 2;
      if (cnt4 <= 0) continue;
      float* dst = p.recv[peer] + ((size_t)parity * W + R) * slot;
      const float4* g4 = reinterpret_cast<const float4*>(p.g + gbeg);
      for (long i0 = t; i0 < cnt4; i0 += (long)PUSH_U * blockDim.x) {
        float4 v[PUSH_U];
#pragma unroll
        bool need[PUSH_U];
#pragma unroll
        for (int u = 0; u < PUSH_U; ++u) {
          const long i = i0 + (long)u * blockDim.x;
          const long gi = gbeg + (i << 2);
          need[u] = i < cnt4 && !(gi >= pp_lo && gi < pp_hi);
          if (need[u]) v[u] = __ldg(g4 + i);
        }
#pragma unroll
        for (int u = 0; u < PUSH_U; ++u) {
          const long i = i0 + (long)u * blockDim.x;
          if (need[u]) slot_store4(dst, lo4 + i, v[u], p.wire_bf16);
        }
      }
    }
    flag_barrier(p, 0, b, e);
    const long own0 = (long)R * slot;
    const long own_cnt4 = (min(own0 + hi, p.n_total) - (own0 + lo)) >> 2;
    for (long i = t; i < own_cnt4; i += blockDim.x) {
      const long li = lo + (i << 2);
      const long gi = own0 + li;
      const float4 s = reduce_group(p, base, slot, gi, lo4 + i);
      const float4 nw = update_group(p, h, gi, li, s, false);
      // all-gather by push: the new values go into every replica (own included)
      for (int k = 0; k < W; ++k) {
        const int peer = (R + k) % W;
        __stcg(reinterpret_cast<float4*>(p.wpeer[peer] + gi), nw);
        if (p.w_bf16) store_bf16_shadow(p.wbf_peer[peer], gi, nw, p.n);
      }
    }
    flag_barrier(p, 1, b, e);
    // every slice's sub-range [lo,hi) has now landed locally: accumulate the metrics tail once.
    if (p.metrics) {
      for (int j = t; j < p.n_extra; j += blockDim.x) {
        const long i = p.n + j;
        const long li = i % slot;
        if (li >= lo && li < hi) p.metrics[j] += __ldcg(p.w + i);
      }
    }
  }

  __syncthreads();
  if (t == 0) {
    p.epoch[b] = e;
    if (atomicAdd(p.done_counter, 1u) == (unsigned)nb - 1u) {
      // last block of this launch: every block has consumed the old hyper-parameters by now
      if (p.opt >= 0) {
        p.hp->beta1_pow *= p.hp->beta1;
        p.hp->beta2_pow *= p.hp->beta2;
        p.hp->step += 1u;
      }
      *p.done_counter = 0u;
    }
  }
}

// ---- NVLS variant: the NVSwitch does the sum and the fan-out ------------------------------------------
// Same ownership as two-shot (rank r owns slice r, optimizer state sharded by owner), but no receive
// slots and no per-peer loops: the owner reads its slice through the MULTICAST mapping of the gradient
// arena with multimem.ld_reduce (the switch fetches the W replicas and returns their sum), applies the
// optimizer, and writes the new weights through the multicast mapping of the parameter arena with
// multimem.st (the switch replicates the store into every GPU, own included). Per GPU and direction that
// is ~S*(1+1/W) bytes on NVLink instead of 2*S*(W-1)/W, and two signal rounds (one multimem.red each).
__device__ __forceinline__ float4 mc_ld_reduce4(const float* mc) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
  return v;
}
__device__ __forceinline__ void mc_st4(float* mc, const float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
               ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
// Arrive on counter `b` of `phase` in EVERY rank's signal pad with one instruction, then wait until all W
// ranks have arrived on ours. Counters only ever grow: after epoch e each holds e*W.
__device__ __forceinline__ void nvls_barrier(const ARParams& p, int phase, int b, uint32_t e) {
  __syncthreads();
  if (threadIdx.x == 0) {
    // phase 0 publishes nothing this kernel wrote (the gradients come from earlier kernels of the stream,
    // already visible at the kernel boundary), so it skips the release fence (a MEMBAR.SYS round trip).
    if (phase == 0)
      asm volatile("multimem.red.relaxed.sys.global.add.u32 [%0], %1;"
                   ::"l"(p.mc_flags + phase * MAX_BLOCKS + b), "r"(1u) : "memory");
    else
      asm volatile("multimem.red.release.sys.global.add.u32 [%0], %1;"
                   ::"l"(p.mc_flags + phase * MAX_BLOCKS + b), "r"(1u) : "memory");
    const uint32_t* mine = p.nv_flags + phase * MAX_BLOCKS + b;
    const uint32_t target = e * (uint32_t)p.world;
    const unsigned long long t0 = globaltimer_ns();
    unsigned spins = 0;
    while ((int)(ld_acquire_sys(mine) - target) < 0) {
      if ((++spins & 0x3ff) == 0 && globaltimer_ns() - t0 > p.timeout_ns) {
        if (atomicCAS(p.status, 0, 2) == 0) { p.status[1] = (int)e; p.status[2] = phase; p.status[3] = b; }
        break;
      }
    }
  }
  __syncthreads();
}

__global__ void __launch_bounds__(AR_THREADS)
allreduce_nvls_kernel(const ARParams p) {
  TFDE_PDL_PROLOGUE();
  const int b = blockIdx.x, nb = gridDim.x, t = threadIdx.x;
  __shared__ uint32_t s_e;
  if (t == 0) s_e = p.epoch[b] + 1u;
  __syncthreads();
  const uint32_t e = s_e;
  const int R = p.rank;

  Hyper h;
  if (p.opt >= 0) {
    h.lr = p.hp->lr; h.mu = p.hp->momentum; h.b1 = p.hp->beta1; h.b2 = p.hp->beta2; h.eps = p.hp->eps;
    h.gs = p.hp->grad_scale;
    const float b1p = p.hp->beta1_pow * h.b1, b2p = p.hp->beta2_pow * h.b2;
    h.lr_t = (p.opt == 3) ? h.lr * sqrtf(1.f - b2p) / (1.f - b1p) : h.lr;
  } else {
    h.lr = h.mu = h.b1 = h.b2 = h.eps = h.lr_t = 0.f; h.gs = 1.f;
  }

  const long slot = p.slot_floats;
  long chunk = (slot + nb - 1) / nb; chunk = (chunk + 3) & ~3L;
  const long lo = min((long)b * chunk, slot), hi = min(lo + chunk, slot);
  const long own0 = (long)R * slot;
  const long own_cnt4 = max(0L, (min(own0 + hi, p.n_total) - (own0 + lo)) >> 2);

  // phase 0: block b of every rank has entered => every replica's gradients (written by kernels that
  // precede this one in each rank's stream) are final and may be read through the switch.
  nvls_barrier(p, 0, b, e);
  for (long i0 = t; i0 < own_cnt4; i0 += (long)PUSH_U * blockDim.x) {
    float4 s[PUSH_U];
#pragma unroll
    for (int u = 0; u < PUSH_U; ++u) {
      const long i = i0 + (long)u * blockDim.x;
      if (i < own_cnt4) s[u] = mc_ld_reduce4(p.mc_g + own0 + lo + (i << 2));
    }
#pragma unroll
    for (int u = 0; u < PUSH_U; ++u) {
      const long i = i0 + (long)u * blockDim.x;
      if (i < own_cnt4) {
        const long li = lo + (i << 2);
        const float4 nw = update_group(p, h, own0 + li, li, s[u], false);
        mc_st4(p.mc_w + own0 + li, nw);
      }
    }
  }
  // phase 1: every rank's block b has issued (and released) its multicast stores => sub-range [lo,hi) of
  // every slice has landed here; the kernel as a whole ends only when all blocks have seen that, which
  // also means no peer is still reading our gradients when the next step starts overwriting them.
  nvls_barrier(p, 1, b, e);
  if (p.metrics) {
    for (int j = t; j < p.n_extra; j += blockDim.x) {
      const long i = p.n + j;
      const long li = i % slot;
      if (li >= lo && li < hi) p.metrics[j] += __ldcg(p.w + i);
    }
  }
  __syncthreads();
  if (t == 0) {
    p.epoch[b] = e;
    if (atomicAdd(p.done_counter, 1u) == (unsigned)nb - 1u) {
      if (p.opt >= 0) {
        p.hp->beta1_pow *= p.hp->beta1;
        p.hp->beta2_pow *= p.hp->beta2;
        p.hp->step += 1u;
      }
      *p.done_counter = 0u;
    }
  }
}

// Pull copy used for the initial broadcast (G13): dst(local) <- src(peer-mapped), float4 granularity.
__global__ void p2p_copy_kernel(float* __restrict__ dst, const float* __restrict__ src, long n) {
  const long n4 = n / 4;
  const float4* s4 = reinterpret_cast<const float4*>(src);
  float4* d4 = reinterpret_cast<float4*>(dst);
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long)gridDim.x * blockDim.x)
    d4[i] = s4[i];
  for (long i = n4 * 4 + (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x)
    dst[i] = src[i];
}

}  // namespace

// ---------------------------------------------------------------------------------------------
// Host side: symmetric allocation + IPC plumbing (setup only) and the launch wrapper.
// ---------------------------------------------------------------------------------------------
TFDE_API int tfde_set_device(int dev) { return (int)cudaSetDevice(dev); }

TFDE_API int tfde_symm_alloc(size_t bytes, void** out) {
  cudaError_t e = cudaMalloc(out, bytes);
  if (e != cudaSuccess) return (int)e;
  return (int)cudaMemset(*out, 0, bytes);
}
TFDE_API int tfde_symm_free(void* p) { return (int)cudaFree(p); }
TFDE_API int tfde_ipc_get_handle(void* p, void* handle64) {
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, p);
  if (e != cudaSuccess) return (int)e;
  memcpy(handle64, &h, sizeof(h));
  return 0;
}
TFDE_API int tfde_ipc_open_handle(const void* handle64, void** out) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, sizeof(h));
  return (int)cudaIpcOpenMemHandle(out, h, cudaIpcMemLazyEnablePeerAccess);
}
TFDE_API int tfde_ipc_close_handle(void* p) { return (int)cudaIpcCloseMemHandle(p); }
TFDE_API int tfde_ipc_handle_bytes() { return (int)sizeof(cudaIpcMemHandle_t); }
TFDE_API int tfde_ar_max_ranks() { return MAX_RANKS; }
TFDE_API int tfde_ar_max_blocks() { return MAX_BLOCKS; }
TFDE_API size_t tfde_ar_nvls_flag_bytes() { return (size_t)2 * MAX_BLOCKS * sizeof(uint32_t); }
TFDE_API size_t tfde_ar_flag_bytes(int world) { return (size_t)2 * world * MAX_BLOCKS * sizeof(uint32_t); }

// Descriptor filled by the Python side once (pointers are stable), then passed by value per launch.
struct TfdeARDesc {
  int rank, world, mode, opt, wire_bf16, n_extra, blocks, reserved;
  long n, n_total, slot_floats;
  const float* g; float* w; float* s1; float* s2; void* w_bf16;
  void* recv[MAX_RANKS]; void* wpeer[MAX_RANKS]; void* wbf_peer[MAX_RANKS]; void* flags[MAX_RANKS];
  void* epoch; float* metrics; void* hp; void* done_counter; void* status;
  unsigned long long timeout_ns;
  long pp_lo, pp_hi; const void* pp_done;     // pre-pushed range (see ARParams)
  const void* mc_g; void* mc_w; void* mc_flags; const void* nv_flags;   // mode 2 (NVLS) only
};

TFDE_API size_t tfde_ar_desc_bytes() { return sizeof(TfdeARDesc); }

TFDE_API int tfde_allreduce_fused(const TfdeARDesc* d, cudaStream_t stream) {
  if (d->world < 1 || d->world > MAX_RANKS || d->blocks < 1 || d->blocks > MAX_BLOCKS) return TFDE_ERR_INVALID;
  if ((d->n_total & 3) || (d->slot_floats & 3) || d->n_extra > 32) return TFDE_ERR_INVALID;
  if (d->opt >= 1 && !d->s1) return TFDE_ERR_INVALID;
  if (d->opt == 3 && !d->s2) return TFDE_ERR_INVALID;
  ARParams p{};
  p.rank = d->rank; p.world = d->world; p.mode = d->mode; p.opt = d->opt; p.wire_bf16 = d->wire_bf16;
  p.n = d->n; p.n_total = d->n_total; p.slot_floats = d->slot_floats;
  p.g = d->g; p.w = d->w; p.s1 = d->s1; p.s2 = d->s2; p.w_bf16 = (__nv_bfloat16*)d->w_bf16;
  for (int i = 0; i < MAX_RANKS; ++i) {
    p.recv[i] = (float*)d->recv[i]; p.wpeer[i] = (float*)d->wpeer[i];
    p.wbf_peer[i] = (__nv_bfloat16*)d->wbf_peer[i]; p.flags[i] = (uint32_t*)d->flags[i];
  }
  p.epoch = (uint32_t*)d->epoch; p.metrics = d->metrics; p.n_extra = d->n_extra;
  p.hp = (TfdeHParams*)d->hp; p.done_counter = (unsigned*)d->done_counter; p.status = (int*)d->status;
  p.timeout_ns = d->timeout_ns ? d->timeout_ns : 20000000000ULL;
  p.pp_lo = d->pp_lo; p.pp_hi = d->pp_hi; p.pp_done = (const uint32_t*)d->pp_done;
  if (d->wire_bf16 || (p.pp_lo & 3) || (p.pp_hi & 3)) p.pp_done = nullptr;   // pushes are fp32, 16-byte groups
  if (d->mode == 2) {
    // NVLS: fp32 on the wire, no bf16 shadow (it would need its own multicast mapping), W >= 2
    if (!d->mc_g || !d->mc_w || !d->mc_flags || !d->nv_flags || d->wire_bf16 || d->w_bf16 || d->world < 2)
      return TFDE_ERR_INVALID;
    p.mc_g = (const float*)d->mc_g; p.mc_w = (float*)d->mc_w;
    p.mc_flags = (uint32_t*)d->mc_flags; p.nv_flags = (const uint32_t*)d->nv_flags;
    tfde_launch(allreduce_nvls_kernel, dim3(d->blocks), dim3(AR_THREADS), 0, stream, p);
    TFDE_RETURN_LAUNCH_STATUS();
  }
  tfde_launch(allreduce_fused_kernel, dim3(d->blocks), dim3(AR_THREADS), 0, stream, p);
  TFDE_RETURN_LAUNCH_STATUS();
}

TFDE_API int tfde_p2p_copy(float* dst, const float* src, long n, cudaStream_t stream) {
  long grid = (n / 4 + 255) / 256;
  if (grid > 148L * 4) grid = 148L * 4;
  if (grid < 1) grid = 1;
  p2p_copy_kernel<<<(int)grid, 256, 0, stream>>>(dst, src, n);
  TFDE_RETURN_LAUNCH_STATUS();
}
