// Persistent variant of the tcgen05 bf16 GEMM (see gemm_tcgen05.cu for the operand layout and the PTX
// helpers' rationale): one CTA per SM loops over output tiles, the TMEM accumulator is DOUBLE-BUFFERED
// (2 x BN columns), so the epilogue of tile i (tcgen05.ld -> bias/ReLU -> coalesced stores) overlaps the
// TMA/MMA main loop of tile i+1, and the per-CTA set-up (barrier init, TMEM allocation, tensor-map
// prefetch) is paid once per SM instead of once per tile. This is the kernel for the many small-K /
// output-bound GEMMs of a convolutional network (1x1 convolutions, K = 64..576), where the one-tile-per-CTA
// kernel spends most of its time in set-up and drain.
//
//   C[M,N] = act( A[M,K] . B[N,K]^T + bias[N] ),  A/B bf16 K-major, C fp32 and/or bf16
//
// Pipelines: smem ring full[s]/empty[s] (TMA <-> MMA, STAGES deep), TMEM ring tmem_full[a]/tmem_empty[a]
// (MMA <-> epilogue, 2 deep). Tile order is static: tile = blockIdx.x + i * gridDim.x, m-major.
#include "common.cuh"
#include <cuda.h>
#include <cuda_bf16.h>
#include <cstdlib>
#include <mutex>

namespace {

constexpr int P_BM = 128;
constexpr int P_BK = 64;
constexpr int P_THREADS = 192;
constexpr int P_UMMA_K = 16;
constexpr int P_STG_PITCH = 144;   // bytes per staged epilogue row: 128 of payload + 16 of padding
constexpr int P_EPI_BYTES = 4 * 2 * 32 * 128;   // epilogue staging: 4 warps x 2 TMA-store slabs (>= 4 x 32 x P_STG_PITCH)

__device__ __forceinline__ uint32_t p_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void p_mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(p_smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void p_mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(p_smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void p_mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(p_smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void p_mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok) : "r"(p_smem_u32(bar)), "r"(parity) : "memory");
  } while (!ok);
}
__device__ __forceinline__ void p_tma_load_2d(void* dst, const CUtensorMap* map, uint64_t* bar, int c_inner, int c_outer) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(p_smem_u32(dst)), "l"(map), "r"(p_smem_u32(bar)), "r"(c_inner), "r"(c_outer) : "memory");
}
// im2col-mode load of an NHWC activation: `pixelsPerColumn` pixels x `channelsPerPixel` channels of the tensor map,
// starting at the base pixel (w, h, n) of the bounding box and walking W, then H, then N with the map's traversal
// strides; (ow, oh) is the filter tap added to every pixel; out-of-image pixels are zero-filled. The box lands as
// one 128-byte (64 bf16 channel) row per pixel, SWIZZLE_128B: the same bytes a tiled {64, rows} box would leave.
__device__ __forceinline__ void p_tma_load_im2col(void* dst, const CUtensorMap* map, uint64_t* bar, int c, int w, int h,
                                                  int n, int ow, int oh) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.im2col.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2], {%7, %8};"
      ::"r"(p_smem_u32(dst)), "l"(map), "r"(p_smem_u32(bar)), "r"(c), "r"(w), "r"(h), "r"(n), "h"((uint16_t)ow),
        "h"((uint16_t)oh) : "memory");
}
// TMA store of one staged {64 x bf16 = 128 B, 32 rows} slab (SWIZZLE_128B in shared memory) into the output tensor;
// rows past the tensor's end are clipped by the tensor map. Bulk-group completion: wait_group.read N = the shared
// memory of all but the N newest groups may be overwritten.
__device__ __forceinline__ void p_tma_store_2d(const CUtensorMap* map, const void* src, int c_inner, int c_outer) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
               ::"l"(map), "r"(p_smem_u32(src)), "r"(c_inner), "r"(c_outer) : "memory");
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
__device__ __forceinline__ uint64_t p_smem_desc(uint32_t addr) {   // K-major, SWIZZLE_128B (see gemm_tcgen05.cu)
  return (uint64_t)((addr & 0x3FFFFu) >> 4) | (1ull << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
// MN-major operand (the non-reduced dimension is the contiguous one in global memory), SWIZZLE_128B:
// a TMA box of {64 MN elements = 128 B, 64 K rows} lands as 64 rows of 128 B; 8 rows form one 1024-byte
// swizzle atom (stride-dimension offset, K direction), consecutive 64-wide MN blocks are whole boxes
// 8192 bytes apart (leading-dimension offset).
__device__ __forceinline__ uint64_t p_smem_desc_mn(uint32_t addr) {
  return (uint64_t)((addr & 0x3FFFFu) >> 4) | ((uint64_t)(8192 >> 4) << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
__device__ __forceinline__ void p_umma(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void p_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(p_smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void p_tmem_ld(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr) : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

__device__ __forceinline__ void p_tmem_ld_nowait(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr) : "memory");
}
// accumulator registers (fp32 bit patterns) += bias[col + j], ReLU; both optional, warp-uniform
__device__ __forceinline__ void p_bias_act(uint32_t (&r)[32], const float* __restrict__ bias, int col, bool has_bias,
                                           bool relu) {
#pragma unroll
  for (int j = 0; j < 32; ++j) {
    float x = __uint_as_float(r[j]);
    if (has_bias) x += __ldg(bias + col + j);
    if (relu) x = fmaxf(x, 0.f);
    r[j] = __float_as_uint(x);
  }
}
// 8 consecutive fp32 accumulator values -> 8 bf16 (round to nearest even) in one 16-byte chunk
__device__ __forceinline__ uint4 p_pack8(const uint32_t (&r)[32], int j) {
  uint4 o;
  __nv_bfloat162 p0 = __floats2bfloat162_rn(__uint_as_float(r[j]), __uint_as_float(r[j + 1]));
  __nv_bfloat162 p1 = __floats2bfloat162_rn(__uint_as_float(r[j + 2]), __uint_as_float(r[j + 3]));
  __nv_bfloat162 p2 = __floats2bfloat162_rn(__uint_as_float(r[j + 4]), __uint_as_float(r[j + 5]));
  __nv_bfloat162 p3 = __floats2bfloat162_rn(__uint_as_float(r[j + 6]), __uint_as_float(r[j + 7]));
  o.x = *reinterpret_cast<uint32_t*>(&p0); o.y = *reinterpret_cast<uint32_t*>(&p1);
  o.z = *reinterpret_cast<uint32_t*>(&p2); o.w = *reinterpret_cast<uint32_t*>(&p3);
  return o;
}

struct PEpilogue {
  float* c_f32; __nv_bfloat16* c_bf16; long ldc;
  const float* bias; int relu;
  int M, N;
  // optional fused batch-norm statistics (bf16 output, staged path, no split-K): per 32-row group partial
  // column sums / sums of squares of the ROUNDED outputs, [ceil(M/32)][N] each -- the consumer of a
  // convolution's output is a batch-norm whose first pass would otherwise re-read the whole tensor
  float* stat_sum; float* stat_sq;
  // bf16 output written by TMA stores (tmC describes c_bf16 as [M, N] with 128-byte x 32-row boxes): the epilogue
  // warps only fill swizzled shared-memory slabs, the copy engine moves them out while the next slab is converted
  int tma_store;
};

// Convolution geometry of the implicit-GEMM modes (CONV != 0): the A operand is never materialised, the producer
// warp fetches it from the NHWC activation with im2col-mode TMA loads.
//   pixel grid = the positions the bounding box of the tensor map can take: [n][gh][gw], linear index = GEMM row
//   (CONV 1, 2) or GEMM reduction index (CONV 3); base coordinate of grid position (p, q) = (p*sh - off_h, q*sw - off_w)
struct PConv {
  int gw, ghw;          // grid width, grid height * width
  int sw, sh;           // traversal strides
  int off_w, off_h;     // -lower corner (the padding seen by tap 0)
  int kw, taps;         // filter width, KH*KW
  int cblk;             // 64-channel blocks per tap along the reduction (CONV 1: Cin/64, CONV 2: Cout/64)
  int cin;              // CONV 2: rows per tap of the [taps*Cin, Cout] weight matrix; CONV 3: Cin
};

// MNMAJ is a bit mask: bit 0 = A is MN-major (stored [K, M], M contiguous), bit 1 = B is MN-major (stored
// [K, N], N contiguous); a clear bit means K-major ([M, K] / [N, K], K contiguous).
//   0: A[M,K] B[N,K]   forward / data-gradient GEMMs with pre-transposed weights
//   2: A[M,K] B[K,N]   forward convolution / dense with the weights exactly as Keras stores them (HWIO, (in,out))
//   3: A[K,M] B[K,N]   filter gradient: activations [rows, Cin] and gradients [rows, Cout], reduced over rows
// so no transposed copies of weights, activations or gradients are needed.
// CONV selects how the A operand (and, for 2, the B operand) is addressed:
//   0: plain GEMM, tiled tensor maps
//   1: convolution forward, MNMAJ 2: A = im2col(x) [pixels, taps*Cin] fetched tap by tap (128 pixels x 64 channels per
//      k-block), B = HWIO weights [taps*Cin, Cout] as stored
//   2: data gradient of a stride-1 convolution, MNMAJ 0: A = im2col(dy) over the input grid with the mirrored padding,
//      B = the SAME weight matrix read K-major as [taps*Cin, Cout] rows (tap mirrored: taps-1-tap), no rotated copy
//   3: filter gradient, MNMAJ 3: A = im2col(x) MN-major (64 pixels x 64 channels per box, two boxes per 128-row
//      tile = two (tap, channel-block) pairs), B = dy [pixels, Cout], reduced over pixels
template <int BN, int MNMAJ, int CONV = 0>
__global__ void __launch_bounds__(P_THREADS, 1)
gemm_bf16_tcgen05_persistent_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                                    const __grid_constant__ CUtensorMap tmC, const PEpilogue ep_in, const int K,
                                    const int tiles_m, const int tiles_n, const int splits, float* __restrict__ ws,
                                    const PConv cv) {
  constexpr uint32_t A_BYTES = P_BM * P_BK * 2;
  constexpr uint32_t B_BYTES = BN * P_BK * 2;
  constexpr uint32_t STAGE_BYTES = A_BYTES + B_BYTES;
  constexpr int STAGES = (BN == 256) ? 4 : 6;
  constexpr uint32_t TMEM_COLS = 2 * BN < 32 ? 32 : 2 * BN;      // 128 / 256 / 512: powers of two
  extern __shared__ __align__(1024) uint8_t p_smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(p_smem_raw) + 1023) & ~(uintptr_t)1023);
  // epilogue staging: 4 warps x 32 rows x 144 bytes (padded, coalesced-store path) or 4 warps x 2 buffers x 32 rows x
  // 128 bytes (swizzled, TMA-store path); P_EPI_BYTES covers the larger
  float* epi_tiles = reinterpret_cast<float*>(smem + (size_t)STAGES * STAGE_BYTES);
  __shared__ __align__(8) uint64_t full_bar[STAGES];
  __shared__ __align__(8) uint64_t empty_bar[STAGES];
  __shared__ __align__(8) uint64_t tmem_full[2];
  __shared__ __align__(8) uint64_t tmem_empty[2];
  __shared__ uint32_t tmem_base_smem;

  // Programmatic dependent launch: let the next kernel of the stream get scheduled now; our own set-up
  // (barriers, TMEM allocation, tensor-map prefetch) overlaps the tail of the previous kernel, and every
  // thread waits for that kernel's memory before the first global access (pdl_wait below).
  pdl_launch_dependents();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int num_k = (K + P_BK - 1) / P_BK;
  // work item = (output tile, k split); split z covers k-blocks [z*kb_per, min(num_k, (z+1)*kb_per)).
  // With splits > 1 every item stores an fp32 partial tile into ws[z]; a second kernel adds them in z order.
  const int kb_per = (num_k + splits - 1) / splits;
  const int num_tiles = tiles_m * tiles_n * splits;

  if (threadIdx.x == 0) {
#pragma unroll
    for (int s = 0; s < STAGES; ++s) { p_mbar_init(&full_bar[s], 1); p_mbar_init(&empty_bar[s], 1); }
    p_mbar_init(&tmem_full[0], 1); p_mbar_init(&tmem_full[1], 1);
    p_mbar_init(&tmem_empty[0], 4); p_mbar_init(&tmem_empty[1], 4);    // one arrival per epilogue warp
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmA) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmB) : "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(p_smem_u32(&tmem_base_smem)), "r"(TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_d = tmem_base_smem;
  pdl_wait();

  if (warp == 0) {
    if (lane == 0) {
      uint32_t it = 0;                                   // k-blocks issued so far (across tiles)
      for (int item = blockIdx.x; item < num_tiles; item += gridDim.x) {
        const int tile = item / splits, z = item % splits;
        const int m0 = (tile / tiles_n) * P_BM, n0 = (tile % tiles_n) * BN;
        const int kb_end = min(num_k, (z + 1) * kb_per);
        // implicit GEMM: where this tile starts on the pixel grid (CONV 1, 2) / which (tap, channel block) pairs its
        // two 64-row halves are (CONV 3)
        int img = 0, bw = 0, bh = 0, tap3[2] = {0, 0}, ch3[2] = {0, 0}, nbox3 = 0;
        if (CONV == 1 || CONV == 2) {
          img = m0 / cv.ghw;
          const int rem = m0 - img * cv.ghw, p = rem / cv.gw;
          bh = p * cv.sh - cv.off_h; bw = (rem - p * cv.gw) * cv.sw - cv.off_w;
        }
        if (CONV == 3) {
#pragma unroll
          for (int j = 0; j < 2; ++j) {
            const int mi = m0 + j * 64;
            if (mi < ep_in.M) { tap3[j] = mi / cv.cin; ch3[j] = mi - tap3[j] * cv.cin; nbox3 = j + 1; }
          }
        }
        for (int kb = z * kb_per; kb < kb_end; ++kb, ++it) {
          const uint32_t s = it % STAGES, ph = (it / STAGES) & 1u;
          p_mbar_wait(&empty_bar[s], ph ^ 1u);
          p_mbar_expect_tx(&full_bar[s], CONV == 3 ? (uint32_t)nbox3 * 8192u + B_BYTES : STAGE_BYTES);
          uint8_t* a_dst = smem + (size_t)s * STAGE_BYTES;
          if (CONV == 1 || CONV == 2) {
            const int tap = kb / cv.cblk, cb = kb - tap * cv.cblk, r = tap / cv.kw;
            p_tma_load_im2col(a_dst, &tmA, &full_bar[s], cb * 64, bw, bh, img, tap - r * cv.kw, r);
            if (CONV == 2) {
              p_tma_load_2d(a_dst + A_BYTES, &tmB, &full_bar[s], cb * 64, (cv.taps - 1 - tap) * cv.cin + n0);
              continue;
            }
          } else if (CONV == 3) {
            const int pix = kb * P_BK, n3 = pix / cv.ghw, rem = pix - n3 * cv.ghw, p = rem / cv.gw;
            const int h3 = p * cv.sh - cv.off_h, w3 = (rem - p * cv.gw) * cv.sw - cv.off_w;
            for (int j = 0; j < nbox3; ++j) {
              const int r = tap3[j] / cv.kw;
              p_tma_load_im2col(a_dst + j * 8192, &tmA, &full_bar[s], ch3[j], w3, h3, n3, tap3[j] - r * cv.kw, r);
            }
          } else if (MNMAJ & 1) {
#pragma unroll
            for (int j = 0; j < P_BM / 64; ++j) p_tma_load_2d(a_dst + j * 8192, &tmA, &full_bar[s], m0 + j * 64, kb * P_BK);
          } else {
            p_tma_load_2d(a_dst, &tmA, &full_bar[s], kb * P_BK, m0);
          }
          if (MNMAJ & 2) {
#pragma unroll
            for (int j = 0; j < BN / 64; ++j) p_tma_load_2d(a_dst + A_BYTES + j * 8192, &tmB, &full_bar[s], n0 + j * 64, kb * P_BK);
          } else {
            p_tma_load_2d(a_dst + A_BYTES, &tmB, &full_bar[s], kb * P_BK, n0);
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(P_BM >> 4) << 24) |
                             ((MNMAJ & 1) ? (1u << 15) : 0u) | ((MNMAJ & 2) ? (1u << 16) : 0u);
      uint32_t it = 0, tl = 0;                           // tl: tiles done by this CTA
      for (int item = blockIdx.x; item < num_tiles; item += gridDim.x, ++tl) {
        const int z = item % splits;
        const int kb_begin = z * kb_per, kb_end = min(num_k, (z + 1) * kb_per);
        const uint32_t acc = tl & 1u, aph = (tl >> 1) & 1u;
        p_mbar_wait(&tmem_empty[acc], aph ^ 1u);         // epilogue has drained this accumulator
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        for (int kb = kb_begin; kb < kb_end; ++kb, ++it) {
          const uint32_t s = it % STAGES, ph = (it / STAGES) & 1u;
          p_mbar_wait(&full_bar[s], ph);
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          const uint32_t a_addr = p_smem_u32(smem + (size_t)s * STAGE_BYTES);
          const uint32_t b_addr = a_addr + A_BYTES;
#pragma unroll
          for (int k = 0; k < P_BK / P_UMMA_K; ++k) {
            // K-major: 16 elements = 32 bytes along the swizzled row; MN-major: 16 K rows = 2 atoms = 2048 bytes
            const uint64_t da = (MNMAJ & 1) ? p_smem_desc_mn(a_addr + k * 2048) : p_smem_desc(a_addr + k * P_UMMA_K * 2);
            const uint64_t db = (MNMAJ & 2) ? p_smem_desc_mn(b_addr + k * 2048) : p_smem_desc(b_addr + k * P_UMMA_K * 2);
            p_umma(tmem_d + acc * BN, da, db, idesc, (kb != kb_begin || k != 0) ? 1u : 0u);
          }
          p_commit(&empty_bar[s]);
        }
        p_commit(&tmem_full[acc]);
      }
    }
  } else {
    const int q = warp & 3;
    uint32_t tl = 0, slab = 0;                          // slab: TMA-store slabs issued by this warp (buffer = slab & 1)
    const bool tma_out = ep_in.tma_store != 0 && splits == 1 && ep_in.c_bf16 != nullptr && ep_in.c_f32 == nullptr;
    for (int item = blockIdx.x; item < num_tiles; item += gridDim.x, ++tl) {
      const int tile = item / splits, z = item % splits;
      const uint32_t acc = tl & 1u, aph = (tl >> 1) & 1u;
      const int m0 = (tile / tiles_n) * P_BM, n0 = (tile % tiles_n) * BN;
      PEpilogue ep = ep_in;
      if (splits > 1) {   // raw fp32 partial into ws[z][M][N]; bias / activation / cast happen in the reduce kernel
        ep.c_f32 = ws + (size_t)z * ep_in.M * ep_in.N; ep.c_bf16 = nullptr; ep.ldc = ep_in.N; ep.bias = nullptr; ep.relu = 0;
      }
      p_mbar_wait(&tmem_full[acc], aph);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const int row0 = m0 + q * 32;
      // Staged epilogue (the common case: one output dtype, 16-byte aligned rows). A tcgen05.ld hands every
      // lane one ROW; storing that directly makes each warp store touch 32 different lines with 16 bytes each
      // (measured on the K=64, N=256 1x1 convolutions: 213 us for 129 MB, the store path saturated by
      // half-sector requests). Instead each warp bounces a 32-row x 128-byte slab through its own padded
      // shared-memory tile (144-byte pitch: conflict-free both ways) and writes it back with 8 lanes per row:
      // every store instruction covers 4 complete 128-byte lines.
      int c_done = 0;
      if (((ep.c_f32 == nullptr) != (ep.c_bf16 == nullptr)) && (ep.ldc & 7) == 0) {
        uint8_t* stg = reinterpret_cast<uint8_t*>(epi_tiles) + (tma_out ? q * 8192 : q * (32 * P_STG_PITCH));
        const bool out16 = ep.c_bf16 != nullptr;
        const int pass = out16 ? 64 : 32;                       // columns per 128-byte row slab
#pragma unroll 1
        for (int c0 = 0; c0 + pass <= BN && n0 + c0 + pass <= ep.N; c0 += pass) {
          const int nbase = n0 + c0;
          uint8_t* sb = stg + (tma_out ? (slab & 1u) * 4096u : 0u);
          if (tma_out) {                                   // the store that last used this buffer has read it
            if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
            __syncwarp();
          }
          // both 32-column TMEM loads of a bf16 slab are in flight before the one wait; bias / activation only
          // when present (warp-uniform), then straight from the accumulator registers into packed 16-byte chunks
          const bool has_bias = ep.bias != nullptr, relu = ep.relu != 0;
          const uint32_t tcol = tmem_d + ((uint32_t)(q * 32) << 16) + acc * BN + (uint32_t)c0;
          if (out16) {
            uint32_t ra[32], rb[32];
            p_tmem_ld_nowait(tcol, ra);
            p_tmem_ld_nowait(tcol + 32, rb);
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            if (has_bias || relu) {
              p_bias_act(ra, ep.bias, nbase, has_bias, relu);
              p_bias_act(rb, ep.bias, nbase + 32, has_bias, relu);
            }
#pragma unroll
            for (int c = 0; c < 8; ++c) {
              const uint4 o = c < 4 ? p_pack8(ra, c * 8) : p_pack8(rb, (c - 4) * 8);
              if (tma_out)                                  // 128-byte rows, 16-byte chunk c of row r at c ^ (r & 7)
                *reinterpret_cast<uint4*>(sb + lane * 128 + ((c ^ (lane & 7)) << 4)) = o;
              else
                *reinterpret_cast<uint4*>(stg + lane * P_STG_PITCH + c * 16) = o;
            }
          } else {
            uint32_t r[32];
            p_tmem_ld(tcol, r);
            if (has_bias || relu) p_bias_act(r, ep.bias, nbase, has_bias, relu);
            uint4* srow = reinterpret_cast<uint4*>(stg + lane * P_STG_PITCH);
#pragma unroll
            for (int j = 0; j < 32; j += 4) srow[j >> 2] = make_uint4(r[j], r[j + 1], r[j + 2], r[j + 3]);
          }
          if (tma_out) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
          __syncwarp();
          if (ep.stat_sum != nullptr && out16) {
            // fused batch-norm statistics: lane owns columns 2*lane, 2*lane+1 of the 64-column slab and walks
            // the 32 staged rows (4-byte conflict-free reads of the rounded bf16 pairs)
            float s0 = 0.f, s1 = 0.f, q0 = 0.f, q1 = 0.f;
            const int nrows = min(32, ep.M - row0);
#pragma unroll 8
            for (int rr = 0; rr < 32; ++rr) {
              if (rr < nrows) {
                const uint32_t pr = tma_out
                    ? *reinterpret_cast<const uint32_t*>(sb + rr * 128 + (((lane >> 2) ^ (rr & 7)) << 4) + (lane & 3) * 4)
                    : *reinterpret_cast<const uint32_t*>(stg + rr * P_STG_PITCH + lane * 4);
                const float a = __uint_as_float(pr << 16), b = __uint_as_float(pr & 0xffff0000u);
                s0 += a; s1 += b; q0 = fmaf(a, a, q0); q1 = fmaf(b, b, q1);
              }
            }
            if (nrows > 0) {
              const size_t o = (size_t)(row0 >> 5) * ep.N + nbase + 2 * lane;
              *reinterpret_cast<float2*>(ep.stat_sum + o) = make_float2(s0, s1);
              *reinterpret_cast<float2*>(ep.stat_sq + o) = make_float2(q0, q1);
            }
          }
          if (tma_out) {
            if (lane == 0 && row0 < ep.M) p_tma_store_2d(&tmC, sb, nbase, row0);
            if (row0 < ep.M) ++slab;
            c_done = c0 + pass;
            continue;
          }
          uint8_t* gbase = out16 ? reinterpret_cast<uint8_t*>(ep.c_bf16 + nbase) : reinterpret_cast<uint8_t*>(ep.c_f32 + nbase);
          const size_t gpitch = (size_t)ep.ldc * (out16 ? 2 : 4);
#pragma unroll
          for (int it = 0; it < 8; ++it) {
            const int rr = it * 4 + (lane >> 3), piece = lane & 7;
            const uint4 val = *reinterpret_cast<const uint4*>(stg + rr * P_STG_PITCH + piece * 16);
            const int row = row0 + rr;
            if (row < ep.M) *reinterpret_cast<uint4*>(gbase + (size_t)row * gpitch + piece * 16) = val;
          }
          __syncwarp();
          c_done = c0 + pass;
        }
      }
#pragma unroll 1
      for (int c0 = c_done; c0 < BN; c0 += 32) {
        const int nbase = n0 + c0;
        if (nbase >= ep.N) break;
        uint32_t r[32];
        p_tmem_ld(tmem_d + ((uint32_t)(q * 32) << 16) + acc * BN + (uint32_t)c0, r);
        // Direct row-segment stores: lane == output row, 32 consecutive columns per thread. Every thread
        // writes whole 16-byte pieces of its own 128-byte (fp32) / 64-byte (bf16) row segment; the pieces of
        // one cache line arrive back to back and merge in L2 (measured faster than a shared-memory transpose).
        const int row = row0 + lane;
        if (row < ep.M) {
          const bool full = (nbase + 32 <= ep.N) && ((ep.ldc & 7) == 0);
          float v[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            float x = __uint_as_float(r[j]);
            if (ep.bias != nullptr) x += (full || nbase + j < ep.N) ? __ldg(ep.bias + nbase + j) : 0.f;
            v[j] = ep.relu ? fmaxf(x, 0.f) : x;
          }
          if (ep.c_f32 != nullptr) {
            float* dst = ep.c_f32 + (size_t)row * ep.ldc + nbase;
            if (full) {
#pragma unroll
              for (int j = 0; j < 32; j += 4) *reinterpret_cast<float4*>(dst + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j) if (nbase + j < ep.N) dst[j] = v[j];
            }
          }
          if (ep.c_bf16 != nullptr) {
            __nv_bfloat16* dst = ep.c_bf16 + (size_t)row * ep.ldc + nbase;
            if (full) {
#pragma unroll
              for (int j = 0; j < 32; j += 8) {
                uint4 o;
                __nv_bfloat162 p0 = __floats2bfloat162_rn(v[j], v[j + 1]), p1 = __floats2bfloat162_rn(v[j + 2], v[j + 3]);
                __nv_bfloat162 p2 = __floats2bfloat162_rn(v[j + 4], v[j + 5]), p3 = __floats2bfloat162_rn(v[j + 6], v[j + 7]);
                o.x = *reinterpret_cast<uint32_t*>(&p0); o.y = *reinterpret_cast<uint32_t*>(&p1);
                o.z = *reinterpret_cast<uint32_t*>(&p2); o.w = *reinterpret_cast<uint32_t*>(&p3);
                *reinterpret_cast<uint4*>(dst + j) = o;
              }
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j) if (nbase + j < ep.N) dst[j] = __float2bfloat16(v[j]);
            }
          }
        }
      }
      // this warp is done reading the accumulator: hand it back to the MMA warp
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      __syncwarp();
      if (lane == 0) p_mbar_arrive(&tmem_empty[acc]);
    }
  }

  if (warp >= 2 && lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");   // TMA stores have landed
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"(TMEM_COLS) : "memory");
  }
}

// out[i] = act(sum_z ws[z][i] + bias[col]) in z order; fp32 and/or bf16 output with row stride ldc.
__global__ void splitk_reduce_kernel(const float* __restrict__ ws, float* __restrict__ c_f32,
                                     __nv_bfloat16* __restrict__ c_bf16, long ldc, const float* __restrict__ bias,
                                     int relu, int M, int N, int splits) {
  TFDE_PDL_PROLOGUE();
  const long total = (long)M * N;
  if ((N & 3) == 0 && (ldc & 3) == 0) {
    // 16-byte path: 4 consecutive columns per thread, 8 partials in flight
    const long total4 = total >> 2;
    const float4* ws4 = reinterpret_cast<const float4*>(ws);
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total4; i += (long)gridDim.x * blockDim.x) {
      float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int z0 = 0; z0 < splits; z0 += 8) {
        float4 v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u)
          v[u] = (z0 + u < splits) ? __ldcg(ws4 + (size_t)(z0 + u) * total4 + i) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int u = 0; u < 8; ++u) { s.x += v[u].x; s.y += v[u].y; s.z += v[u].z; s.w += v[u].w; }
      }
      const int col = (int)((i << 2) % N);
      const long row = (i << 2) / N;
      if (bias != nullptr) {
        const float4 bv = __ldg(reinterpret_cast<const float4*>(bias + col));
        s.x += bv.x; s.y += bv.y; s.z += bv.z; s.w += bv.w;
      }
      if (relu) { s.x = fmaxf(s.x, 0.f); s.y = fmaxf(s.y, 0.f); s.z = fmaxf(s.z, 0.f); s.w = fmaxf(s.w, 0.f); }
      if (c_f32 != nullptr) *reinterpret_cast<float4*>(c_f32 + row * ldc + col) = s;
      if (c_bf16 != nullptr) {
        __nv_bfloat162 p0 = __floats2bfloat162_rn(s.x, s.y), p1 = __floats2bfloat162_rn(s.z, s.w);
        uint2 o; o.x = *reinterpret_cast<uint32_t*>(&p0); o.y = *reinterpret_cast<uint32_t*>(&p1);
        *reinterpret_cast<uint2*>(c_bf16 + row * ldc + col) = o;
      }
    }
    return;
  }
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    float s = 0.f;
    for (int z0 = 0; z0 < splits; z0 += 4) {
      float v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) v[u] = (z0 + u < splits) ? __ldcg(ws + (size_t)(z0 + u) * total + i) : 0.f;
#pragma unroll
      for (int u = 0; u < 4; ++u) s += v[u];
    }
    const int col = (int)(i % N);
    const long row = i / N;
    if (bias != nullptr) s += __ldg(bias + col);
    if (relu) s = fmaxf(s, 0.f);
    if (c_f32 != nullptr) c_f32[row * ldc + col] = s;
    if (c_bf16 != nullptr) c_bf16[row * ldc + col] = __float2bfloat16(s);
  }
}

// Small outputs with many splits (the 1x1 filter gradients: 64x256 outputs, 64 partials) are latency-bound in
// the kernel above: few threads, each walking all partials. Here ZP adjacent lanes share one 4-column group,
// each sums every ZP-th partial with all loads in flight, and a fixed xor-shuffle tree combines them
// (fixed order => still deterministic). Requires N % 4 == 0 and ldc % 4 == 0.
template <int ZP>
__global__ void splitk_reduce_zpar_kernel(const float* __restrict__ ws, float* __restrict__ c_f32,
                                          __nv_bfloat16* __restrict__ c_bf16, long ldc, const float* __restrict__ bias,
                                          int relu, int M, int N, int splits) {
  TFDE_PDL_PROLOGUE();
  const long total4 = ((long)M * N) >> 2;
  const float4* ws4 = reinterpret_cast<const float4*>(ws);
  const long tid = (long)blockIdx.x * blockDim.x + threadIdx.x;
  const long i = tid / ZP;
  const int zl = (int)(tid % ZP);
  float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
  if (i < total4) {
    for (int z0 = zl; z0 < splits; z0 += 8 * ZP) {
      float4 v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u)
        v[u] = (z0 + u * ZP < splits) ? __ldcg(ws4 + (size_t)(z0 + u * ZP) * total4 + i) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int u = 0; u < 8; ++u) { s.x += v[u].x; s.y += v[u].y; s.z += v[u].z; s.w += v[u].w; }
    }
  }
#pragma unroll
  for (int off = ZP / 2; off >= 1; off >>= 1) {
    s.x += __shfl_xor_sync(0xffffffffu, s.x, off); s.y += __shfl_xor_sync(0xffffffffu, s.y, off);
    s.z += __shfl_xor_sync(0xffffffffu, s.z, off); s.w += __shfl_xor_sync(0xffffffffu, s.w, off);
  }
  if (i < total4 && zl == 0) {
    const int col = (int)((i << 2) % N);
    const long row = (i << 2) / N;
    if (bias != nullptr) {
      const float4 bv = __ldg(reinterpret_cast<const float4*>(bias + col));
      s.x += bv.x; s.y += bv.y; s.z += bv.z; s.w += bv.w;
    }
    if (relu) { s.x = fmaxf(s.x, 0.f); s.y = fmaxf(s.y, 0.f); s.z = fmaxf(s.z, 0.f); s.w = fmaxf(s.w, 0.f); }
    if (c_f32 != nullptr) *reinterpret_cast<float4*>(c_f32 + row * ldc + col) = s;
    if (c_bf16 != nullptr) {
      __nv_bfloat162 p0 = __floats2bfloat162_rn(s.x, s.y), p1 = __floats2bfloat162_rn(s.z, s.w);
      uint2 o; o.x = *reinterpret_cast<uint32_t*>(&p0); o.y = *reinterpret_cast<uint32_t*>(&p1);
      *reinterpret_cast<uint2*>(c_bf16 + row * ldc + col) = o;
    }
  }
}

typedef CUresult (*PEncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                              const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                              CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
PEncodeFn p_encode = nullptr;
std::mutex p_encode_mu;

int p_resolve_encode() {
  std::lock_guard<std::mutex> lock(p_encode_mu);
  if (p_encode == nullptr) {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    const cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
    if (e != cudaSuccess) return (int)e;
    if (q != cudaDriverEntryPointSuccess || fn == nullptr) return (int)cudaErrorNotSupported;
    p_encode = reinterpret_cast<PEncodeFn>(fn);
  }
  return 0;
}

int p_make_map(CUtensorMap* map, const void* ptr, long rows, long cols, int box_rows, int box_cols = P_BK) {
  {
    const int rc0 = p_resolve_encode();
    if (rc0 != 0) return rc0;
  }
  cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t gstride[1] = {(cuuint64_t)cols * 2};
  cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  const CUresult r = p_encode(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), gdim, gstride, box, estr,
                              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : (int)cudaErrorInvalidValue;
}

// rows x cols view with row stride `ld` elements
int p_make_map_ld(CUtensorMap* map, const void* ptr, long rows, long cols, long ld, int box_rows, int box_cols) {
  const int rc = p_resolve_encode();
  if (rc != 0) return rc;
  cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t gstride[1] = {(cuuint64_t)ld * 2};
  cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  const CUresult r = p_encode(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), gdim, gstride, box, estr,
                              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : (int)cudaErrorInvalidValue;
}

// TMA-store epilogue (bf16-only output in whole 64-column slabs): build the output map, set ep.tma_store.
// TFDE_TMA_STORE=0 keeps the coalesced-store epilogue.
bool p_tma_store_wanted() {
  static const bool on = [] { const char* e = getenv("TFDE_TMA_STORE"); return e == nullptr || e[0] != '0'; }();
  return on;
}
int p_prepare_store(PEpilogue& ep, CUtensorMap* tc) {
  ep.tma_store = 0;
  if (!p_tma_store_wanted() || ep.c_bf16 == nullptr || ep.c_f32 != nullptr || (ep.N & 63) || (ep.ldc & 7) ||
      (reinterpret_cast<uintptr_t>(ep.c_bf16) & 15))
    return 0;
  const int rc = p_make_map_ld(tc, ep.c_bf16, ep.M, ep.N, ep.ldc, 32, 64);
  if (rc != 0) return rc;
  ep.tma_store = 1;
  return 0;
}

template <int BN, int MNMAJ, int CONV = 0>
int p_launch(const CUtensorMap& ta, const CUtensorMap& tb, const PEpilogue& ep, int M, int N, int K, float* ws,
             size_t ws_bytes, cudaStream_t stream, const PConv& cv = PConv{}, const CUtensorMap* tc = nullptr) {
  constexpr int STAGES = (BN == 256) ? 4 : 6;
  constexpr size_t smem = (size_t)STAGES * (P_BM * P_BK * 2 + BN * P_BK * 2) + P_EPI_BYTES + 1024;
  static size_t configured[TFDE_MAX_DEVICES] = {};
  { const cudaError_t e = tfde_ensure_dyn_smem(gemm_bf16_tcgen05_persistent_kernel<BN, MNMAJ, CONV>, smem, configured); if (e != cudaSuccess) return (int)e; }
  const int tiles_m = (M + P_BM - 1) / P_BM, tiles_n = (N + BN - 1) / BN;
  const int tiles = tiles_m * tiles_n;
  const int num_k = (K + P_BK - 1) / P_BK;
  // split K when the output has too few tiles to fill the machine and the reduction is long (the filter-
  // gradient GEMMs of a convolution: output [KH*KW*Cin, Cout], reduction over B*Ho*Wo)
  int splits = 1;
  if (ws != nullptr && tiles < 148 && num_k >= 8) {
    // choose the split count by a cost model in k-block units: waves of 148 CTAs x (k-blocks + epilogue) per item,
    // plus the second kernel that adds the partials. (A plain ceil(148 / tiles) leaves 150..216 items for the
    // filter-gradient shapes of a ResNet: a second wave for a handful of CTAs, twice the time.)
    const size_t per = (size_t)M * N * sizeof(float);
    int smax = num_k / 4;
    if (smax > 148) smax = 148;
    if ((size_t)smax * per > ws_bytes) smax = (int)(ws_bytes / per);
    double best = 1e30;
    for (int s = 1; s <= smax; ++s) {
      const int kb_per = (num_k + s - 1) / s;
      const int s_eff = (num_k + kb_per - 1) / kb_per;            // no empty split
      if (s_eff != s) continue;
      const int waves = (tiles * s + 147) / 148;
      const double cost = (double)waves * (kb_per + BN / 32) + (s > 1 ? 4.0 + (double)s * (double)per / 2.0e6 : 0.0);
      if (cost < best) { best = cost; splits = s; }
    }
  }
  const int items = tiles * splits;
  const int grid = items < 148 ? items : 148;
  tfde_launch(gemm_bf16_tcgen05_persistent_kernel<BN, MNMAJ, CONV>, dim3(grid), dim3(P_THREADS), smem, stream, ta, tb,
              tc != nullptr ? *tc : ta, ep, K, tiles_m, tiles_n, splits, ws, cv);
  if (splits > 1) {
    const long total4 = (long)M * N / 4;
    const bool vec = (N & 3) == 0 && (ep.ldc & 3) == 0;
    if (vec && splits >= 8 && total4 * 8 <= 148L * 8 * 256) {      // small output, many partials: z-parallel lanes
      const long zblocks = (total4 * 8 + 255) / 256;
      tfde_launch(splitk_reduce_zpar_kernel<8>, dim3((int)zblocks), dim3(256), 0, stream, (const float*)ws, ep.c_f32,
                  ep.c_bf16, ep.ldc, ep.bias, ep.relu, M, N, splits);
      TFDE_RETURN_LAUNCH_STATUS();
    }
    long blocks = (total4 + 255) / 256;
    if (blocks > 148L * 8) blocks = 148L * 8;
    if (blocks < 1) blocks = 1;
    tfde_launch(splitk_reduce_kernel, dim3((int)blocks), dim3(256), 0, stream, (const float*)ws, ep.c_f32, ep.c_bf16, ep.ldc,
                ep.bias, ep.relu, M, N, splits);
  }
  TFDE_RETURN_LAUNCH_STATUS();
}

}  // namespace

// `ws` (optional, ws_bytes): fp32 scratch for deterministic split-K (partials summed in split order by a
// second kernel); pass NULL to disable splitting.
// Same contract as tfde_gemm_bf16_tn (K % 8 == 0, 16-byte aligned K-major bf16 operands); persistent
// one-CTA-per-SM schedule with a double-buffered TMEM accumulator. First call per tile shape sets a
// function attribute: make it outside of stream capture.
TFDE_API int tfde_gemm_bf16_tn_persistent(const void* A, const void* B, float* c_f32, void* c_bf16, long ldc,
                                          const float* bias, int relu, int M, int N, int K, float* ws,
                                          size_t ws_bytes, cudaStream_t stream) {
  if (M <= 0 || N <= 0 || K <= 0 || (K & 7) || !A || !B || (!c_f32 && !c_bf16)) return TFDE_ERR_INVALID;
  if ((reinterpret_cast<uintptr_t>(A) & 15) || (reinterpret_cast<uintptr_t>(B) & 15)) return TFDE_ERR_INVALID;
  const int BN = N <= 64 ? 64 : (N <= 128 ? 128 : 256);
  CUtensorMap ta, tb;
  int rc = p_make_map(&ta, A, M, K, P_BM);
  if (rc != 0) return rc;
  rc = p_make_map(&tb, B, N, K, BN);
  if (rc != 0) return rc;
  PEpilogue ep{c_f32, (__nv_bfloat16*)c_bf16, ldc, bias, relu, M, N};
  CUtensorMap tc;
  rc = p_prepare_store(ep, &tc);
  if (rc != 0) return rc;
  const CUtensorMap* tcp = ep.tma_store ? &tc : nullptr;
  if (BN == 64) return p_launch<64, 0>(ta, tb, ep, M, N, K, ws, ws_bytes, stream, PConv{}, tcp);
  if (BN == 128) return p_launch<128, 0>(ta, tb, ep, M, N, K, ws, ws_bytes, stream, PConv{}, tcp);
  return p_launch<256, 0>(ta, tb, ep, M, N, K, ws, ws_bytes, stream, PConv{}, tcp);
}

// C[M,N] = A[K,M]^T . B[K,N]: both operands "MN-major" (row-major with the REDUCED dimension as rows), which
// is how a filter-gradient GEMM finds them: A = activations / im2col [rows, Kc], B = output gradients
// [rows, Cout], reduced over rows. No transposed copies; M % 8 == 0 and N % 8 == 0 (16-byte TMA strides),
// K arbitrary (rows past K are zero-filled by TMA). lda / ldb: row strides in elements (multiples of 8).
TFDE_API int tfde_gemm_bf16_nt_persistent(const void* A, const void* B, float* c_f32, void* c_bf16, long ldc,
                                          int M, int N, int K, long lda, long ldb, float* ws, size_t ws_bytes,
                                          cudaStream_t stream) {
  if (M <= 0 || N <= 0 || K <= 0 || (M & 7) || (N & 7) || (lda & 7) || (ldb & 7) || lda < M || ldb < N || !A || !B ||
      (!c_f32 && !c_bf16))
    return TFDE_ERR_INVALID;
  if ((reinterpret_cast<uintptr_t>(A) & 15) || (reinterpret_cast<uintptr_t>(B) & 15)) return TFDE_ERR_INVALID;
  const int BN = N <= 64 ? 64 : (N <= 128 ? 128 : 256);
  CUtensorMap ta, tb;
  int rc = p_make_map_ld(&ta, A, K, M, lda, P_BK, 64);
  if (rc != 0) return rc;
  rc = p_make_map_ld(&tb, B, K, N, ldb, P_BK, 64);
  if (rc != 0) return rc;
  PEpilogue ep{c_f32, (__nv_bfloat16*)c_bf16, ldc, nullptr, 0, M, N};
  CUtensorMap tc;
  rc = p_prepare_store(ep, &tc);
  if (rc != 0) return rc;
  const CUtensorMap* tcp = ep.tma_store ? &tc : nullptr;
  if (BN == 64) return p_launch<64, 3>(ta, tb, ep, M, N, K, ws, ws_bytes, stream, PConv{}, tcp);
  if (BN == 128) return p_launch<128, 3>(ta, tb, ep, M, N, K, ws, ws_bytes, stream, PConv{}, tcp);
  return p_launch<256, 3>(ta, tb, ep, M, N, K, ws, ws_bytes, stream, PConv{}, tcp);
}

// C[M,N] = act(A[M,K] . B[K,N] + bias): A K-major, B MN-major -- a convolution / dense forward with the weight
// matrix exactly as Keras stores it ([KH*KW*Cin, Cout] = HWIO flattened, or (in, out)): no per-step
// transposed weight copy. K % 8 == 0 is required of A's row stride only (lda); B may have fewer than lda rows
// (`kb` valid rows: padded im2col columns meet zero-filled weight rows). N % 8 == 0.
TFDE_API int tfde_gemm_bf16_nn_persistent(const void* A, const void* B, float* c_f32, void* c_bf16, long ldc,
                                          const float* bias, int relu, int M, int N, int K, int kb, long ldb,
                                          float* ws, size_t ws_bytes, cudaStream_t stream) {
  if (M <= 0 || N <= 0 || K <= 0 || (K & 7) || (N & 7) || (ldb & 7) || ldb < N || kb <= 0 || kb > K || !A || !B ||
      (!c_f32 && !c_bf16))
    return TFDE_ERR_INVALID;
  if ((reinterpret_cast<uintptr_t>(A) & 15) || (reinterpret_cast<uintptr_t>(B) & 15)) return TFDE_ERR_INVALID;
  const int BN = N <= 64 ? 64 : (N <= 128 ? 128 : 256);
  CUtensorMap ta, tb;
  int rc = p_make_map(&ta, A, M, K, P_BM);
  if (rc != 0) return rc;
  rc = p_make_map_ld(&tb, B, kb, N, ldb, P_BK, 64);
  if (rc != 0) return rc;
  PEpilogue ep{c_f32, (__nv_bfloat16*)c_bf16, ldc, bias, relu, M, N};
  CUtensorMap tc;
  rc = p_prepare_store(ep, &tc);
  if (rc != 0) return rc;
  const CUtensorMap* tcp = ep.tma_store ? &tc : nullptr;
  if (BN == 64) return p_launch<64, 2>(ta, tb, ep, M, N, K, ws, ws_bytes, stream, PConv{}, tcp);
  if (BN == 128) return p_launch<128, 2>(ta, tb, ep, M, N, K, ws, ws_bytes, stream, PConv{}, tcp);
  return p_launch<256, 2>(ta, tb, ep, M, N, K, ws, ws_bytes, stream, PConv{}, tcp);
}

// Same GEMM with bf16 output only, plus the batch-norm statistics of that output fused into the epilogue:
// stat_sum / stat_sq [ceil(M/32)][N] receive per-32-row-group column sums and sums of squares of the ROUNDED
// values (what the batch-norm will read back). No split-K (partial tiles have no statistics): use it when the
// output has enough tiles to fill the machine. N % 64 == 0, ldc % 8 == 0.
TFDE_API int tfde_gemm_bf16_nn_stats_persistent(const void* A, const void* B, void* c_bf16, long ldc, const float* bias,
                                                int M, int N, int K, int kb, long ldb, float* stat_sum,
                                                float* stat_sq, cudaStream_t stream) {
  if (M <= 0 || N <= 0 || K <= 0 || (K & 7) || (N & 63) || (ldc & 7) || (ldb & 7) || ldb < N || kb <= 0 || kb > K ||
      !A || !B || !c_bf16 || !stat_sum || !stat_sq)
    return TFDE_ERR_INVALID;
  if ((reinterpret_cast<uintptr_t>(A) & 15) || (reinterpret_cast<uintptr_t>(B) & 15) ||
      (reinterpret_cast<uintptr_t>(c_bf16) & 15) || (reinterpret_cast<uintptr_t>(stat_sum) & 7) ||
      (reinterpret_cast<uintptr_t>(stat_sq) & 7))
    return TFDE_ERR_INVALID;
  const int BN = N <= 64 ? 64 : (N <= 128 ? 128 : 256);
  CUtensorMap ta, tb;
  int rc = p_make_map(&ta, A, M, K, P_BM);
  if (rc != 0) return rc;
  rc = p_make_map_ld(&tb, B, kb, N, ldb, P_BK, 64);
  if (rc != 0) return rc;
  PEpilogue ep{nullptr, (__nv_bfloat16*)c_bf16, ldc, bias, 0, M, N, stat_sum, stat_sq};
  CUtensorMap tc;
  rc = p_prepare_store(ep, &tc);
  if (rc != 0) return rc;
  const CUtensorMap* tcp = ep.tma_store ? &tc : nullptr;
  if (BN == 64) return p_launch<64, 2>(ta, tb, ep, M, N, K, nullptr, 0, stream, PConv{}, tcp);
  if (BN == 128) return p_launch<128, 2>(ta, tb, ep, M, N, K, nullptr, 0, stream, PConv{}, tcp);
  return p_launch<256, 2>(ta, tb, ep, M, N, K, nullptr, 0, stream, PConv{}, tcp);
}

// ---------------------------------------------------------------------------------------------------------------
// Implicit-GEMM convolutions (reference layers: /root/reference/mnist_keras_distributed.py:83-98 Conv2D,
// /root/reference/distributed_with_keras.py:34): the A operand is fetched from the NHWC bf16 activation by
// im2col-mode TMA loads (cuTensorMapEncodeIm2col), nothing is gathered into HBM first.
// ---------------------------------------------------------------------------------------------------------------
namespace {

typedef CUresult (*PEncodeIm2colFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const int*, const int*, cuuint32_t, cuuint32_t, const cuuint32_t*,
                                    CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                    CUtensorMapFloatOOBfill);
PEncodeIm2colFn p_encode_im2col = nullptr;

int p_resolve_encode_im2col() {
  std::lock_guard<std::mutex> lock(p_encode_mu);
  if (p_encode_im2col == nullptr) {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    const cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &fn, cudaEnableDefault, &q);
    if (e != cudaSuccess) return (int)e;
    if (q != cudaDriverEntryPointSuccess || fn == nullptr) return (int)cudaErrorNotSupported;
    p_encode_im2col = reinterpret_cast<PEncodeIm2colFn>(fn);
  }
  return 0;
}

// NHWC bf16 tensor [B, H, W, C] as an im2col map: boxes of `pixels` x 64 channels; the bounding box of base pixels
// is [lo, dim + hi) per spatial dimension, walked with strides (sw, sh).
int p_make_im2col_map(CUtensorMap* map, const void* ptr, int B, int H, int W, int C, int lo_w, int lo_h, int hi_w,
                      int hi_h, int sw, int sh, int pixels) {
  const int rc = p_resolve_encode_im2col();
  if (rc != 0) return rc;
  if (lo_w < -128 || lo_w > 127 || lo_h < -128 || lo_h > 127 || hi_w < -128 || hi_w > 127 || hi_h < -128 || hi_h > 127 ||
      sw < 1 || sw > 8 || sh < 1 || sh > 8 || pixels < 1 || pixels > 1024)
    return TFDE_ERR_INVALID;
  cuuint64_t gdim[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B};
  cuuint64_t gstride[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
  int lower[2] = {lo_w, lo_h}, upper[2] = {hi_w, hi_h};
  cuuint32_t estr[4] = {1, (cuuint32_t)sw, (cuuint32_t)sh, 1};
  const CUresult r = p_encode_im2col(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(ptr), gdim, gstride, lower,
                                     upper, 64, (cuuint32_t)pixels, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return (int)cudaErrorInvalidValue;
  // im2col maps of tensors smaller than 128 KB: drivers up to 13.1 set a descriptor bit that makes the load misbehave
  // (the fix-up CUTLASS applies in cute/atom/copy_traits_sm90_im2col.hpp)
  int drv = 0;
  if (cudaDriverGetVersion(&drv) == cudaSuccess && drv <= 13010 && (size_t)B * H * W * C * 2 < 131072)
    reinterpret_cast<uint64_t*>(map)[1] &= ~(1ull << 21);
  return 0;
}

bool p_conv_ok(int B, int H, int W, int C, int KH, int KW, int Ho, int Wo, int Cout) {
  return B > 0 && H > 0 && W > 0 && Ho > 0 && Wo > 0 && KH > 0 && KW > 0 && KH * KW > 1 && KH <= 16 && KW <= 16 && C > 0 &&
         Cout > 0 && (C & 63) == 0 && (Cout & 63) == 0 && (long)B * Ho * Wo < (1L << 31) && (long)B * H * W < (1L << 31);
}

}  // namespace

// 1 when the three implicit-GEMM entry points below take this convolution (dgrad additionally needs stride 1).
TFDE_API int tfde_conv_igemm_supported(int B, int H, int W, int C, int KH, int KW, int Ho, int Wo, int Cout) {
  return p_conv_ok(B, H, W, C, KH, KW, Ho, Wo, Cout) ? 1 : 0;
}

// y[B,Ho,Wo,Cout] (bf16) = conv(x[B,H,W,C] bf16, w[KH,KW,C,Cout] bf16 as stored) + bias; pads = (top, bottom, left,
// right). stat_sum / stat_sq (optional, both or none): fused batch-norm partial statistics, [ceil(M/32)][Cout].
TFDE_API int tfde_conv_igemm_fwd(const void* x, const void* w16, void* y, const float* bias, int B, int H, int W, int C,
                                 int KH, int KW, int Ho, int Wo, int sh, int sw, int pt, int pb, int pl, int pr, int Cout,
                                 float* stat_sum, float* stat_sq, float* ws, size_t ws_bytes, cudaStream_t stream) {
  if (!x || !w16 || !y || !p_conv_ok(B, H, W, C, KH, KW, Ho, Wo, Cout) || ((stat_sum == nullptr) != (stat_sq == nullptr)))
    return TFDE_ERR_INVALID;
  if ((H + pt + pb - KH) / sh + 1 != Ho || (W + pl + pr - KW) / sw + 1 != Wo) return TFDE_ERR_INVALID;
  if ((reinterpret_cast<uintptr_t>(x) & 15) || (reinterpret_cast<uintptr_t>(w16) & 15) || (reinterpret_cast<uintptr_t>(y) & 15))
    return TFDE_ERR_INVALID;
  const int M = B * Ho * Wo, N = Cout, K = KH * KW * C;
  const int BN = N <= 64 ? 64 : (N <= 128 ? 128 : 256);
  CUtensorMap ta, tb;
  // base pixels [-pad, dim + pad_far - (k-1)): exactly Ho x Wo positions at the traversal strides
  int rc = p_make_im2col_map(&ta, x, B, H, W, C, -pl, -pt, pr - (KW - 1), pb - (KH - 1), sw, sh, P_BM);
  if (rc != 0) return rc;
  rc = p_make_map_ld(&tb, w16, K, N, N, P_BK, 64);
  if (rc != 0) return rc;
  PEpilogue ep{nullptr, (__nv_bfloat16*)y, N, bias, 0, M, N, stat_sum, stat_sq};
  const PConv cv{Wo, Ho * Wo, sw, sh, pl, pt, KW, KH * KW, C / 64, C};
  if (stat_sum != nullptr) { ws = nullptr; ws_bytes = 0; }     // partial tiles have no statistics
  CUtensorMap tc;
  rc = p_prepare_store(ep, &tc);
  if (rc != 0) return rc;
  const CUtensorMap* tcp = ep.tma_store ? &tc : nullptr;
  if (BN == 64) return p_launch<64, 2, 1>(ta, tb, ep, M, N, K, ws, ws_bytes, stream, cv, tcp);
  if (BN == 128) return p_launch<128, 2, 1>(ta, tb, ep, M, N, K, ws, ws_bytes, stream, cv, tcp);
  return p_launch<256, 2, 1>(ta, tb, ep, M, N, K, ws, ws_bytes, stream, cv, tcp);
}

// dx[B,H,W,C] (bf16) of a STRIDE-1 convolution: a convolution of dy[B,Ho,Wo,Cout] with the mirrored filter and the
// complementary padding; the weights are the forward pass's bf16 copy, addressed tap-mirrored and K-major.
TFDE_API int tfde_conv_igemm_dgrad(const void* dy, const void* w16, void* dx, int B, int H, int W, int C, int KH, int KW,
                                   int Ho, int Wo, int pt, int pl, int Cout, float* ws, size_t ws_bytes,
                                   cudaStream_t stream) {
  if (!dy || !w16 || !dx || !p_conv_ok(B, H, W, C, KH, KW, Ho, Wo, Cout)) return TFDE_ERR_INVALID;
  if ((reinterpret_cast<uintptr_t>(dy) & 15) || (reinterpret_cast<uintptr_t>(w16) & 15) || (reinterpret_cast<uintptr_t>(dx) & 15))
    return TFDE_ERR_INVALID;
  const int M = B * H * W, N = C, K = KH * KW * Cout;
  const int BN = (N % 256 == 0) ? 256 : ((N % 128 == 0) ? 128 : 64);   // a B tile stays inside one tap's Cin rows
  const int lo_w = -(KW - 1 - pl), lo_h = -(KH - 1 - pt);
  CUtensorMap ta, tb;
  // W positions of the base pixel: Wo + hi - lo = W
  int rc = p_make_im2col_map(&ta, dy, B, Ho, Wo, Cout, lo_w, lo_h, W - Wo + lo_w, H - Ho + lo_h, 1, 1, P_BM);
  if (rc != 0) return rc;
  rc = p_make_map(&tb, w16, (long)KH * KW * C, Cout, BN);
  if (rc != 0) return rc;
  PEpilogue ep{nullptr, (__nv_bfloat16*)dx, N, nullptr, 0, M, N};
  const PConv cv{W, H * W, 1, 1, -lo_w, -lo_h, KW, KH * KW, Cout / 64, C};
  CUtensorMap tc;
  rc = p_prepare_store(ep, &tc);
  if (rc != 0) return rc;
  const CUtensorMap* tcp = ep.tma_store ? &tc : nullptr;
  if (BN == 64) return p_launch<64, 0, 2>(ta, tb, ep, M, N, K, ws, ws_bytes, stream, cv, tcp);
  if (BN == 128) return p_launch<128, 0, 2>(ta, tb, ep, M, N, K, ws, ws_bytes, stream, cv, tcp);
  return p_launch<256, 0, 2>(ta, tb, ep, M, N, K, ws, ws_bytes, stream, cv, tcp);
}

// dw[KH,KW,C,Cout] (fp32) = im2col(x)^T . dy, reduced over the B*Ho*Wo output pixels (split-K through ws).
TFDE_API int tfde_conv_igemm_wgrad(const void* x, const void* dy, float* dw, int B, int H, int W, int C, int KH, int KW,
                                   int Ho, int Wo, int sh, int sw, int pt, int pb, int pl, int pr, int Cout, float* ws,
                                   size_t ws_bytes, cudaStream_t stream) {
  if (!x || !dy || !dw || !p_conv_ok(B, H, W, C, KH, KW, Ho, Wo, Cout)) return TFDE_ERR_INVALID;
  if ((H + pt + pb - KH) / sh + 1 != Ho || (W + pl + pr - KW) / sw + 1 != Wo) return TFDE_ERR_INVALID;
  if ((reinterpret_cast<uintptr_t>(x) & 15) || (reinterpret_cast<uintptr_t>(dy) & 15)) return TFDE_ERR_INVALID;
  const int M = KH * KW * C, N = Cout, K = B * Ho * Wo;
  const int BN = N <= 64 ? 64 : (N <= 128 ? 128 : 256);
  CUtensorMap ta, tb;
  int rc = p_make_im2col_map(&ta, x, B, H, W, C, -pl, -pt, pr - (KW - 1), pb - (KH - 1), sw, sh, P_BK);
  if (rc != 0) return rc;
  rc = p_make_map_ld(&tb, dy, K, N, N, P_BK, 64);
  if (rc != 0) return rc;
  PEpilogue ep{dw, nullptr, N, nullptr, 0, M, N};
  const PConv cv{Wo, Ho * Wo, sw, sh, pl, pt, KW, KH * KW, 1, C};
  CUtensorMap tc;
  rc = p_prepare_store(ep, &tc);
  if (rc != 0) return rc;
  const CUtensorMap* tcp = ep.tma_store ? &tc : nullptr;
  if (BN == 64) return p_launch<64, 3, 3>(ta, tb, ep, M, N, K, ws, ws_bytes, stream, cv, tcp);
  if (BN == 128) return p_launch<128, 3, 3>(ta, tb, ep, M, N, K, ws, ws_bytes, stream, cv, tcp);
  return p_launch<256, 3, 3>(ta, tb, ep, M, N, K, ws, ws_bytes, stream, cv, tcp);
}
