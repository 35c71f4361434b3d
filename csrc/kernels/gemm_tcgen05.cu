// bf16 GEMM on the 5th-generation tensor cores: tcgen05.mma with the accumulator in TMEM, operands
// staged in shared memory by TMA (cp.async.bulk.tensor, 128-byte swizzle), warp-specialised roles.
//
//   C[M,N] = act( A[M,K] . B[N,K]^T + bias[N] )        A, B bf16 K-major ("TN"), C fp32 or bf16
//
// This is the tensor-core path for the dense layers of the reference when they are big enough to feed
// it -- Dense(64) over the 5408-wide flattened activation at large batch, BASELINE.json config 3
// ("distributed_with_keras.py model at large global batch (8192) bf16"), i.e. the G4/G5 rows of
// SURVEY.md section 2.7 (cuBLAS GEMM + Eigen bias/relu in TensorFlow: distributed_with_keras.py:37-38).
// The three GEMMs of a dense layer all map onto this one TN kernel:
//   forward  y  = x . W          A = x  [M,K],        B = W^T [N,K]   (+bias, ReLU in the epilogue)
//   dgrad    dx = dy . W^T       A = dy [M,N],        B = W   [K,N]   (Keras (in,out) layout as stored)
//   wgrad    dW = x^T . dy       A = x^T [K,M],       B = dy^T [N,M]
//
// Structure (one (128*MT) x BN output tile per CTA, 192 threads; MT = 2 for large problems: two M=128
// accumulators share every B tile, which cuts L2->SM operand traffic per FLOP by a third -- at 8192^3 the
// MT=1 kernel measured 0.90 PFLOP/s and was bound by exactly that traffic):
//   warp 0      TMA producer: one elected lane arms full[s] with expect_tx and issues two bulk tensor
//               loads per k-block into a 4-stage ring (A: 128x64 bf16 = 16 KB, B: BNx64 bf16)
//   warp 1      allocates BN TMEM columns; one elected lane issues 4 x tcgen05.mma (K=16 each) per
//               k-block and releases the stage with tcgen05.commit -> empty[s]; final commit -> tmem_full
//   warps 2..5  epilogue: tcgen05.ld (32 lanes x 32 columns per instruction) -> bias/ReLU -> global
// M/N/K tails are handled by TMA zero fill on loads and by bounds checks on stores.
#include "common.cuh"
#include <cuda.h>
#include <cuda_bf16.h>
#include <mutex>

namespace {

constexpr int TC_BM = 128;
constexpr int TC_BK = 64;          // 64 bf16 = 128 bytes = one swizzle row
constexpr int TC_STAGES = 4;
constexpr int TC_THREADS = 192;
constexpr int UMMA_K = 16;

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) { }
}
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* map, uint64_t* bar, int c_inner, int c_outer) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(map), "r"(smem_u32(bar)), "r"(c_inner), "r"(c_outer) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// K-major, SWIZZLE_128B shared-memory matrix descriptor (cute::UMMA::SmemDescriptor bit layout):
// [0,14) start address >> 4 | [16,30) leading byte offset >> 4 (unused for swizzled K-major, 1) |
// [32,46) stride byte offset >> 4 (8 rows x 128 B = 1024 B) | [46,48) version = 1 | [61,64) layout = 2.
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t smem_addr) {
  return (uint64_t)((smem_addr & 0x3FFFFu) >> 4) | (1ull << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) |
         (2ull << 61);
}
// kind::f16 instruction descriptor: fp32 accumulate, bf16 A/B, both K-major, N>>3 at [17,23), M>>4 at [24,29).
__device__ __forceinline__ uint32_t make_instr_desc(int umma_m, int umma_n) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(umma_n >> 3) << 17) | ((uint32_t)(umma_m >> 4) << 24);
}
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_ld_32x32(uint32_t taddr, uint32_t (&r)[32]) {
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

struct TcEpilogue {
  float* c_f32; __nv_bfloat16* c_bf16; long ldc;
  const float* bias; int relu;
  int M, N;
};

template <int BN, int MT>
__global__ void __launch_bounds__(TC_THREADS, 1)
gemm_bf16_tcgen05_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                         const TcEpilogue ep, const int K) {
  constexpr uint32_t A_SUB_BYTES = TC_BM * TC_BK * 2;   // 16 KB per 128-row sub-tile
  constexpr uint32_t A_BYTES = A_SUB_BYTES * MT;
  constexpr uint32_t B_BYTES = BN * TC_BK * 2;
  constexpr uint32_t STAGE_BYTES = A_BYTES + B_BYTES;
  constexpr int STAGES = (MT == 2) ? 3 : TC_STAGES;
  constexpr uint32_t TMEM_COLS = (BN * MT <= 32) ? 32 : (BN * MT <= 64 ? 64 : (BN * MT <= 128 ? 128 : (BN * MT <= 256 ? 256 : 512)));
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // 1024-byte alignment is required by the 128-byte swizzle atoms (base_offset field left at 0)
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  __shared__ __align__(8) uint64_t full_bar[STAGES];
  __shared__ __align__(8) uint64_t empty_bar[STAGES];
  __shared__ __align__(8) uint64_t tmem_full_bar;
  __shared__ uint32_t tmem_base_smem;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m0 = blockIdx.y * (TC_BM * MT), n0 = blockIdx.x * BN;
  const int num_k = (K + TC_BK - 1) / TC_BK;

  if (threadIdx.x == 0) {
#pragma unroll
    for (int s = 0; s < STAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    mbar_init(&tmem_full_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmA) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmB) : "memory");
  }
  if (warp == 1) {   // whole warp: allocate BN TMEM columns (power of two >= 32), publish the base address
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_smem)), "r"(TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_d = tmem_base_smem;

  if (warp == 0) {
    if (lane == 0) {
      for (int kb = 0; kb < num_k; ++kb) {
        const int s = kb % STAGES;
        const uint32_t ph = (uint32_t)(kb / STAGES) & 1u;
        mbar_wait(&empty_bar[s], ph ^ 1u);            // slot free (fresh barrier: passes immediately)
        mbar_expect_tx(&full_bar[s], STAGE_BYTES);
        uint8_t* a_dst = smem + (size_t)s * STAGE_BYTES;
        tma_load_2d(a_dst, &tmA, &full_bar[s], kb * TC_BK, m0);
        tma_load_2d(a_dst + A_BYTES, &tmB, &full_bar[s], kb * TC_BK, n0);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc = make_instr_desc(TC_BM, BN);
      for (int kb = 0; kb < num_k; ++kb) {
        const int s = kb % STAGES;
        const uint32_t ph = (uint32_t)(kb / STAGES) & 1u;
        mbar_wait(&full_bar[s], ph);                  // TMA bytes have landed
        tc_fence_after();
        const uint32_t a_addr = smem_u32(smem + (size_t)s * STAGE_BYTES);
        const uint32_t b_addr = a_addr + A_BYTES;
#pragma unroll
        for (int mt = 0; mt < MT; ++mt) {
#pragma unroll
          for (int k = 0; k < TC_BK / UMMA_K; ++k) {
            // advancing K by 16 bf16 = 32 bytes inside the 128-byte swizzle row: bump the start address
            const uint64_t da = make_smem_desc(a_addr + mt * A_SUB_BYTES + k * UMMA_K * 2);
            const uint64_t db = make_smem_desc(b_addr + k * UMMA_K * 2);
            umma_bf16(tmem_d + (uint32_t)(mt * BN), da, db, idesc, (kb | k) != 0 ? 1u : 0u);
          }
        }
        umma_commit(&empty_bar[s]);                   // stage reusable once these MMAs have read it
      }
      umma_commit(&tmem_full_bar);                    // accumulator complete
    }
  } else {
    // epilogue warps 2..5: TMEM lane quarter = warp % 4. Once tmem_full fires every MMA has finished
    // reading shared memory, so the pipeline stages are free: each warp transposes its 32x32 chunk through
    // a private padded tile there and stores full 128-byte (fp32) / 64-byte (bf16) row segments.
    mbar_wait(&tmem_full_bar, 0);
    tc_fence_after();
    const int q = warp & 3;
#pragma unroll 1
    for (int mt = 0; mt < MT; ++mt) {
      const int row0 = m0 + mt * TC_BM + q * 32;
      // Staged path (one output dtype, 16-byte aligned rows): each warp bounces a 32-row x 128-byte slab
      // through a private padded tile in the (now idle) pipeline shared memory and stores it with 8 lanes per
      // row, so every store instruction covers 4 complete 128-byte lines instead of 32 half-sectors
      // (see gemm_tcgen05_persistent.cu; profiles/ncu_resnet_gemm_direct_epilogue.json for the measurement).
      int c_done = 0;
      if (((ep.c_f32 == nullptr) != (ep.c_bf16 == nullptr)) && (ep.ldc & 7) == 0) {
        constexpr int PITCH = 144;
        uint8_t* stg = smem + q * (32 * PITCH);
        const bool out16 = ep.c_bf16 != nullptr;
        const int pass = out16 ? 64 : 32;
#pragma unroll 1
        for (int c0 = 0; c0 + pass <= BN && n0 + c0 + pass <= ep.N; c0 += pass) {
          const int nbase = n0 + c0;
#pragma unroll 1
          for (int h = 0; h * 32 < pass; ++h) {
            uint32_t r[32];
            tmem_ld_32x32(tmem_d + ((uint32_t)(q * 32) << 16) + (uint32_t)(mt * BN + c0 + h * 32), r);
            float v[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              float x = __uint_as_float(r[j]);
              if (ep.bias != nullptr) x += __ldg(ep.bias + nbase + h * 32 + j);
              v[j] = ep.relu ? fmaxf(x, 0.f) : x;
            }
            uint4* srow = reinterpret_cast<uint4*>(stg + lane * PITCH);
            if (out16) {
#pragma unroll
              for (int j = 0; j < 32; j += 8) {
                uint4 o;
                __nv_bfloat162 p0 = __floats2bfloat162_rn(v[j], v[j + 1]), p1 = __floats2bfloat162_rn(v[j + 2], v[j + 3]);
                __nv_bfloat162 p2 = __floats2bfloat162_rn(v[j + 4], v[j + 5]), p3 = __floats2bfloat162_rn(v[j + 6], v[j + 7]);
                o.x = *reinterpret_cast<uint32_t*>(&p0); o.y = *reinterpret_cast<uint32_t*>(&p1);
                o.z = *reinterpret_cast<uint32_t*>(&p2); o.w = *reinterpret_cast<uint32_t*>(&p3);
                srow[h * 4 + (j >> 3)] = o;
              }
            } else {
#pragma unroll
              for (int j = 0; j < 32; j += 4)
                srow[j >> 2] = make_uint4(__float_as_uint(v[j]), __float_as_uint(v[j + 1]), __float_as_uint(v[j + 2]),
                                          __float_as_uint(v[j + 3]));
            }
          }
          __syncwarp();
          uint8_t* gbase = out16 ? reinterpret_cast<uint8_t*>(ep.c_bf16 + nbase) : reinterpret_cast<uint8_t*>(ep.c_f32 + nbase);
          const size_t gpitch = (size_t)ep.ldc * (out16 ? 2 : 4);
#pragma unroll
          for (int it = 0; it < 8; ++it) {
            const int rr = it * 4 + (lane >> 3), piece = lane & 7;
            const uint4 val = *reinterpret_cast<const uint4*>(stg + rr * PITCH + piece * 16);
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
        tmem_ld_32x32(tmem_d + ((uint32_t)(q * 32) << 16) + (uint32_t)(mt * BN + c0), r);
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
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"(TMEM_COLS) : "memory");
  }
}

// ---- elementwise helpers for the bf16 path ------------------------------------------------------
__global__ void cast_f32_to_bf16_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ y, long n) {
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x)
    y[i] = __float2bfloat16(__ldg(x + i));
}
// y[C,R] (bf16) = transpose(x[R,C] (fp32)), optionally masked by (mask[R,C] > 0): 32x32 tiles via smem.
__global__ void cast_transpose_bf16_kernel(const float* __restrict__ x, const float* __restrict__ mask,
                                           __nv_bfloat16* __restrict__ y, int R, int C) {
  __shared__ float tile[32][33];
  const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int r = r0 + i, c = c0 + threadIdx.x;
    float v = 0.f;
    if (r < R && c < C) {
      v = __ldg(x + (size_t)r * C + c);
      if (mask != nullptr && !(__ldg(mask + (size_t)r * C + c) > 0.f)) v = 0.f;
    }
    tile[i][threadIdx.x] = v;
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int c = c0 + i, r = r0 + threadIdx.x;
    if (c < C && r < R) y[(size_t)c * R + r] = __float2bfloat16(tile[threadIdx.x][i]);
  }
}

// ---- host: tensor maps through the driver entry point (no link-time dependency on libcuda) ---------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn g_encode = nullptr;
std::mutex g_encode_mu;

int get_encode(EncodeTiledFn* out) {
  std::lock_guard<std::mutex> lock(g_encode_mu);
  if (g_encode == nullptr) {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
    if (e != cudaSuccess) return (int)e;
    if (qres != cudaDriverEntryPointSuccess || fn == nullptr) return (int)cudaErrorNotSupported;
    g_encode = reinterpret_cast<EncodeTiledFn>(fn);
  }
  *out = g_encode;
  return 0;
}

// 2-D bf16 tensor [rows, cols] (cols contiguous), box = [box_rows, 64 cols], 128-byte swizzle.
int make_map(CUtensorMap* map, const void* ptr, long rows, long cols, int box_rows) {
  EncodeTiledFn enc;
  const int rc = get_encode(&enc);
  if (rc != 0) return rc;
  cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t gstride[1] = {(cuuint64_t)cols * 2};
  cuuint32_t box[2] = {(cuuint32_t)TC_BK, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  const CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), gdim, gstride, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : (int)cudaErrorInvalidValue;
}

template <int BN, int MT>
int launch_tc(const CUtensorMap& ta, const CUtensorMap& tb, const TcEpilogue& ep, int M, int N, int K, cudaStream_t stream) {
  constexpr int STAGES = (MT == 2) ? 3 : TC_STAGES;
  constexpr size_t smem = (size_t)STAGES * (MT * TC_BM * TC_BK * 2 + BN * TC_BK * 2) + 1024;
  static size_t configured[TFDE_MAX_DEVICES] = {};
  { const cudaError_t e = tfde_ensure_dyn_smem(gemm_bf16_tcgen05_kernel<BN, MT>, smem, configured); if (e != cudaSuccess) return (int)e; }
  dim3 grid((N + BN - 1) / BN, (M + TC_BM * MT - 1) / (TC_BM * MT));
  gemm_bf16_tcgen05_kernel<BN, MT><<<grid, TC_THREADS, smem, stream>>>(ta, tb, ep, K);
  TFDE_RETURN_LAUNCH_STATUS();
}

}  // namespace

// C[M,N] = act(A[M,K] . B[N,K]^T + bias). A, B: bf16, row-major with K contiguous, 16-byte aligned,
// K % 8 == 0 (TMA global strides are multiples of 16 bytes). Exactly one of c_f32 / c_bf16 may be NULL.
// The first call per tile shape sets a function attribute: make it outside of stream capture.
TFDE_API int tfde_gemm_bf16_tn(const void* A, const void* B, float* c_f32, void* c_bf16, long ldc,
                               const float* bias, int relu, int M, int N, int K, cudaStream_t stream) {
  if (M <= 0 || N <= 0 || K <= 0 || (K & 7) || !A || !B || (!c_f32 && !c_bf16)) return TFDE_ERR_INVALID;
  if ((reinterpret_cast<uintptr_t>(A) & 15) || (reinterpret_cast<uintptr_t>(B) & 15)) return TFDE_ERR_INVALID;
  const int BN = N <= 64 ? 64 : (N <= 128 ? 128 : 256);
  // two stacked M=128 accumulators per CTA when the grid still fills the machine (>= 2 waves of 148 SMs)
  const long tiles256 = (long)((M + 255) / 256) * ((N + BN - 1) / BN);
  const int MT = (BN == 256 && tiles256 >= 2 * 148) ? 2 : 1;
  CUtensorMap ta, tb;
  int rc = make_map(&ta, A, M, K, TC_BM * MT);
  if (rc != 0) return rc;
  rc = make_map(&tb, B, N, K, BN);
  if (rc != 0) return rc;
  TcEpilogue ep{c_f32, (__nv_bfloat16*)c_bf16, ldc, bias, relu, M, N};
  if (BN == 64) return launch_tc<64, 1>(ta, tb, ep, M, N, K, stream);
  if (BN == 128) return launch_tc<128, 1>(ta, tb, ep, M, N, K, stream);
  if (MT == 2) return launch_tc<256, 2>(ta, tb, ep, M, N, K, stream);
  return launch_tc<256, 1>(ta, tb, ep, M, N, K, stream);
}

TFDE_API int tfde_cast_bf16(const float* x, void* y, long n, cudaStream_t stream) {
  long grid = (n + 255) / 256;
  if (grid > 148L * 8) grid = 148L * 8;
  if (grid < 1) grid = 1;
  cast_f32_to_bf16_kernel<<<(int)grid, 256, 0, stream>>>(x, (__nv_bfloat16*)y, n);
  TFDE_RETURN_LAUNCH_STATUS();
}

// y[C,R] bf16 = transpose(x[R,C] fp32 * (mask > 0))
TFDE_API int tfde_cast_transpose_bf16(const float* x, const float* mask, void* y, int R, int C, cudaStream_t stream) {
  dim3 grid((C + 31) / 32, (R + 31) / 32), block(32, 8);
  cast_transpose_bf16_kernel<<<grid, block, 0, stream>>>(x, mask, (__nv_bfloat16*)y, R, C);
  TFDE_RETURN_LAUNCH_STATUS();
}
