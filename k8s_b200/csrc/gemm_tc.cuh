// tcgen05 GEMM for sm_100a:  C[M,N] (fp32) = A(M x K) * B(N x K)^T with bf16 (kind::f16) or
// tf32 (kind::tf32) operands and fp32 accumulation in TMEM.  Each operand may be stored
// K-major ([rows, K], K contiguous) or MN-major ([K, rows], rows contiguous), so every
// product of two row-major matrices, transposed or not, runs without a transposed copy.
//
//   * operands reach shared memory by TMA (cp.async.bulk.tensor.2d, 128-byte swizzle,
//     out-of-bounds rows/columns zero-filled, so M, N and K need no padding);
//   * one elected thread issues tcgen05.mma (M = 128, N = BLOCK_N, 32 bytes of K per
//     instruction) over a 4-stage mbarrier ring; tcgen05.commit frees each stage;
//   * the accumulator (128 lanes x BLOCK_N columns) is read back with tcgen05.ld by four
//     epilogue warps (one TMEM lane quarter each), which apply bias / relu / mask and store.
//   * gridDim.z > 1 = split-K: split z handles a contiguous range of K blocks and writes
//     its partial product to C + z*M*ldc... (dense [z][M][N]); a fixed-order reduce follows.
//
// Warp roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM allocator + MMA issuer,
// warps 2..5 = epilogue (warp w owns TMEM lanes 32*(w%4) .. +31).
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace tfg {
namespace tc {

constexpr int kBlockM = 128;
constexpr int kBlockKBytes = 128;      // one 128-B swizzle row of K per stage row
constexpr int kMaxStages = 4;
constexpr int kThreads = 192;
constexpr int kStatsThreads = 256;
constexpr int kThreadsPersistent = kThreads + kStatsThreads;   // + eight warps that take batch-norm column sums off the epilogue's path
constexpr long long kWaitLimitCycles = 4000000000ll;   // ~2 s: a wait that long is a bug -> trap, never hang

struct Epilogue {
  const float* bias;    // [N] added to every row, or NULL
  const float* mask;    // [M, ldc]: keep the value where mask > 0, else 0; or NULL
  int relu;
  void* out_bf16;       // when set, the result is stored as bf16 [M, ldc] here instead of fp32 in C
  int direct;           // 1: store straight from registers (thread = row), 0: via the smem transpose
  int tma_out;          // persistent kernels, bf16 output through the kernel's output tensor map:
                        // 1 = TMA store of the staged tile, 2 = TMA reduce-add into the destination
  float* col_stats;     // with tma_out: per-tile column sums of the bf16 output and of its squares,
                        // [m_tile * parts + slice][2][N] (parts = 256 / BLOCK_N row slices), or NULL
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// Low-order plane of the 3xTF32 split (gemm_x3.cuh): kind::tf32 reads an fp32 word's top 19 bits
// (truncation), so hi = x with the low 13 mantissa bits cleared, x - hi is exact, and lo is that
// remainder rounded to tf32 (round-to-nearest) so that the hardware's truncation of lo loses nothing.
__device__ __forceinline__ float tf32_lo(float x) {
  const float hi = __uint_as_float(__float_as_uint(x) & 0xffffe000u);
  uint32_t o;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(o) : "f"(x - hi));
  return __uint_as_float(o);
}

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count) : "memory");
}

__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;"
               :: "r"(smem_u32(bar)), "r"(bytes) : "memory");
}

// Wait with a suspend-time hint: the warp sleeps in hardware until the phase completes (or the
// hint expires) instead of spinning.  A CTA here holds up to 14 warps of which ONE thread issues
// all the MMAs; twelve warps polling without the hint took three quarters of that thread's issue
// slots (tensor pipe 40 % active with every operand already in shared memory).
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  const uint32_t addr = smem_u32(bar);
  long long t0 = 0;
  for (;;) {
    uint32_t done;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done) : "r"(addr), "r"(parity), "r"(1000000u) : "memory");
    if (done) return;
    if (t0 == 0) t0 = clock64();
    else if (clock64() - t0 > kWaitLimitCycles) __trap();
  }
}

__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* map, uint64_t* bar,
                                            int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      :: "r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}

// Shared-memory matrix descriptor, 128-byte swizzle, version 1 (Blackwell), layout type 2.
//   K-major : a tile row (128 B of K) every 128 B, 8-row atoms 1024 B apart = stride byte
//             offset; the leading byte offset is not used.
//   MN-major: a 128-B line holds 64 bf16 / 32 tf32 consecutive rows of ONE k; 8 consecutive
//             k form a 1024-B atom (stride byte offset = 1024); the next 128 B worth of rows
//             is a separate TMA box `lbo_bytes` further on (leading byte offset).
//   MN-major with 32-bit elements (tf32) uses the "128B swizzle, 32-byte base" layout
//             (type 1): 32-B chunks are swizzled, 4 consecutive k form a 512-B atom.
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t smem_addr, uint32_t lbo_bytes,
                                                   uint32_t sbo_bytes = 1024, uint32_t layout_type = 2) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFFu) >> 4);        // start address   [0,14)
  d |= static_cast<uint64_t>((lbo_bytes >> 4) & 0x3FFFu) << 16;   // leading byte offset [16,30)
  d |= static_cast<uint64_t>(sbo_bytes >> 4) << 32;               // stride byte offset [32,46)
  d |= static_cast<uint64_t>(1) << 46;                            // descriptor version [46,48)
  d |= static_cast<uint64_t>(layout_type) << 61;                  // 2 = SWIZZLE_128B, 1 = SWIZZLE_128B_BASE32B [61,64)
  return d;
}

// Instruction descriptor: D = F32, A/B format `fmt` (1 = BF16, 2 = TF32), operand majors at
// bits 15/16 (0 = K-major, 1 = MN-major), N >> 3 at bit 17, M >> 4 at bit 24.
__host__ __device__ constexpr uint32_t make_idesc(int fmt, int n, bool a_mn, bool b_mn) {
  return (1u << 4) | (static_cast<uint32_t>(fmt) << 7) | (static_cast<uint32_t>(fmt) << 10) |
         (a_mn ? (1u << 15) : 0u) | (b_mn ? (1u << 16) : 0u) |
         (static_cast<uint32_t>(n >> 3) << 17) | (static_cast<uint32_t>(kBlockM >> 4) << 24);
}

template <int KIND>  // 0 = bf16 (kind::f16), 1 = tf32
__device__ __forceinline__ void umma(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                     uint32_t accumulate) {
  if (KIND == 0) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        :: "r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
  } else {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        :: "r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
  }
}

// ---- lean issue path -----------------------------------------------------------------------------
// ONE thread issues every MMA of a CTA, so the instructions between two tcgen05.mma set the pace of
// the tensor pipe (measured: ~100 instructions per K block held it at 40 %).  The MMA warp therefore
// runs converged (values stay warp-uniform), waits as a warp, and elects a lane only for the issue;
// descriptors are kept as a constant high word and a 32-bit low word that is bumped by constants.
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
  return pred != 0;
}
// low word: start address >> 4 in [0,14), leading byte offset >> 4 in [16,30)
__device__ __forceinline__ uint32_t desc_lo(uint32_t smem_addr, uint32_t lbo_bytes) {
  return ((smem_addr & 0x3FFFFu) >> 4) | (((lbo_bytes >> 4) & 0x3FFFu) << 16);
}
// high word: stride byte offset >> 4 in [0,14), version 1 at bit 14, layout type at bit 29
__host__ __device__ constexpr uint32_t desc_hi(uint32_t sbo_bytes = 1024, uint32_t layout_type = 2) {
  return (sbo_bytes >> 4) | (1u << 14) | (layout_type << 29);
}
template <int KIND>
__device__ __forceinline__ void umma2(uint32_t tmem_d, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo, uint32_t b_hi,
                                      uint32_t idesc, bool accumulate) {
  if (KIND == 0) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\tsetp.ne.b32 p, %6, 0;\n\t"
        "mov.b64 da, {%1, %2};\n\tmov.b64 db, {%3, %4};\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %5, p;\n\t}"
        :: "r"(tmem_d), "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(accumulate ? 1u : 0u) : "memory");
  } else {
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\tsetp.ne.b32 p, %6, 0;\n\t"
        "mov.b64 da, {%1, %2};\n\tmov.b64 db, {%3, %4};\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], da, db, %5, p;\n\t}"
        :: "r"(tmem_d), "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(accumulate ? 1u : 0u) : "memory");
  }
}
__device__ __forceinline__ void umma_commit_addr(uint32_t bar_addr) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(bar_addr) : "memory");
}
// mbar_wait on a shared-memory address (no generic-to-shared conversion in the loop)
__device__ __forceinline__ void mbar_wait_addr(uint32_t addr, uint32_t parity) {
  long long t0 = 0;
  for (;;) {
    uint32_t done;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done) : "r"(addr), "r"(parity), "r"(1000000u) : "memory");
    if (done) return;
    if (t0 == 0) t0 = clock64();
    else if (clock64() - t0 > kWaitLimitCycles) __trap();
  }
}

__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];"
               :: "r"(smem_u32(bar)) : "memory");
}

template <int BLOCK_N, int STAGES>
constexpr size_t smem_bytes() {
  return 1024 /*alignment slack*/ + static_cast<size_t>(STAGES) * (kBlockM * kBlockKBytes + BLOCK_N * kBlockKBytes) +
         (2 * STAGES + 1) * sizeof(uint64_t) + 16 + 4 * 32 * 33 * sizeof(float);
}

// Epilogue of one 128 x BLOCK_N tile by the four epilogue warps.  tcgen05.ld hands thread t row t
// of its lane quarter and 32 consecutive columns per pass.  Default (epi.direct): each thread
// stores its 64 B (bf16) / 128 B (fp32) row piece straight from registers with 16-byte stores --
// measured faster than the alternative kept for ragged tails, which bounces 32x32 blocks through a
// per-warp shared-memory tile so that 8 lanes write one 128-byte row segment.
// `have_acc` = the accumulator holds a product (false: an empty K range writes zeros).
template <int BLOCK_N>
__device__ __forceinline__ void drain_tile(uint32_t tmem_acc, bool have_acc, float* stage_base, int warp, int lane,
                                           float* __restrict__ C, int M, int N, int ldc, int m0, int n0, int z,
                                           const Epilogue& epi) {
  const int q = warp & 3;                            // TMEM lane quarter this warp may touch
  float (*stage)[33] = reinterpret_cast<float (*)[33]>(stage_base + (warp - 2) * 32 * 33);
  float* cbase = C + static_cast<size_t>(z) * M * static_cast<size_t>(ldc);
  // 128-bit accesses need 16-B aligned rows (C, bias and mask come from cudaMalloc / 128-B
  // aligned variable offsets; the leading dimension decides the rest)
  const bool vec_ok = (ldc & 3) == 0 && (reinterpret_cast<uintptr_t>(C) & 15) == 0 &&
                      (!epi.bias || (reinterpret_cast<uintptr_t>(epi.bias) & 15) == 0) &&
                      (!epi.mask || (reinterpret_cast<uintptr_t>(epi.mask) & 15) == 0);
#pragma unroll 1
  for (int c0 = 0; c0 < BLOCK_N; c0 += 32) {
    uint32_t r[32];
    if (have_acc) {
      const uint32_t taddr = tmem_acc + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(c0);
      asm volatile(
          "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
          "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
          : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
            "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
          : "r"(taddr) : "memory");
      asm volatile(
          "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
          "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
          : "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]),
            "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]),
            "=r"(r[30]), "=r"(r[31])
          : "r"(taddr + 16) : "memory");
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    } else {
#pragma unroll
      for (int u = 0; u < 32; ++u) r[u] = 0u;
    }
    if (epi.direct && vec_ok && n0 + c0 + 32 <= N && (!epi.out_bf16 || (ldc & 7) == 0)) {
      // thread t owns row t of the quarter and 32 consecutive columns: 64 B (bf16) or 128 B (fp32)
      const int row = m0 + q * 32 + lane;
      if (row < M) {
        const int colb = n0 + c0;
#pragma unroll
        for (int u = 0; u < 32; u += 8) {
          float v[8];
#pragma unroll
          for (int k = 0; k < 8; ++k) v[k] = __uint_as_float(r[u + k]);
          if (epi.bias) {
#pragma unroll
            for (int k = 0; k < 8; ++k) v[k] += epi.bias[colb + u + k];
          }
          if (epi.relu) {
#pragma unroll
            for (int k = 0; k < 8; ++k) v[k] = fmaxf(v[k], 0.f);
          }
          if (epi.mask) {
#pragma unroll
            for (int k = 0; k < 8; ++k)
              v[k] = epi.mask[static_cast<size_t>(row) * ldc + colb + u + k] > 0.f ? v[k] : 0.f;
          }
          if (epi.out_bf16) {
            uint4 pk;
            __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
            for (int k = 0; k < 4; ++k) h[k] = __floats2bfloat162_rn(v[2 * k], v[2 * k + 1]);
            *reinterpret_cast<uint4*>(static_cast<__nv_bfloat16*>(epi.out_bf16) +
                                      (static_cast<size_t>(z) * M + row) * static_cast<size_t>(ldc) + colb + u) = pk;
          } else {
            float* dst = cbase + static_cast<size_t>(row) * ldc + colb + u;
            *reinterpret_cast<float4*>(dst) = make_float4(v[0], v[1], v[2], v[3]);
            *reinterpret_cast<float4*>(dst + 4) = make_float4(v[4], v[5], v[6], v[7]);
          }
        }
      }
      continue;
    }
#pragma unroll
    for (int u = 0; u < 32; ++u) stage[lane][u] = __uint_as_float(r[u]);   // bank (lane + u) % 32
    __syncwarp();
    const int cc = (lane & 7) * 4;
    const int col = n0 + c0 + cc;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int rr = i * 4 + (lane >> 3);
      const int row = m0 + q * 32 + rr;
      float4 v = make_float4(stage[rr][cc], stage[rr][cc + 1], stage[rr][cc + 2], stage[rr][cc + 3]);
      if (row >= M || col >= N) continue;
      float* dst = cbase + static_cast<size_t>(row) * ldc + col;
      if (vec_ok && col + 3 < N) {
        if (epi.bias) {
          const float4 b = *reinterpret_cast<const float4*>(epi.bias + col);
          v.x += b.x; v.y += b.y; v.z += b.z; v.w += b.w;
        }
        if (epi.relu) { v.x = fmaxf(v.x, 0.f); v.y = fmaxf(v.y, 0.f); v.z = fmaxf(v.z, 0.f); v.w = fmaxf(v.w, 0.f); }
        if (epi.mask) {
          const float4 k = *reinterpret_cast<const float4*>(epi.mask + static_cast<size_t>(row) * ldc + col);
          v.x = k.x > 0.f ? v.x : 0.f; v.y = k.y > 0.f ? v.y : 0.f;
          v.z = k.z > 0.f ? v.z : 0.f; v.w = k.w > 0.f ? v.w : 0.f;
        }
        if (epi.out_bf16) {
          const __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
          uint2 pk;
          pk.x = *reinterpret_cast<const uint32_t*>(&lo);
          pk.y = *reinterpret_cast<const uint32_t*>(&hi);
          *reinterpret_cast<uint2*>(static_cast<__nv_bfloat16*>(epi.out_bf16) +
                                    (static_cast<size_t>(z) * M + row) * static_cast<size_t>(ldc) + col) = pk;
        } else {
          *reinterpret_cast<float4*>(dst) = v;
        }
      } else {
        const float vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          if (col + u >= N) break;
          float x = vv[u];
          if (epi.bias) x += epi.bias[col + u];
          if (epi.relu) x = fmaxf(x, 0.f);
          if (epi.mask) x = epi.mask[static_cast<size_t>(row) * ldc + col + u] > 0.f ? x : 0.f;
          if (epi.out_bf16)
            static_cast<__nv_bfloat16*>(epi.out_bf16)[(static_cast<size_t>(z) * M + row) * static_cast<size_t>(ldc) +
                                                      col + u] = __float2bfloat16_rn(x);
          else
            dst[u] = x;
        }
      }
    }
    __syncwarp();
  }
}

// STAGES = depth of the TMA ring.  Short-K GEMMs (one or two K blocks per CTA) take 1 or 2
// stages so that several CTAs fit on an SM (shared memory and TMEM columns permitting) and one
// CTA's epilogue overlaps its neighbours' loads.
template <int BLOCK_N, int KIND, bool A_MN, bool B_MN, int STAGES>
__global__ void __launch_bounds__(kThreads)
gemm_tc_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
               float* __restrict__ C, int M, int N, int ldc, int k_blocks, int k_blocks_per_split,
               Epilogue epi) {
  constexpr int kElemBytes = KIND == 0 ? 2 : 4;
  constexpr int kBlockK = kBlockKBytes / kElemBytes;           // elements of K per stage
  constexpr int kABytes = kBlockM * kBlockKBytes, kBBytes = BLOCK_N * kBlockKBytes;
  constexpr int kMnElems = 128 / kElemBytes;                   // rows per 128-B line of an MN-major box
  constexpr int kChunkBytes = kBlockK * 128;                   // one MN-major box: kBlockK lines of 128 B
  constexpr int kUmmaK = 32 / kElemBytes;                      // K per instruction: 16 (bf16), 8 (tf32)
  // descriptor start-address step (16-B units) per instruction along K
  constexpr uint32_t kStepA = A_MN ? (kUmmaK * 128) >> 4 : 32 >> 4;
  constexpr uint32_t kStepB = B_MN ? (kUmmaK * 128) >> 4 : 32 >> 4;
  static_assert(BLOCK_N == 64 || BLOCK_N == 128, "TMEM columns must be a power of two >= 32");
  constexpr int kStages = STAGES;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + kStages * kABytes;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + kStages * (kABytes + kBBytes));
  uint64_t* empty = full + kStages;
  uint64_t* tmem_full = empty + kStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);
  float* stage_base = reinterpret_cast<float*>(tmem_slot + 4);   // 4 epilogue warps x [32][33] floats

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m0 = blockIdx.y * kBlockM, n0 = blockIdx.x * BLOCK_N;
  const int kb_begin = blockIdx.z * k_blocks_per_split;
  const int kb_end = min(k_blocks, kb_begin + k_blocks_per_split);
  const int n_kb = max(kb_end - kb_begin, 0);

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" :: "l"(reinterpret_cast<uint64_t>(&map_a)) : "memory");
    asm volatile("prefetch.tensormap [%0];" :: "l"(reinterpret_cast<uint64_t>(&map_b)) : "memory");
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int s = 0; s < kStages; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
      mbar_init(tmem_full, 1);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                 :: "r"(smem_u32(tmem_slot)), "n"(BLOCK_N) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      for (int i = 0; i < n_kb; ++i) {
        const int s = i % kStages;
        const uint32_t phase = (i / kStages) & 1;
        mbar_wait(&empty[s], phase ^ 1);
        mbar_expect_tx(&full[s], kABytes + kBBytes);
        const int k_elem = (kb_begin + i) * kBlockK;
        if (!A_MN) {
          tma_load_2d(smem_a + s * kABytes, &map_a, &full[s], k_elem, m0);
        } else {
#pragma unroll
          for (int c = 0; c < kBlockM / kMnElems; ++c)
            tma_load_2d(smem_a + s * kABytes + c * kChunkBytes, &map_a, &full[s], m0 + c * kMnElems, k_elem);
        }
        if (!B_MN) {
          tma_load_2d(smem_b + s * kBBytes, &map_b, &full[s], k_elem, n0);
        } else {
#pragma unroll
          for (int c = 0; c < BLOCK_N / kMnElems; ++c)
            tma_load_2d(smem_b + s * kBBytes + c * kChunkBytes, &map_b, &full[s], n0 + c * kMnElems, k_elem);
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: converged warp, one elected lane issues ("lean issue path") =====
    constexpr uint32_t idesc = make_idesc(KIND == 0 ? 1 : 2, BLOCK_N, A_MN, B_MN);
    constexpr bool kBase32 = KIND == 1;   // MN-major tf32: 32-byte-base swizzle, 4-line atoms
    constexpr uint32_t hi_a = desc_hi((A_MN && kBase32) ? 512 : 1024, (A_MN && kBase32) ? 1 : 2);
    constexpr uint32_t hi_b = desc_hi((B_MN && kBase32) ? 512 : 1024, (B_MN && kBase32) ? 1 : 2);
    const uint32_t a0 = smem_u32(smem_a), b0 = smem_u32(smem_b);
    const uint32_t full0 = smem_u32(full), empty0 = smem_u32(empty);
    uint32_t st = 0, ph = 0;
    for (int i = 0; i < n_kb; ++i) {
      mbar_wait_addr(full0 + st * 8, ph);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const uint32_t a_lo = desc_lo(a0 + st * kABytes, A_MN ? kChunkBytes : 16);
      const uint32_t b_lo = desc_lo(b0 + st * kBBytes, B_MN ? kChunkBytes : 16);
      if (elect_one()) {
#pragma unroll
        for (int k = 0; k < kBlockK / kUmmaK; ++k)    // 32 bytes of K per instruction
          umma2<KIND>(tmem_base, a_lo + kStepA * k, hi_a, b_lo + kStepB * k, hi_b, idesc, (i | k) != 0);
        umma_commit_addr(empty0 + st * 8);             // frees the stage when these MMAs retire
      }
      __syncwarp();
      if (++st == kStages) { st = 0; ph ^= 1; }
    }
    if (elect_one()) umma_commit_addr(smem_u32(tmem_full));   // accumulator complete
    __syncwarp();
  } else {
    // ===== epilogue =====
    if (n_kb > 0) {
      mbar_wait(tmem_full, 0);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    }
    drain_tile<BLOCK_N>(tmem_base, n_kb > 0, stage_base, warp, lane, C, M, N, ldc, m0, n0, blockIdx.z, epi);
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem_base), "n"(BLOCK_N) : "memory");
  }
}

// ---- bf16 output through TMA -----------------------------------------------------------------------
// The register-direct stores write 16 B per lane into 32 different rows: one L1 wavefront and one
// L2 request per lane, which caps the output-heavy GEMMs (1x1 convolutions that widen the channel
// count) at about 40 % of HBM bandwidth.  Here the four epilogue warps pack the tile as bf16 into a
// 128-byte-swizzled shared-memory image (row r, 16-byte piece j at r * 128 + ((j ^ (r & 7)) << 4):
// conflict-free for thread = row) and ONE thread hands each 64-column half to the TMA unit, which
// writes whole lines, clips rows / columns outside the tensor and can add into the destination
// (bf16 add, one rounding -- what a separate accumulate kernel would do).
constexpr int kOutHalfBytes = kBlockM * 128;   // 128 rows x 64 bf16

__device__ __forceinline__ void epi_barrier() { asm volatile("bar.sync 1, 128;" ::: "memory"); }
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void bulk_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" :: "n"(N) : "memory");
}
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

__device__ __forceinline__ void tma_store_2d(const CUtensorMap* map, const void* smem_src, int c0, int c1, bool add) {
  if (add)
    asm volatile("cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3}], [%1];"
                 :: "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1) : "memory");
  else
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.tile.bulk_group [%0, {%2, %3}], [%1];"
                 :: "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tma_store_4d(const CUtensorMap* map, const void* smem_src, int c0, int c1, int c2,
                                             int c3, bool add) {
  if (add)
    asm volatile("cp.reduce.async.bulk.tensor.4d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3, %4, %5}], [%1];"
                 :: "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
                 : "memory");
  else
    asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.tile.bulk_group [%0, {%2, %3, %4, %5}], [%1];"
                 :: "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
                 : "memory");
}

// TMEM accumulator -> bf16 tile image in `out_stage` (BLOCK_N / 64 halves of kOutHalfBytes);
// bias and relu applied on the way.  Called by the four epilogue warps.
// keep = false: this thread's row is not part of the output (halo convolution) and is staged as zeros.
// row_base / half_bytes: the tile image may be taller than 128 rows (two M tiles of one CTA): this
// accumulator's rows start at row_base and the 64-column halves are half_bytes apart.
template <int BLOCK_N>
__device__ __forceinline__ void drain_tile_staged(uint32_t tmem_acc, uint8_t* out_stage, int warp, int lane, int n0,
                                                  const Epilogue& epi, bool keep = true, int row_base = 0,
                                                  int half_bytes = kOutHalfBytes) {
  const int q = warp & 3;
  const int row = row_base + q * 32 + lane;
#pragma unroll 1
  for (int c0 = 0; c0 < BLOCK_N; c0 += 32) {
    uint32_t r[32];
    const uint32_t taddr = tmem_acc + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(c0);
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
    uint8_t* half = out_stage + (c0 >> 6) * half_bytes + row * 128;
    const int j0 = (c0 & 63) >> 3;
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      float v[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) v[k] = keep ? __uint_as_float(r[u * 8 + k]) : 0.f;
      if (epi.bias) {
#pragma unroll
        for (int k = 0; k < 8; ++k) v[k] += epi.bias[n0 + c0 + u * 8 + k];
      }
      if (epi.relu) {
#pragma unroll
        for (int k = 0; k < 8; ++k) v[k] = fmaxf(v[k], 0.f);
      }
      uint4 pk;
      __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
      for (int k = 0; k < 4; ++k) h[k] = __floats2bfloat162_rn(v[2 * k], v[2 * k + 1]);
      *reinterpret_cast<uint4*>(half + (((j0 + u) ^ (row & 7)) << 4)) = pk;
    }
  }
}

// Column sums of the staged bf16 tile images (batch-norm statistics of a convolution output, taken
// while each tile is still in shared memory) by the eight column-sum warps: thread et owns the
// column pair et % (BLOCK_N / 2) and one of 256 / (BLOCK_N / 2) row slices, reads its two columns
// through the swizzle (a warp reads one whole 128-byte row per instruction: conflict-free) and
// keeps two independent chains per column.  The launch makes the grid a multiple of tiles_n, so
// with the n-fastest tile order a CTA keeps its column block: the sums run over ALL its tiles in
// registers and leave as one record per CTA row, stats[(blockIdx.x / tiles_n) * parts + slice][2][N].
template <int BLOCK_N> constexpr int col_stats_parts() { return kStatsThreads / (BLOCK_N / 2); }   // 4 or 8

struct ColStats {
  float s0[2] = {0.f, 0.f}, s1[2] = {0.f, 0.f}, q0[2] = {0.f, 0.f}, q1[2] = {0.f, 0.f};
};

template <int BLOCK_N, int ROWS = kBlockM>
__device__ __forceinline__ void tile_col_stats(const uint8_t* img, int et, int valid, ColStats& a) {
  constexpr int kPairs = BLOCK_N / 2;
  constexpr int kRows = ROWS / col_stats_parts<BLOCK_N>();
  const int c = (et % kPairs) * 2, part = et / kPairs;
  const uint8_t* col = img + (c >> 6) * (ROWS * 128) + (c & 7) * 2;
  const int j = (c & 63) >> 3;
  const int r_end = min(valid, (part + 1) * kRows);
#pragma unroll 1
  for (int r0 = part * kRows; r0 < r_end; r0 += 8) {   // r0 is a multiple of 8: row r0 + i has swizzle phase i
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (r0 + i < r_end) {
        const uint32_t w = *reinterpret_cast<const uint32_t*>(col + (r0 + i) * 128 + ((j ^ i) << 4));
        const float lo = __uint_as_float(w << 16), hi = __uint_as_float(w & 0xffff0000u);
        a.s0[i & 1] += lo; a.q0[i & 1] = fmaf(lo, lo, a.q0[i & 1]);
        a.s1[i & 1] += hi; a.q1[i & 1] = fmaf(hi, hi, a.q1[i & 1]);
      }
    }
  }
}

template <int BLOCK_N>
__device__ __forceinline__ void store_col_stats(const ColStats& a, int et, int n0, int N, int cta_row,
                                                float* __restrict__ stats) {
  constexpr int kPairs = BLOCK_N / 2;
  const int c = (et % kPairs) * 2, part = et / kPairs;
  if (n0 + c < N) {   // N is even
    float* dst = stats + (static_cast<size_t>(cta_row) * col_stats_parts<BLOCK_N>() + part) * 2 * N + n0 + c;
    *reinterpret_cast<float2*>(dst) = make_float2(a.s0[0] + a.s0[1], a.s1[0] + a.s1[1]);
    *reinterpret_cast<float2*>(dst + N) = make_float2(a.q0[0] + a.q0[1], a.q1[0] + a.q1[1]);
  }
}

// grid of a persistent launch: one CTA per SM, rounded down to a multiple of tiles_n when column
// statistics are taken (see above)
inline int persistent_grid(long long total_tiles, int tiles_n, bool col_stats, int sms) {
  long long g = total_tiles < sms ? total_tiles : sms;
  if (col_stats && g >= tiles_n) g -= g % tiles_n;
  return static_cast<int>(g);
}

// ---- persistent variant ----------------------------------------------------------------------------
// One CTA per SM walks the tiles t = blockIdx.x, blockIdx.x + gridDim.x, ...  The TMA ring keeps
// running across tile boundaries and the accumulator is double-buffered in TMEM (2 x BLOCK_N
// columns): while the epilogue warps drain tile i, the MMA warp already accumulates tile i + 1.
// Used when the grid is at least two waves deep (the ResNet-50 layers); tile = (n fastest, m, z), so
// the CTAs that share an A tile run at the same time and the tile comes from DRAM once.
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(smem_u32(bar)) : "memory");
}

// Ring depth of the persistent kernels: whatever fits beside the two output tile images.  The TMA
// round trip is 2 us and more under load, so bytes in flight per SM = stages x stage size is what
// sets the bandwidth of the memory-bound GEMMs (1x1 convolutions): 8 x 24 KB / 5 x 32 KB.
template <int BLOCK_N> constexpr int persistent_stages() { return BLOCK_N == 64 ? 8 : 5; }

// [ring stages][2 output tile images (the fp32 transpose staging of the fallback epilogue aliases
// them)][barriers][TMEM slot]
template <int BLOCK_N>
constexpr size_t smem_bytes_persistent() {
  return 1024 + static_cast<size_t>(persistent_stages<BLOCK_N>()) * (kBlockM * kBlockKBytes + BLOCK_N * kBlockKBytes) +
         2 * (BLOCK_N / 64) * kOutHalfBytes + (2 * persistent_stages<BLOCK_N>() + 8) * sizeof(uint64_t) + 16;
}
static_assert(2 * kOutHalfBytes >= 4 * 32 * 33 * sizeof(float), "the fp32 staging must fit inside the output images");
static_assert(smem_bytes_persistent<128>() <= 232448 && smem_bytes_persistent<64>() <= 232448, "shared memory budget");

template <int BLOCK_N, int KIND, bool A_MN, bool B_MN>
__global__ void __launch_bounds__(kThreadsPersistent, 1)
gemm_tc_persistent_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                          const __grid_constant__ CUtensorMap map_c, float* __restrict__ C, int M, int N, int ldc, int k_blocks, int k_blocks_per_split,
                          int tiles_m, int tiles_n, int splits, Epilogue epi) {
  constexpr int kStages = persistent_stages<BLOCK_N>();
  constexpr int kElemBytes = KIND == 0 ? 2 : 4;
  constexpr int kBlockK = kBlockKBytes / kElemBytes;
  constexpr int kABytes = kBlockM * kBlockKBytes, kBBytes = BLOCK_N * kBlockKBytes;
  constexpr int kMnElems = 128 / kElemBytes;
  constexpr int kChunkBytes = kBlockK * 128;
  constexpr int kUmmaK = 32 / kElemBytes;
  constexpr uint32_t kStepA = A_MN ? (kUmmaK * 128) >> 4 : 32 >> 4;
  constexpr uint32_t kStepB = B_MN ? (kUmmaK * 128) >> 4 : 32 >> 4;
  static_assert(BLOCK_N == 64 || BLOCK_N == 128, "two accumulators of BLOCK_N columns must fit a power-of-two allocation");

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + kStages * kABytes;
  constexpr int kOutTileBytes = (BLOCK_N / 64) * kOutHalfBytes;
  uint8_t* out_stage = smem + kStages * (kABytes + kBBytes);           // 1024-aligned: every stage is
  float* stage_base = reinterpret_cast<float*>(out_stage);             // fallback epilogue only
  uint64_t* full = reinterpret_cast<uint64_t*>(out_stage + 2 * kOutTileBytes);
  uint64_t* empty = full + kStages;
  uint64_t* tmem_full = empty + kStages;     // [2]
  uint64_t* tmem_empty = tmem_full + 2;      // [2]
  uint64_t* img_full = tmem_empty + 2;       // [2] output tile image staged (column-sum warps may read it)
  uint64_t* img_free = img_full + 2;         // [2] column sums of the image taken
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(img_free + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int total_tiles = tiles_m * tiles_n * splits;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" :: "l"(reinterpret_cast<uint64_t>(&map_a)) : "memory");
    asm volatile("prefetch.tensormap [%0];" :: "l"(reinterpret_cast<uint64_t>(&map_b)) : "memory");
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int s = 0; s < kStages; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
      for (int a = 0; a < 2; ++a) {
        mbar_init(&tmem_full[a], 1); mbar_init(&tmem_empty[a], 4);
        mbar_init(&img_full[a], 1); mbar_init(&img_free[a], kStatsThreads / 32);
      }
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                 :: "r"(smem_u32(tmem_slot)), "n"(2 * BLOCK_N) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===== TMA producer: one continuous stream of K blocks over all of this CTA's tiles =====
    if (lane == 0) {
      uint32_t it = 0;
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
        const int n0 = (t % tiles_n) * BLOCK_N, m0 = ((t / tiles_n) % tiles_m) * kBlockM, z = t / (tiles_m * tiles_n);
        const int kb_begin = z * k_blocks_per_split, kb_end = min(k_blocks, kb_begin + k_blocks_per_split);
        for (int kb = kb_begin; kb < kb_end; ++kb, ++it) {
          const int s = it % kStages;
          const uint32_t phase = (it / kStages) & 1;
          mbar_wait(&empty[s], phase ^ 1);
          mbar_expect_tx(&full[s], kABytes + kBBytes);
          const int k_elem = kb * kBlockK;
          if (!A_MN) {
            tma_load_2d(smem_a + s * kABytes, &map_a, &full[s], k_elem, m0);
          } else {
#pragma unroll
            for (int c = 0; c < kBlockM / kMnElems; ++c)
              tma_load_2d(smem_a + s * kABytes + c * kChunkBytes, &map_a, &full[s], m0 + c * kMnElems, k_elem);
          }
          if (!B_MN) {
            tma_load_2d(smem_b + s * kBBytes, &map_b, &full[s], k_elem, n0);
          } else {
#pragma unroll
            for (int c = 0; c < BLOCK_N / kMnElems; ++c)
              tma_load_2d(smem_b + s * kBBytes + c * kChunkBytes, &map_b, &full[s], n0 + c * kMnElems, k_elem);
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: accumulator `acc` alternates between the two TMEM halves; converged warp,
    // one elected lane issues ("lean issue path") =====
    constexpr uint32_t idesc = make_idesc(KIND == 0 ? 1 : 2, BLOCK_N, A_MN, B_MN);
    constexpr bool kBase32 = KIND == 1;
    constexpr uint32_t hi_a = desc_hi((A_MN && kBase32) ? 512 : 1024, (A_MN && kBase32) ? 1 : 2);
    constexpr uint32_t hi_b = desc_hi((B_MN && kBase32) ? 512 : 1024, (B_MN && kBase32) ? 1 : 2);
    const uint32_t a0 = smem_u32(smem_a), b0 = smem_u32(smem_b);
    const uint32_t full0 = smem_u32(full), empty0 = smem_u32(empty);
    const uint32_t tmem_full0 = smem_u32(tmem_full), tmem_empty0 = smem_u32(tmem_empty);
    uint32_t st = 0, ph = 0, acc = 0, acc_ph = 0;
    for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
      const int z = t / (tiles_m * tiles_n);
      const int kb_begin = z * k_blocks_per_split, kb_end = min(k_blocks, kb_begin + k_blocks_per_split);
      mbar_wait_addr(tmem_empty0 + acc * 8, acc_ph ^ 1);      // the epilogue has drained this half
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const uint32_t d_tmem = tmem_base + acc * BLOCK_N;
      for (int kb = kb_begin; kb < kb_end; ++kb) {
        mbar_wait_addr(full0 + st * 8, ph);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint32_t a_lo = desc_lo(a0 + st * kABytes, A_MN ? kChunkBytes : 16);
        const uint32_t b_lo = desc_lo(b0 + st * kBBytes, B_MN ? kChunkBytes : 16);
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < kBlockK / kUmmaK; ++k)
            umma2<KIND>(d_tmem, a_lo + kStepA * k, hi_a, b_lo + kStepB * k, hi_b, idesc, kb > kb_begin || k > 0);
          umma_commit_addr(empty0 + st * 8);
        }
        __syncwarp();
        if (++st == kStages) { st = 0; ph ^= 1; }
      }
      if (elect_one()) umma_commit_addr(tmem_full0 + acc * 8);
      __syncwarp();
      acc ^= 1;
      if (acc == 0) acc_ph ^= 1;
    }
  } else if (warp < 6) {
    // ===== epilogue warps =====
    uint32_t tc = 0;
    for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, ++tc) {
      const int n0 = (t % tiles_n) * BLOCK_N, m0 = ((t / tiles_n) % tiles_m) * kBlockM, z = t / (tiles_m * tiles_n);
      const uint32_t acc = tc & 1, acc_phase = (tc >> 1) & 1;
      if (epi.tma_out) {
        uint8_t* img = out_stage + acc * kOutTileBytes;
        if (warp == 2 && lane == 0) bulk_wait_read<1>();    // the store of two tiles ago has read this image
        if (epi.col_stats && tc >= 2) mbar_wait(&img_free[acc], ((tc >> 1) - 1) & 1);   // ... and so have the column sums
        epi_barrier();
        mbar_wait(&tmem_full[acc], acc_phase);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        drain_tile_staged<BLOCK_N>(tmem_base + acc * BLOCK_N, img, warp, lane, n0, epi);
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncwarp();
        if (lane == 0) mbar_arrive(&tmem_empty[acc]);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        epi_barrier();
        if (warp == 2 && lane == 0) {
#pragma unroll
          for (int h = 0; h < BLOCK_N / 64; ++h)
            if (n0 + h * 64 < N) tma_store_2d(&map_c, img + h * kOutHalfBytes, n0 + h * 64, m0, epi.tma_out == 2);
          bulk_commit();
          if (epi.col_stats) mbar_arrive(&img_full[acc]);
        }
        continue;
      }
      mbar_wait(&tmem_full[acc], acc_phase);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      drain_tile<BLOCK_N>(tmem_base + acc * BLOCK_N, true, stage_base, warp, lane, C, M, N, ldc, m0, n0, z, epi);
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty[acc]);       // 4 arrivals (one per epilogue warp) free the half
    }
    if (epi.tma_out && warp == 2 && lane == 0) bulk_wait_all();   // shared memory must outlive the last store
  } else if (epi.tma_out && epi.col_stats) {
    // ===== column-sum warps: batch-norm statistics of the staged tiles, off the epilogue's critical path =====
    uint32_t tc = 0;
    ColStats sums;
    const int et = threadIdx.x - kThreads;
    for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, ++tc) {
      const int m0 = ((t / tiles_n) % tiles_m) * kBlockM;
      const uint32_t acc = tc & 1;
      mbar_wait(&img_full[acc], (tc >> 1) & 1);
      tile_col_stats<BLOCK_N>(out_stage + acc * kOutTileBytes, et, min(kBlockM, M - m0), sums);
      __syncwarp();
      if (lane == 0) mbar_arrive(&img_free[acc]);
    }
    store_col_stats<BLOCK_N>(sums, et, (blockIdx.x % tiles_n) * BLOCK_N, N, blockIdx.x / tiles_n, epi.col_stats);
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem_base), "n"(2 * BLOCK_N) : "memory");
  }
}

}  // namespace tc
}  // namespace tfg
