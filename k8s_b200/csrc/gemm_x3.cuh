// fp32-equivalent GEMM on tcgen05 ("3xTF32"):  C[M,N] (fp32) = A(M x K) * B(N x K)^T with fp32 operands.
//
// SURVEY.md par. 8d fixes the sample model of BASELINE configs 1-3 at fp32 and allows the tensor
// pipe only "if the tolerance is met".  One kind::tf32 MMA reads 10 mantissa bits per operand
// (1e-3 relative); this kernel recovers fp32 accuracy from three of them per K step:
//
//   * split.  Each operand X comes as two fp32 tensors of the same shape: X itself -- kind::tf32
//     reads the top 19 bits of an fp32 word, i.e. it sees hi = trunc_tf32(X) -- and a low-order
//     plane lo = rn_tf32(X - hi) written by whichever kernel produced X (tf32_lo(), gemm_tc.cuh).
//     The accumulator receives A_lo*B + A*B_lo + A*B; what is dropped is lo*lo (< 2^-20 relative)
//     and the rounding of lo (2^-22).
//   * chunked accumulation.  The tensor core adds into its fp32 accumulator with TRUNCATION: every
//     MMA loses about one ulp of the accumulator's current magnitude, always in the same direction
//     (measured here, K = 3136, same-sign operands, all 1176 MMAs into one TMEM accumulator:
//     6.2e-5 relative -- sixty times the operand error; chunks of 24 MMAs: 1.2e-6).  Two measures:
//       - the two small products (A_lo*B, A*B_lo) go to their own accumulator S, whose magnitude --
//         and therefore whose ulp -- is 2^-11 of the main sum's: their truncation vanishes;
//       - the main product A*B accumulates in chunks of `chunk_kb` K blocks (32 fp32 of K each) that
//         start from zero in one of two TMEM accumulators M0 / M1; the four epilogue warps read a
//         finished chunk back (tcgen05.ld) and add it to the running sum in REGISTERS with
//         round-to-nearest while the MMA warp fills the other one.  A chunk of 2 K blocks holds 8
//         truncating adds at the magnitude of a 64-term partial sum: ~4e-7 worst case (same-sign
//         data), the level of a sequential fp32 sum.
//     S is read back every kSPeriodKb K blocks and at the end.  Drain cost: 128 x BLOCK_N x 4 B per
//     chunk at 64 B/clk of TMEM read = 512 clk (BLOCK_N = 64) against 768 clk of MMAs (24
//     instructions of 32 clk), so the tensor pipe stays the limiter.
//
// Layout and roles as gemm_tc_kernel: warp 0 = TMA producer (a stage holds [A | A_lo] and
// [B | B_lo], 128-byte swizzle, K-major or MN-major operands), warp 1 = TMEM allocator + MMA
// issuer (converged warp, one elected lane), warps 2..5 = accumulate + epilogue (warp w owns TMEM
// lanes 32*(w%4)..+31; thread = output row).  gridDim.z > 1 = split-K into C + z*M*ldc.
// TMEM columns: [S | M0 | M1] = 3 x BLOCK_N, allocated as 4 x BLOCK_N.
#pragma once
#include "gemm_tc.cuh"

namespace tfg {
namespace tc {

constexpr int kSPeriodKb = 128;   // K blocks (K = 4096) between two read-backs of the small-term accumulator

template <int BLOCK_N, int STAGES>
constexpr size_t smem_bytes_x3() {
  return 1024 + static_cast<size_t>(STAGES) * 2 * (kBlockM * kBlockKBytes + BLOCK_N * kBlockKBytes) +
         (2 * STAGES + 6) * sizeof(uint64_t) + 16;
}

// 32 consecutive accumulator columns of this warp's lane quarter -> r[0..31] (thread = row)
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
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

// One output row held in registers -> C (fp32), with bias / relu / mask.  16-byte stores where the
// row piece is aligned and inside N, scalar stores for ragged tails.
template <int BLOCK_N>
__device__ __forceinline__ void store_row_regs(const float (&v)[BLOCK_N], float* __restrict__ crow,
                                               const float* __restrict__ mrow, int n0, int N, bool vec_ok,
                                               const Epilogue& epi) {
#pragma unroll
  for (int u = 0; u < BLOCK_N; u += 4) {
    const int col = n0 + u;
    if (col >= N) break;
    float x[4] = {v[u], v[u + 1], v[u + 2], v[u + 3]};
    if (vec_ok && col + 3 < N) {
      if (epi.bias) {
        const float4 b = *reinterpret_cast<const float4*>(epi.bias + col);
        x[0] += b.x; x[1] += b.y; x[2] += b.z; x[3] += b.w;
      }
      if (epi.relu) {
#pragma unroll
        for (int k = 0; k < 4; ++k) x[k] = fmaxf(x[k], 0.f);
      }
      if (mrow) {
        const float4 k4 = *reinterpret_cast<const float4*>(mrow + col);
        x[0] = k4.x > 0.f ? x[0] : 0.f; x[1] = k4.y > 0.f ? x[1] : 0.f;
        x[2] = k4.z > 0.f ? x[2] : 0.f; x[3] = k4.w > 0.f ? x[3] : 0.f;
      }
      *reinterpret_cast<float4*>(crow + col) = make_float4(x[0], x[1], x[2], x[3]);
    } else {
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        if (col + k < N) {
          float y = x[k];
          if (epi.bias) y += epi.bias[col + k];
          if (epi.relu) y = fmaxf(y, 0.f);
          if (mrow) y = mrow[col + k] > 0.f ? y : 0.f;
          crow[col + k] = y;
        }
      }
    }
  }
}

template <int BLOCK_N, bool A_MN, bool B_MN, int STAGES>
__global__ void __launch_bounds__(kThreads)
gemm_x3_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
               const __grid_constant__ CUtensorMap map_a_lo, const __grid_constant__ CUtensorMap map_b_lo,
               float* __restrict__ C, int M, int N, int ldc, int k_blocks, int k_blocks_per_split, int chunk_kb,
               Epilogue epi) {
  constexpr int kBlockK = kBlockKBytes / 4;                    // 32 fp32 of K per stage
  constexpr int kAPlane = kBlockM * kBlockKBytes, kBPlane = BLOCK_N * kBlockKBytes;
  constexpr int kABytes = 2 * kAPlane, kBBytes = 2 * kBPlane;  // [X | X_lo] per stage
  constexpr int kMnElems = 32;                                 // rows per 128-B line of an MN-major box
  constexpr int kChunkBytes = kBlockK * 128;                   // one MN-major box
  constexpr int kUmmaK = 8;
  constexpr uint32_t kStepA = A_MN ? (kUmmaK * 128) >> 4 : 32 >> 4;
  constexpr uint32_t kStepB = B_MN ? (kUmmaK * 128) >> 4 : 32 >> 4;
  static_assert(BLOCK_N == 64 || BLOCK_N == 128, "[S | M0 | M1] of BLOCK_N columns each, allocated as 4 x BLOCK_N <= 512");

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + STAGES * kABytes;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + STAGES * (kABytes + kBBytes));
  uint64_t* empty = full + STAGES;
  uint64_t* tmem_full = empty + STAGES;      // [2]
  uint64_t* tmem_empty = tmem_full + 2;      // [2]
  uint64_t* s_full = tmem_empty + 2;         // small-term accumulator ready to be read back
  uint64_t* s_empty = s_full + 1;            // ... and read
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(s_empty + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m0 = blockIdx.y * kBlockM, n0 = blockIdx.x * BLOCK_N;
  const int kb_begin = blockIdx.z * k_blocks_per_split;
  const int kb_end = min(k_blocks, kb_begin + k_blocks_per_split);
  const int n_kb = max(kb_end - kb_begin, 0);
  const int n_chunks = (n_kb + chunk_kb - 1) / chunk_kb;
  const int s_period = max(1, kSPeriodKb / chunk_kb);          // chunks per epoch of S

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" :: "l"(reinterpret_cast<uint64_t>(&map_a)) : "memory");
    asm volatile("prefetch.tensormap [%0];" :: "l"(reinterpret_cast<uint64_t>(&map_b)) : "memory");
    asm volatile("prefetch.tensormap [%0];" :: "l"(reinterpret_cast<uint64_t>(&map_a_lo)) : "memory");
    asm volatile("prefetch.tensormap [%0];" :: "l"(reinterpret_cast<uint64_t>(&map_b_lo)) : "memory");
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int s = 0; s < STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
      for (int a = 0; a < 2; ++a) { mbar_init(&tmem_full[a], 1); mbar_init(&tmem_empty[a], 4); }
      mbar_init(s_full, 1); mbar_init(s_empty, 4);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                 :: "r"(smem_u32(tmem_slot)), "n"(4 * BLOCK_N) : "memory");
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
        const int s = i % STAGES;
        const uint32_t phase = (i / STAGES) & 1;
        mbar_wait(&empty[s], phase ^ 1);
        mbar_expect_tx(&full[s], kABytes + kBBytes);
        const int k_elem = (kb_begin + i) * kBlockK;
#pragma unroll
        for (int pl = 0; pl < 2; ++pl) {                         // plane 1 = the low-order parts
          const CUtensorMap* ma = pl ? &map_a_lo : &map_a;
          const CUtensorMap* mb = pl ? &map_b_lo : &map_b;
          uint8_t* sa = smem_a + s * kABytes + pl * kAPlane;
          uint8_t* sb = smem_b + s * kBBytes + pl * kBPlane;
          if (!A_MN) {
            tma_load_2d(sa, ma, &full[s], k_elem, m0);
          } else {
#pragma unroll
            for (int c = 0; c < kBlockM / kMnElems; ++c)
              tma_load_2d(sa + c * kChunkBytes, ma, &full[s], m0 + c * kMnElems, k_elem);
          }
          if (!B_MN) {
            tma_load_2d(sb, mb, &full[s], k_elem, n0);
          } else {
#pragma unroll
            for (int c = 0; c < BLOCK_N / kMnElems; ++c)
              tma_load_2d(sb + c * kChunkBytes, mb, &full[s], n0 + c * kMnElems, k_elem);
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: chunk c accumulates the main product from zero in TMEM accumulator M[c & 1],
    // the small products run on in S =====
    constexpr uint32_t idesc = make_idesc(2, BLOCK_N, A_MN, B_MN);
    constexpr uint32_t hi_a = desc_hi(A_MN ? 512 : 1024, A_MN ? 1 : 2);   // MN-major tf32: 32-byte-base swizzle
    constexpr uint32_t hi_b = desc_hi(B_MN ? 512 : 1024, B_MN ? 1 : 2);
    const uint32_t a0 = smem_u32(smem_a), b0 = smem_u32(smem_b);
    const uint32_t full0 = smem_u32(full), empty0 = smem_u32(empty);
    const uint32_t tmem_full0 = smem_u32(tmem_full), tmem_empty0 = smem_u32(tmem_empty);
    const uint32_t s_full0 = smem_u32(s_full), s_empty0 = smem_u32(s_empty);
    uint32_t st = 0, ph = 0;
    int kb = 0, epoch = 0;
    bool s_fresh = true;                                         // the next small product overwrites S
    for (int c = 0; c < n_chunks; ++c) {
      if (c > 0 && c % s_period == 0) {                          // hand S over, go on when it has been read
        if (elect_one()) umma_commit_addr(s_full0);
        __syncwarp();
        mbar_wait_addr(s_empty0, epoch & 1);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        ++epoch;
        s_fresh = true;
      }
      const uint32_t acc = c & 1;
      mbar_wait_addr(tmem_empty0 + acc * 8, ((c >> 1) & 1) ^ 1);      // the adders have read this accumulator's last chunk
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const uint32_t d_main = tmem_base + (1 + acc) * BLOCK_N, d_small = tmem_base;
      const int kb_stop = min(n_kb, kb + chunk_kb);
      for (bool first = true; kb < kb_stop; ++kb, first = false) {
        mbar_wait_addr(full0 + st * 8, ph);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint32_t a_lo = desc_lo(a0 + st * kABytes, A_MN ? kChunkBytes : 16);
        const uint32_t b_lo = desc_lo(b0 + st * kBBytes, B_MN ? kChunkBytes : 16);
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < kBlockK / kUmmaK; ++k) {
            umma2<1>(d_small, a_lo + (kAPlane >> 4) + kStepA * k, hi_a, b_lo + kStepB * k, hi_b, idesc, !(s_fresh && k == 0));
            umma2<1>(d_small, a_lo + kStepA * k, hi_a, b_lo + (kBPlane >> 4) + kStepB * k, hi_b, idesc, true);
            umma2<1>(d_main, a_lo + kStepA * k, hi_a, b_lo + kStepB * k, hi_b, idesc, !(first && k == 0));
          }
          umma_commit_addr(empty0 + st * 8);             // frees the stage when these MMAs retire
        }
        __syncwarp();
        s_fresh = false;
        if (++st == STAGES) { st = 0; ph ^= 1; }
      }
      if (elect_one()) {
        umma_commit_addr(tmem_full0 + acc * 8);          // chunk complete
        if (c == n_chunks - 1) umma_commit_addr(s_full0);
      }
      __syncwarp();
    }
  } else {
    // ===== accumulate (fp32, round-to-nearest, in registers) + epilogue =====
    const int q = warp & 3;
    const uint32_t lane_base = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
    float sum[BLOCK_N];
#pragma unroll
    for (int u = 0; u < BLOCK_N; ++u) sum[u] = 0.f;
    int epoch = 0;
    for (int c = 0; c < n_chunks; ++c) {
      const uint32_t acc = c & 1;
      mbar_wait(&tmem_full[acc], (c >> 1) & 1);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
      for (int c0 = 0; c0 < BLOCK_N; c0 += 32) {
        uint32_t r[32];
        tmem_ld32(lane_base + (1 + acc) * BLOCK_N + c0, r);
#pragma unroll
        for (int u = 0; u < 32; ++u) sum[c0 + u] = __fadd_rn(sum[c0 + u], __uint_as_float(r[u]));
      }
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty[acc]);      // 4 arrivals (one per warp) hand the accumulator back
      if (c == n_chunks - 1 || (c + 1) % s_period == 0) {   // the small products of this epoch
        mbar_wait(s_full, epoch & 1);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
        for (int c0 = 0; c0 < BLOCK_N; c0 += 32) {
          uint32_t r[32];
          tmem_ld32(lane_base + c0, r);
#pragma unroll
          for (int u = 0; u < 32; ++u) sum[c0 + u] = __fadd_rn(sum[c0 + u], __uint_as_float(r[u]));
        }
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncwarp();
        if (lane == 0) mbar_arrive(s_empty);
        ++epoch;
      }
    }
    const int row = m0 + q * 32 + lane;
    if (row < M) {
      const bool vec_ok = (ldc & 3) == 0 && (reinterpret_cast<uintptr_t>(C) & 15) == 0 &&
                          (!epi.bias || (reinterpret_cast<uintptr_t>(epi.bias) & 15) == 0) &&
                          (!epi.mask || (reinterpret_cast<uintptr_t>(epi.mask) & 15) == 0);
      float* crow = C + (static_cast<size_t>(blockIdx.z) * M + row) * static_cast<size_t>(ldc);
      const float* mrow = epi.mask ? epi.mask + static_cast<size_t>(row) * ldc : nullptr;
      store_row_regs<BLOCK_N>(sum, crow, mrow, n0, N, vec_ok, epi);
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem_base), "n"(4 * BLOCK_N) : "memory");
  }
}

}  // namespace tc
}  // namespace tfg
