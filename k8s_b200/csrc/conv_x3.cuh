// conv2 of the sample model (5x5, stride 1, SAME, 14x14 images, 32 -> 64 channels) as implicit GEMMs
// in fp32-equivalent 3xTF32 arithmetic (gemm_x3.cuh): forward, data gradient and weight gradient with
// NO im2col matrix, no dcol and no col2im.  fp32 NHWC activations; a pixel's 32 channels are one
// 128-byte line, so the layouts are those of a 64-channel bf16 layer (conv_tc.cuh).
//
// The image lives in shared memory once, as a zero-padded slab: position q = hh * 18 + ww of an
// 18-column grid holds input pixel (hh - 2, ww - 2); the TMA unit zero-fills the border.  An MMA row
// m stands for output position (m / 18, m % 18); its input under filter tap (r, s) is slab position
// m + r * 18 + s, so the A operand of EVERY tap is the same slab read from a start address
// (r * 18 + s) * 128 B further on (the MMA unit swizzles on absolute shared-memory address bits, so
// the shifted start is consistent with what the TMA unit wrote).  Rows with m % 18 >= 14 are not
// outputs and are never stored.
//
//   forward  y[pix, n]   = sum_{r,s,c} x[pix + (r,s) - 2, c] W[r,s,c,n] + b[n]     A = x slab (K-major), B = W (MN-major)
//   dgrad    dx[pix, c]  = sum_{r,s,n} dy[pix - (r,s) + 2, n] W[r,s,c,n], masked    A = dy slab (K-major, 2 channel blocks), B = W (K-major)
//   wgrad    dW[(r,s,c), n] = sum_pix x[pix + (r,s) - 2, c] dy[pix, n]              A = x slab (MN-major views), B = dy (MN-major)
//
// Every operand comes with its low-order plane.  These MMAs are bound by shared-memory reads of the A
// operand (N = 64 / 32 output channels: 4 KB of A per instruction against 2 / 1 KB of B), so forward and
// data gradient put W and W_lo side by side along N: ONE instruction A x [W | W_lo] yields the main
// product (columns [0, N)) and the small product A*W_lo (columns [N, 2N)) for one read of A, and a second
// one adds A_lo x W into the small columns -- two reads of A-sized operands per K step instead of three.
// The accumulator [main | small] alternates between two TMEM buffers per chunk of taps and is summed in
// registers by the epilogue warps (gemm_x3.cuh explains why); the small columns accumulate at their own
// magnitude, so the tensor core's truncation costs them nothing.
#pragma once
#include "gemm_x3.cuh"
#include "conv_tc.cuh"

namespace tfg {
namespace tc {

constexpr int kC2H = 14, kC2W = 14, kC2PW = 18;            // image and padded-grid width
constexpr int kC2Pos = 256;                                // MMA rows per image: two M tiles cover the 252 grid positions of 14 rows
constexpr int kC2SlabLines = 336;                          // >= 255 + 4 * 18 + 4 + 1 lines of 128 B; the TMA box writes 18 * 18 = 324
constexpr int kC2SlabBytes = kC2SlabLines * 128;           // 43008 = 42 * 1024: every slab stays 1024-B aligned
constexpr int kC2Threads = 64 + 256;                       // TMA warp, MMA warp, eight accumulate/epilogue warps

// Development aid (make EXTRA=-DTFG_TRACE, scripts/conv_trace.py): cycles each role of a CTA spends waiting
// on each barrier, read back with tfg_debug_c2_trace.  Not part of the shipped build.
#ifdef TFG_TRACE
__device__ long long g_c2_trace[2][160][16];
#define C2T_DECL(...) long long __VA_ARGS__
#define C2T_NOW() clock64()
#define C2T_SET(v) v = clock64()
#define C2T_ADD(acc, t0) acc += clock64() - (t0)
#define C2T_PUT(slot, v) do { if (lane == 0) g_c2_trace[DGRAD ? 1 : 0][blockIdx.x][slot] = (v); } while (0)
#else
#define C2T_DECL(...)
#define C2T_NOW() 0
#define C2T_SET(v)
#define C2T_ADD(acc, t0)
#define C2T_PUT(slot, v)
#endif

// ---- forward (DGRAD = false) and data gradient (DGRAD = true) -----------------------------------------
//   N_OUT   = output channels of the GEMM: 64 (forward) / 32 (data gradient)
//   CBLOCKS = 32-channel blocks of the gathered tensor: 1 (x has 32 channels) / 2 (dy has 64)
// map_x / map_x_lo: 4-D maps over the gathered tensor [Nimg, 14, 14, 32 * CBLOCKS], box {32, 18, 18, 1}.
// map_w / map_w_lo: W2 viewed [800, 64]; forward reads boxes {32 n, 32 k-lines} (MN-major), the data
//                   gradient boxes {32 k, 32 rows = input channels of the tap} (K-major).
// Output out[pixels, N_OUT] fp32 (+ bias[N_OUT]) (kept where mask[pixels, N_OUT] > 0).
// One image per tile, persistent over images; B_STAGES weight stages of one tap each:
//   forward  [W n 0..31 | W n 32..63 | W_lo n 0..31 | W_lo n 32..63], MN-major boxes of 4 KB (group stride 4 KB)
//   dgrad    per K block (32 of the 64 dy channels): [W rows c 0..31 | W_lo rows c 0..31], K-major, 8 KB
template <bool DGRAD> constexpr int c2_b_stage_bytes() { return 2 /*planes*/ * 64 * 32 * 4; }   // 16 KB: [W | W_lo] of one tap
template <bool DGRAD> constexpr int c2_b_stages() { return DGRAD ? 3 : 4; }
// forward only: the image's 14 x 14 x 64 outputs staged for the 2x2 max-pool of the epilogue, one position per
// line of 64 + 4 floats (the pad spreads the row-per-thread stores over the banks)
constexpr int kC2PoolPitch = 68;
template <bool DGRAD> constexpr int c2_pool_bytes() { return DGRAD ? 0 : kC2H * kC2W * kC2PoolPitch * 4; }
template <bool DGRAD> constexpr int c2_slabs() { return DGRAD ? 4 : 2; }                        // planes x channel blocks
template <bool DGRAD>
constexpr size_t smem_bytes_c2() {
  return 1024 + static_cast<size_t>(c2_slabs<DGRAD>()) * kC2SlabBytes + c2_b_stages<DGRAD>() * c2_b_stage_bytes<DGRAD>() +
         c2_pool_bytes<DGRAD>() + (2 * c2_b_stages<DGRAD>() + 8) * sizeof(uint64_t) + 16;
}
static_assert(smem_bytes_c2<false>() <= 232448 && smem_bytes_c2<true>() <= 232448, "shared memory budget");

template <bool DGRAD>
__global__ void __launch_bounds__(kC2Threads, 1)
conv2_x3_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_x_lo,
                const __grid_constant__ CUtensorMap map_w, const __grid_constant__ CUtensorMap map_w_lo,
                float* __restrict__ out, const float* __restrict__ bias, const float* __restrict__ mask, int Nimg,
                int chunk_taps, float* __restrict__ pool_a, float* __restrict__ pool_lo, uint8_t* __restrict__ pool_idx) {
  constexpr int N_OUT = DGRAD ? 32 : 64;
  constexpr int CBLOCKS = DGRAD ? 2 : 1;
  constexpr int kBStages = c2_b_stages<DGRAD>();
  constexpr int kBStage = c2_b_stage_bytes<DGRAD>();
  constexpr int kAccCols = 4 * N_OUT;                        // one chunk accumulator: [main | small] x two M tiles
  constexpr int kTmemCols = 2 * kAccCols;                    // two of them: 512 (forward) / 256 (data gradient)
  constexpr uint32_t idesc_wide = make_idesc(2, 2 * N_OUT, false, !DGRAD);   // A x [W | W_lo]
  constexpr uint32_t idesc_narrow = make_idesc(2, N_OUT, false, !DGRAD);     // A_lo x W

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* slab = smem;                                      // [plane][cblock] slabs
  uint8_t* smem_b = slab + c2_slabs<DGRAD>() * kC2SlabBytes;
  float* pool_stage = reinterpret_cast<float*>(smem_b + kBStages * kBStage);          // forward only (c2_pool_bytes)
  uint64_t* slab_full = reinterpret_cast<uint64_t*>(smem_b + kBStages * kBStage + c2_pool_bytes<DGRAD>());
  uint64_t* slab_empty = slab_full + 1;
  uint64_t* b_full = slab_empty + 1;
  uint64_t* b_empty = b_full + kBStages;
  uint64_t* tmem_full = b_empty + kBStages;   // [2]
  uint64_t* tmem_empty = tmem_full + 2;       // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int chunks_per_tile = (25 + chunk_taps - 1) / chunk_taps;
  C2T_DECL(t_entry = C2T_NOW(), tw0 = 0, tw1 = 0, tw2 = 0, tq = 0);

  // lines 324..335 of every slab are read by the last taps of rows that are not outputs; the TMA
  // box never writes them, so they are zeroed once (any finite value would do)
  for (int i = threadIdx.x; i < c2_slabs<DGRAD>() * (kC2SlabLines - 324) * 32; i += kC2Threads) {
    const int sl = i / ((kC2SlabLines - 324) * 32), w = i % ((kC2SlabLines - 324) * 32);
    reinterpret_cast<float*>(slab + sl * kC2SlabBytes + 324 * 128)[w] = 0.f;
  }
  // loads of one weight stage (tap) into ring slot s
  auto load_tap = [&](int tap, int s) {
    mbar_expect_tx(&b_full[s], kBStage);
#pragma unroll
    for (int pl = 0; pl < 2; ++pl) {
      const CUtensorMap* mw = pl ? &map_w_lo : &map_w;
      uint8_t* dst = smem_b + s * kBStage;
      if (!DGRAD) {            // rows = K (the tap's 32 input channels), columns = n: two MN-major boxes of 32 n per plane
        tma_load_2d(dst + pl * 8192, mw, &b_full[s], 0, tap * 32);
        tma_load_2d(dst + pl * 8192 + 4096, mw, &b_full[s], 32, tap * 32);
      } else {                 // rows = the tap's 32 input channels (N of this GEMM), columns = K: per K block [W | W_lo]
        tma_load_2d(dst + pl * 4096, mw, &b_full[s], 0, tap * 32);
        tma_load_2d(dst + 8192 + pl * 4096, mw, &b_full[s], 32, tap * 32);
      }
    }
  };
  auto load_slabs = [&](int img) {
    mbar_expect_tx(slab_full, c2_slabs<DGRAD>() * 324 * 128);
#pragma unroll
    for (int pl = 0; pl < 2; ++pl)
#pragma unroll
      for (int cb = 0; cb < CBLOCKS; ++cb)
        tma_load_4d(slab + (pl * CBLOCKS + cb) * kC2SlabBytes, pl ? &map_x_lo : &map_x, slab_full, cb * 32, -2, -2, img);
  };
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // the zeroed lines, before any MMA reads them
  if (warp == 0) {
    // The producer warp sets up the barriers and only ARRIVES at the block's start barrier: its loads are
    // under way while the TMEM allocation and everybody else's wait are still going on.
    if (lane == 0) {
      mbar_init(slab_full, 1); mbar_init(slab_empty, 1);
      for (int s = 0; s < kBStages; ++s) { mbar_init(&b_full[s], 1); mbar_init(&b_empty[s], 1); }
      for (int a = 0; a < 2; ++a) { mbar_init(&tmem_full[a], 1); mbar_init(&tmem_empty[a], 8); }
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    asm volatile("bar.arrive 1, %0;" :: "n"(kC2Threads) : "memory");
  } else {
    if (warp == 1) {
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                   :: "r"(smem_u32(tmem_slot)), "n"(kTmemCols) : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    asm volatile("bar.sync 1, %0;" :: "n"(kC2Threads) : "memory");
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  }
  const uint32_t tmem_base = *tmem_slot;
  // TMEM columns: chunk accumulator a, M tile h at a * kAccCols + h * 2 * N_OUT: [main N_OUT | small N_OUT]

  if (warp == 0) {
    // ===== TMA producer: one slab set per image, then the 25 taps' weight tiles =====
    if (lane == 0) {
      uint32_t tile = 0, bit = 0;
      for (int img = blockIdx.x; img < Nimg; img += gridDim.x, ++tile) {
        C2T_SET(tq);
        mbar_wait(slab_empty, (tile & 1) ^ 1);
        C2T_ADD(tw0, tq);
        load_slabs(img);
        for (int tap = 0; tap < 25; ++tap, ++bit) {
          const int s = bit % kBStages;
          C2T_SET(tq);
          mbar_wait(&b_empty[s], ((bit / kBStages) & 1) ^ 1);
          C2T_ADD(tw1, tq);
          load_tap(tap, s);
        }
      }
      C2T_PUT(0, tw0); C2T_PUT(1, tw1); C2T_PUT(2, C2T_NOW() - t_entry);
    }
  } else if (warp == 1) {
    // ===== MMA issuer (converged warp, one elected lane) =====
    constexpr uint32_t hi_a = desc_hi();                                         // K-major, 128-byte swizzle
    constexpr uint32_t hi_b = DGRAD ? desc_hi() : desc_hi(512, 1);               // forward: MN-major tf32, 32-byte-base swizzle
    constexpr uint32_t kStepB = DGRAD ? (32 >> 4) : ((8 * 128) >> 4);
    const uint32_t slab0 = smem_u32(slab), b0 = smem_u32(smem_b);
    const uint32_t b_full0 = smem_u32(b_full), b_empty0 = smem_u32(b_empty);
    const uint32_t tmem_full0 = smem_u32(tmem_full), tmem_empty0 = smem_u32(tmem_empty);
    const uint32_t slab_full0 = smem_u32(slab_full), slab_empty0 = smem_u32(slab_empty);
    uint32_t bs = 0, bph = 0, tile = 0, chunk = 0;
    for (int img = blockIdx.x; img < Nimg; img += gridDim.x, ++tile) {
      C2T_SET(tq);
      C2T_PUT(3, tq - t_entry);                              // prologue: entry -> ready to wait for the slab
      mbar_wait_addr(slab_full0, tile & 1);
      C2T_ADD(tw0, tq);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      for (int tap = 0; tap < 25; ++tap) {
        const bool chunk_start = tap % chunk_taps == 0;
        const uint32_t acc = chunk & 1;
        if (chunk_start) {
          C2T_SET(tq);
          mbar_wait_addr(tmem_empty0 + acc * 8, ((chunk >> 1) & 1) ^ 1);
          C2T_ADD(tw1, tq);
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        }
        C2T_SET(tq);
        mbar_wait_addr(b_full0 + bs * 8, bph);
        C2T_ADD(tw2, tq);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const int r = tap / 5, sx = tap - r * 5;
        const uint32_t shift = static_cast<uint32_t>(DGRAD ? (4 - r) * kC2PW + (4 - sx) : r * kC2PW + sx) * 8u;   // 16-B units
        const uint32_t bw = desc_lo(b0 + bs * kBStage, DGRAD ? 16 : 4096);
        if (elect_one()) {
#pragma unroll
          for (int cb = 0; cb < CBLOCKS; ++cb) {
            // A: slab [plane 0 | plane 1][cb]; B: [W | W_lo] of the tap (forward) / of K block cb of the tap (dgrad)
            const uint32_t ax = desc_lo(slab0 + cb * kC2SlabBytes, 16) + shift;
            const uint32_t ax_lo = desc_lo(slab0 + (CBLOCKS + cb) * kC2SlabBytes, 16) + shift;
            const uint32_t bx = bw + (DGRAD ? cb * (8192 >> 4) : 0);
#pragma unroll
            for (int h = 0; h < 2; ++h) {
              const uint32_t d = tmem_base + acc * kAccCols + h * 2 * N_OUT;
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                const uint32_t ak = h * 1024 + 2 * k, bk = kStepB * k;
                umma2<1>(d, ax + ak, hi_a, bx + bk, hi_b, idesc_wide, !(chunk_start && cb == 0 && k == 0));   // main | A*W_lo
                umma2<1>(d + N_OUT, ax_lo + ak, hi_a, bx + bk, hi_b, idesc_narrow, true);                      // += A_lo*W
              }
            }
          }
          umma_commit_addr(b_empty0 + bs * 8);
        }
        __syncwarp();
        if (++bs == kBStages) { bs = 0; bph ^= 1; }
        if ((tap + 1) % chunk_taps == 0 || tap == 24) {
          if (elect_one()) {
            umma_commit_addr(tmem_full0 + acc * 8);
            if (tap == 24) umma_commit_addr(slab_empty0);
          }
          __syncwarp();
          ++chunk;
        }
      }
    }
    C2T_PUT(4, tw0); C2T_PUT(5, tw1); C2T_PUT(6, tw2); C2T_PUT(7, C2T_NOW() - t_entry);
  } else {
    // ===== accumulate in registers + epilogue: warp = (TMEM lane quarter q, M tile h), thread = one grid position =====
    const int q = warp & 3, h = (warp - 2) >> 2;
    const uint32_t lane_base = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + h * 2 * N_OUT;
    const int m = h * 128 + q * 32 + lane;
    const int oh = m / kC2PW, ow = m - oh * kC2PW;
    const bool valid = oh < kC2H && ow < kC2W;
    uint32_t chunk = 0;
    for (int img = blockIdx.x; img < Nimg; img += gridDim.x) {
      // Nothing but stores is left for after the last chunk: the running sum starts at the bias, and the
      // mask (data gradient: input activation > 0) is fetched now, as one bit per channel.
      const size_t row = valid ? (static_cast<size_t>(img) * kC2H + oh) * kC2W + ow : 0;
      float sum[N_OUT];
#pragma unroll
      for (int u = 0; u < N_OUT; ++u) sum[u] = bias ? __ldg(bias + u) : 0.f;
      uint32_t keep = 0xffffffffu;
      if (DGRAD && mask && valid) {
        keep = 0;
#pragma unroll
        for (int u = 0; u < N_OUT && u < 32; u += 4) {
          const float4 k4 = *reinterpret_cast<const float4*>(mask + row * N_OUT + u);
          keep |= (k4.x > 0.f ? 1u : 0u) << u | (k4.y > 0.f ? 2u : 0u) << u | (k4.z > 0.f ? 4u : 0u) << u | (k4.w > 0.f ? 8u : 0u) << u;
        }
      }
      for (int c = 0; c < chunks_per_tile; ++c, ++chunk) {
        const uint32_t acc = chunk & 1;
        C2T_SET(tq);
        mbar_wait(&tmem_full[acc], (chunk >> 1) & 1);
        C2T_ADD(tw0, tq);
        if (c == chunks_per_tile - 1) { C2T_SET(tw2); }         // last chunk of the image complete
        C2T_SET(tq);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
        for (int c0 = 0; c0 < 2 * N_OUT; c0 += 32) {           // main columns, then the small ones
          uint32_t rr[32];
          tmem_ld32(lane_base + acc * kAccCols + c0, rr);
#pragma unroll
          for (int u = 0; u < 32; ++u) sum[(c0 + u) % N_OUT] = __fadd_rn(sum[(c0 + u) % N_OUT], __uint_as_float(rr[u]));
        }
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncwarp();
        if (lane == 0) mbar_arrive(&tmem_empty[acc]);
        C2T_ADD(tw1, tq);
      }
      if (!DGRAD && pool_a) {
        // Forward with the 2x2 max-pool + relu in the epilogue (a launch of its own until the end of round 2; same
        // comparisons in the same order as pool_relu_kernel of the other math modes): the outputs go to shared memory, one line per position, and the eight
        // epilogue warps then share out the 49 pooled positions x 16 channel quads.  The convolution output itself is
        // never written: the backward pass needs the pooled activations and the arg-max codes only.
        if (valid) {
          float* line = pool_stage + (oh * kC2W + ow) * kC2PoolPitch;
#pragma unroll
          for (int u = 0; u < N_OUT; u += 4) *reinterpret_cast<float4*>(line + u) = make_float4(sum[u], sum[u + 1], sum[u + 2], sum[u + 3]);
        }
        asm volatile("bar.sync 2, 256;" ::: "memory");
        for (int i = threadIdx.x - 64; i < 49 * 16; i += 256) {
          const int q4 = i & 15, pos = i >> 4, ph = pos / 7, pw = pos - ph * 7;
          float best[4];
          int arg[4];
#pragma unroll
          for (int dy = 0; dy < 2; ++dy)
#pragma unroll
            for (int dx = 0; dx < 2; ++dx) {
              const float4 v4 = *reinterpret_cast<const float4*>(pool_stage + ((2 * ph + dy) * kC2W + 2 * pw + dx) * kC2PoolPitch + q4 * 4);
              const float v[4] = {v4.x, v4.y, v4.z, v4.w};
#pragma unroll
              for (int u = 0; u < 4; ++u)
                if ((dy | dx) == 0 || v[u] > best[u]) { best[u] = v[u]; arg[u] = dy * 2 + dx; }
            }
          float r[4];
#pragma unroll
          for (int u = 0; u < 4; ++u) r[u] = fmaxf(best[u], 0.f);
          const size_t o = (static_cast<size_t>(img) * 49 + pos) * 64 + q4 * 4;
          *reinterpret_cast<float4*>(pool_a + o) = make_float4(r[0], r[1], r[2], r[3]);
          *reinterpret_cast<float4*>(pool_lo + o) = make_float4(tf32_lo(r[0]), tf32_lo(r[1]), tf32_lo(r[2]), tf32_lo(r[3]));
          *reinterpret_cast<uchar4*>(pool_idx + o) = make_uchar4((unsigned char)arg[0], (unsigned char)arg[1], (unsigned char)arg[2],
                                                                (unsigned char)arg[3]);
        }
        asm volatile("bar.sync 2, 256;" ::: "memory");     // the staging lines are free for the next image
      } else if (valid) {
        float* orow = out + row * N_OUT;
#pragma unroll
        for (int u = 0; u < N_OUT; u += 4) {
          float4 o = make_float4(sum[u], sum[u + 1], sum[u + 2], sum[u + 3]);
          if (DGRAD) {
            o.x = (keep >> (u & 31)) & 1u ? o.x : 0.f; o.y = (keep >> ((u + 1) & 31)) & 1u ? o.y : 0.f;
            o.z = (keep >> ((u + 2) & 31)) & 1u ? o.z : 0.f; o.w = (keep >> ((u + 3) & 31)) & 1u ? o.w : 0.f;
          }
          *reinterpret_cast<float4*>(orow + u) = o;
        }
      }
    }
    if (warp == 2) { C2T_PUT(8, tw0); C2T_PUT(9, tw1); C2T_PUT(10, tw2 - t_entry); C2T_PUT(11, C2T_NOW() - t_entry); }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem_base), "n"(kTmemCols) : "memory");
  }
  if (threadIdx.x == 0) C2T_PUT(12, C2T_NOW() - t_entry);
}

// ---- weight gradient ---------------------------------------------------------------------------------
// dW[(r,s,c), n] = sum over pixels: the pixels are the K dimension and both operands are MN-major
// (a 128-byte line = the 32 channels of ONE pixel; tf32 MN-major = 128-byte swizzle with 32-byte base,
// 4 lines per 512-byte atom).  A pixel tile is 5 image rows of the 18-column grid = 90 positions, padded
// to 96 (12 K steps of 8): dy arrives as boxes {32 n, 18 columns, 5 rows} -- columns 14..17 and rows
// past the image zero-filled, so grid positions that are not pixels contribute nothing -- and x as
// ONE slab {32 c, 18 columns from -2, 9 rows from h0 - 2} per plane.  The A operand of tap (r, s) is
// that slab from line r * 18 + s on; an M tile of 128 rows = four taps = four 32-row groups whose
// start lines differ by a constant, which is exactly the descriptor's leading byte offset:
//     tile r (r = 0..4): taps (r, 0..3), first line r * 18, group stride 1 line
//     tile 5:            taps (0..3, 4), first line 4,      group stride 18 lines
//     tile 6:            tap  (4, 4),    first line 76      (one group)
// A CTA owns one group of M tiles {0,1} {2,3} {4,5} {6} and every nz-th pixel tile; it writes its
// partial sum to partial[z][800][64], and wgrad_reduce_kernel adds the partials in z order.
// The main product alternates between two accumulators per pixel tile (12 truncating adds each).
constexpr int kWTileRows = 5, kWTilesPerImage = 3;
constexpr int kWSlabLines = 176;                            // box 9 * 18 = 162; reads reach line 95 + 76
constexpr int kWSlabBytes = kWSlabLines * 128;              // 22 * 1024
constexpr int kWDyLines = 96;                               // box 5 * 18 = 90, padded to 12 K steps
constexpr int kWDyBytes = kWDyLines * 128;                  // 12 * 1024
constexpr int kWStageBytes = 2 * kWSlabBytes + 4 * kWDyBytes;   // [x | x_lo] [dy n0, dy n1 | dy_lo n0, dy_lo n1]
constexpr int kWStages = 2;
constexpr size_t smem_bytes_c2_wgrad() { return 1024 + kWStages * kWStageBytes + (2 * kWStages + 8) * sizeof(uint64_t) + 16; }
static_assert(smem_bytes_c2_wgrad() <= 232448, "shared memory budget");

struct WgradPlan {
  int cta0[5];        // CTAs [cta0[g], cta0[g + 1]) work on M-tile group g
};

__host__ __device__ inline int c2_wgrad_group_of_row(int row) {   // dW row (tap * 32 + c) -> M-tile group
  const int tap = row >> 5, r = tap / 5, s = tap - r * 5;
  const int t = s < 4 ? r : (r < 4 ? 5 : 6);
  return t >> 1;
}

__global__ void __launch_bounds__(kC2Threads, 1)
conv2_wgrad_x3_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_x_lo,
                      const __grid_constant__ CUtensorMap map_dy, const __grid_constant__ CUtensorMap map_dy_lo,
                      float* __restrict__ partial, int Nimg, const __grid_constant__ WgradPlan plan) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + kWStages * kWStageBytes);
  uint64_t* empty = full + kWStages;
  uint64_t* tmem_full = empty + kWStages;     // [2]
  uint64_t* tmem_empty = tmem_full + 2;       // [2]
  uint64_t* s_full = tmem_empty + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(s_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int g = 0;
  while (g < 3 && static_cast<int>(blockIdx.x) >= plan.cta0[g + 1]) ++g;
  const int z = blockIdx.x - plan.cta0[g], nz = plan.cta0[g + 1] - plan.cta0[g];
  const int n_mt = g < 3 ? 2 : 1;
  const int total_tiles = Nimg * kWTilesPerImage;
  const int my_tiles = z < total_tiles ? (total_tiles - z + nz - 1) / nz : 0;

  // lines the TMA boxes never write but the MMAs read: zero them once
  for (int i = threadIdx.x; i < kWStages * (2 * (kWSlabLines - 162) + 4 * (kWDyLines - 90)) * 32; i += kC2Threads) {
    constexpr int kPerStage = (2 * (kWSlabLines - 162) + 4 * (kWDyLines - 90)) * 32;
    const int st = i / kPerStage;
    int w = i % kPerStage;
    uint8_t* base = smem + st * kWStageBytes;
    float* dst;
    if (w < 2 * (kWSlabLines - 162) * 32) {
      const int pl = w / ((kWSlabLines - 162) * 32);
      w -= pl * (kWSlabLines - 162) * 32;
      dst = reinterpret_cast<float*>(base + pl * kWSlabBytes + 162 * 128) + w;
    } else {
      w -= 2 * (kWSlabLines - 162) * 32;
      const int b = w / ((kWDyLines - 90) * 32);
      w -= b * (kWDyLines - 90) * 32;
      dst = reinterpret_cast<float*>(base + 2 * kWSlabBytes + b * kWDyBytes + 90 * 128) + w;
    }
    *dst = 0.f;
  }
  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" :: "l"(reinterpret_cast<uint64_t>(&map_x)) : "memory");
    asm volatile("prefetch.tensormap [%0];" :: "l"(reinterpret_cast<uint64_t>(&map_x_lo)) : "memory");
    asm volatile("prefetch.tensormap [%0];" :: "l"(reinterpret_cast<uint64_t>(&map_dy)) : "memory");
    asm volatile("prefetch.tensormap [%0];" :: "l"(reinterpret_cast<uint64_t>(&map_dy_lo)) : "memory");
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int s = 0; s < kWStages; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
      for (int a = 0; a < 2; ++a) { mbar_init(&tmem_full[a], 1); mbar_init(&tmem_empty[a], 4 * n_mt); }
      mbar_init(s_full, 1);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                 :: "r"(smem_u32(tmem_slot)), "n"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = *tmem_slot;
  // TMEM columns: S of M tile i at i * 64; main accumulator a of M tile i at 128 * (1 + a) + i * 64

  if (warp == 0) {
    if (lane == 0) {
      for (int it = 0; it < my_tiles; ++it) {
        const int t = z + it * nz, img = t / kWTilesPerImage, h0 = (t % kWTilesPerImage) * kWTileRows;
        const int s = it % kWStages;
        mbar_wait(&empty[s], ((it / kWStages) & 1) ^ 1);
        mbar_expect_tx(&full[s], 2 * 162 * 128 + 4 * 90 * 128);
        uint8_t* base = smem + s * kWStageBytes;
        tma_load_4d(base, &map_x, &full[s], 0, -2, h0 - 2, img);
        tma_load_4d(base + kWSlabBytes, &map_x_lo, &full[s], 0, -2, h0 - 2, img);
#pragma unroll
        for (int b = 0; b < 4; ++b)
          tma_load_4d(base + 2 * kWSlabBytes + b * kWDyBytes, (b & 2) ? &map_dy_lo : &map_dy, &full[s], (b & 1) * 32, 0, h0, img);
      }
    }
  } else if (warp == 1) {
    constexpr uint32_t idesc = make_idesc(2, 64, true, true);
    constexpr uint32_t hi = desc_hi(512, 1);                 // MN-major tf32: 32-byte-base swizzle, 4-line atoms
    const uint32_t smem0 = smem_u32(smem);
    const uint32_t full0 = smem_u32(full), empty0 = smem_u32(empty);
    const uint32_t tmem_full0 = smem_u32(tmem_full), tmem_empty0 = smem_u32(tmem_empty);
    uint32_t a_shift[2], a_lbo[2];                           // first line and group stride (bytes) of this CTA's M tiles
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const int t = 2 * g + i;
      a_shift[i] = t < 5 ? t * kC2PW : (t == 5 ? 4 : 76);
      a_lbo[i] = t == 5 ? kC2PW * 128 : 128;
    }
    for (int it = 0; it < my_tiles; ++it) {
      const uint32_t st = it % kWStages, acc = it & 1;
      mbar_wait_addr(tmem_empty0 + acc * 8, ((it >> 1) & 1) ^ 1);
      mbar_wait_addr(full0 + st * 8, (it / kWStages) & 1);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const uint32_t xs = smem0 + st * kWStageBytes;
      const uint32_t b_raw = desc_lo(xs + 2 * kWSlabBytes, kWDyBytes), b_low = desc_lo(xs + 2 * kWSlabBytes + 2 * kWDyBytes, kWDyBytes);
      if (elect_one()) {
        for (int i = 0; i < n_mt; ++i) {
          const uint32_t a_raw = desc_lo(xs + a_shift[i] * 128, a_lbo[i]), a_low = desc_lo(xs + kWSlabBytes + a_shift[i] * 128, a_lbo[i]);
          const uint32_t d_small = tmem_base + i * 64, d_main = tmem_base + 128 * (1 + acc) + i * 64;
#pragma unroll 4
          for (int ks = 0; ks < kWDyLines / 8; ++ks) {       // 8 pixels = 8 lines = 1024 B per instruction
            const uint32_t o = 64u * ks;
            umma2<1>(d_small, a_low + o, hi, b_raw + o, hi, idesc, !(it == 0 && ks == 0));
            umma2<1>(d_small, a_raw + o, hi, b_low + o, hi, idesc, true);
            umma2<1>(d_main, a_raw + o, hi, b_raw + o, hi, idesc, ks != 0);
          }
        }
        umma_commit_addr(empty0 + st * 8);
        umma_commit_addr(tmem_full0 + acc * 8);
        if (it == my_tiles - 1) umma_commit_addr(smem_u32(s_full));
      }
      __syncwarp();
    }
  } else if (((warp - 2) >> 2) < n_mt) {
    // ===== accumulate + store: warp = (row group q = one tap, M tile i), lane = input channel =====
    const int q = warp & 3, i = (warp - 2) >> 2;
    const uint32_t lane_base = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + i * 64;
    float sum[64];
#pragma unroll
    for (int u = 0; u < 64; ++u) sum[u] = 0.f;
    for (int it = 0; it < my_tiles; ++it) {
      const uint32_t acc = it & 1;
      mbar_wait(&tmem_full[acc], (it >> 1) & 1);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
      for (int c0 = 0; c0 < 64; c0 += 32) {
        uint32_t rr[32];
        tmem_ld32(lane_base + 128 * (1 + acc) + c0, rr);
#pragma unroll
        for (int u = 0; u < 32; ++u) sum[c0 + u] = __fadd_rn(sum[c0 + u], __uint_as_float(rr[u]));
      }
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty[acc]);
    }
    if (my_tiles > 0) {
      mbar_wait(s_full, 0);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
      for (int c0 = 0; c0 < 64; c0 += 32) {
        uint32_t rr[32];
        tmem_ld32(lane_base + c0, rr);
#pragma unroll
        for (int u = 0; u < 32; ++u) sum[c0 + u] = __fadd_rn(sum[c0 + u], __uint_as_float(rr[u]));
      }
    }
    const int t = 2 * g + i;
    const int tap = t < 5 ? t * 5 + q : (t == 5 ? q * 5 + 4 : (q == 0 ? 24 : -1));
    if (tap >= 0) {
      float* dst = partial + (static_cast<size_t>(z) * 800 + tap * 32 + lane) * 64;
#pragma unroll
      for (int u = 0; u < 64; u += 4) *reinterpret_cast<float4*>(dst + u) = make_float4(sum[u], sum[u + 1], sum[u + 2], sum[u + 3]);
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem_base), "n"(512) : "memory");
  }
}

// dW[row, :] = sum over the partials of the row's M-tile group.  Four lanes share an output float4: lane l
// adds the partials of its quarter of the z range in z order, and the four lane sums are combined as
// (l0 + l1) + (l2 + l3) -- a fixed order, so the result is reproducible; four times the CTAs and a
// quarter of the dependent-load chain of one thread per output (this reduction is latency-bound).
__global__ void __launch_bounds__(256)
conv2_wgrad_reduce_kernel(const float* __restrict__ partial, float* __restrict__ dw, const __grid_constant__ WgradPlan plan) {
  const int gi = blockIdx.x * 256 + threadIdx.x;
  const int i = gi >> 2, l = gi & 3;                          // float4 index into [800][64], lane of the group
  const bool live = i < 800 * 16;
  float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
  if (live) {
    const int g = c2_wgrad_group_of_row(i >> 4);
    const int nz = plan.cta0[g + 1] - plan.cta0[g];
    const int per = (nz + 3) >> 2;
    for (int zz = l * per; zz < min(nz, (l + 1) * per); ++zz) {
      const float4 v = reinterpret_cast<const float4*>(partial)[static_cast<size_t>(zz) * (800 * 16) + i];
      s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
    }
  }
  // (l0 + l1), (l2 + l3), then their sum; lanes of a group are adjacent threads of one warp
  float4 o;
  o.x = __shfl_xor_sync(0xffffffffu, s.x, 1); o.y = __shfl_xor_sync(0xffffffffu, s.y, 1);
  o.z = __shfl_xor_sync(0xffffffffu, s.z, 1); o.w = __shfl_xor_sync(0xffffffffu, s.w, 1);
  if (l & 1) { float4 t = s; s = o; o = t; }                   // s = the even lane's sum, o = the odd lane's: same operand order in both lanes
  s.x += o.x; s.y += o.y; s.z += o.z; s.w += o.w;
  o.x = __shfl_xor_sync(0xffffffffu, s.x, 2); o.y = __shfl_xor_sync(0xffffffffu, s.y, 2);
  o.z = __shfl_xor_sync(0xffffffffu, s.z, 2); o.w = __shfl_xor_sync(0xffffffffu, s.w, 2);
  if (l & 2) { float4 t = s; s = o; o = t; }
  s.x += o.x; s.y += o.y; s.z += o.z; s.w += o.w;
  if (live && l == 0) reinterpret_cast<float4*>(dw)[i] = s;
}

}  // namespace tc
}  // namespace tfg
