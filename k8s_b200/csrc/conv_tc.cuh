// Implicit-GEMM convolutions on tcgen05: the im2col matrix is never materialised.  The A operand of
// each K block is fetched by ONE 4-D TMA box over the NHWC activation -- {64 channels, TW columns,
// TH rows, TN images} at the coordinates of filter tap (r, s), element strides = convolution
// stride, out-of-bounds (padding) zero-filled by the TMA unit -- and lands in shared memory as the
// K-major (forward / dgrad) or MN-major (wgrad) 128-byte-swizzled tile tcgen05.mma reads.
//
//   forward  z[pix, k]      = sum_{r,s,c} x[pix + (r,s), c] * W[(r,s,c), k]        A = x boxes, B = W MN-major
//   dgrad    dx[pix, c]     = sum_{r,s,k} dz[pix - (r,s), k] * W[(r,s,c), k]       A = dz boxes, B = W K-major  (stride 1)
//   wgrad    dW[(r,s,c), k] = sum_{pix}   x[pix + (r,s), c] * dz[pix, k]           A = x boxes (MN), B = dz boxes (MN)
//
// A tile of output pixels is TN images x TH rows x the full row width (TW = OW), at most 128, so its
// rows are contiguous rows of the [pixels, channels] output and the GEMM epilogue applies unchanged.
// Persistent CTAs, 4-stage (3 for wgrad) TMA ring, two TMEM accumulators (see gemm_tc.cuh).
#pragma once
#include "gemm_tc.cuh"

namespace tfg {
namespace tc {

struct ConvShape {
  int Nimg, OH, OW;          // extent of the pixel space the tiles walk (output for fwd/wgrad, input for dgrad)
  int TN, TH;                // tile = TN images x TH rows x OW columns
  int h_tiles, m_tiles;      // ceil(OH / TH), img_groups * h_tiles
  int R, S, stride, pad;
  int C;                     // channels of the gathered tensor (K dimension per tap), multiple of 64
  int flip;                  // 0: tap offset (+r - pad, +s - pad) scaled by stride (forward / wgrad); 1: (pad - r, pad - s) (dgrad)
  // conv_gemm_kernel only.  ntaps > 0: an explicit tap list replaces the R x S walk -- tap t gathers
  // at offset (tap_dh[t], tap_dw[t]) and multiplies by filter tap tap_w[t] (row block tap_w[t] of the
  // weight matrix).  The tile at pixel rows h0.. is stored at rows h0 * out_stride + out_h0 (columns
  // from out_w0) of the output map, whose element strides are out_stride: one parity class of a
  // stride-2 data gradient, dx[2i + ph, 2j + pw] = sum_t dz[i + dh_t, j + dw_t] W[tap_t].
  int ntaps;
  signed char tap_dh[9], tap_dw[9], tap_w[9];
  int out_stride, out_h0, out_w0;
};

__device__ __forceinline__ void tma_load_4d(void* smem_dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1,
                                            int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
      :: "r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2),
         "r"(c3)
      : "memory");
}

__device__ __forceinline__ void conv_tile_origin(const ConvShape& cs, int mt, int* n0, int* h0) {
  *n0 = (mt / cs.h_tiles) * cs.TN;
  *h0 = (mt % cs.h_tiles) * cs.TH;
}

template <int BLOCK_N>
constexpr size_t smem_bytes_conv() { return smem_bytes_persistent<BLOCK_N>(); }   // same layout as the persistent GEMM

// ---- forward (B_MN = true) and stride-1 data gradient (B_MN = false) ---------------------------------
// map_a: 4-D map over the gathered activation [Nimg_in, H_in, W_in, C], box {64, TW*stride, TH*stride, TN}.
// map_b: 2-D map over W viewed [R*S*C_w, Kout]: forward reads it MN-major (rows = K), dgrad K-major (rows = N).
// Output [pixels, N] row-major (bf16 through epi.out_bf16, or fp32 C), ldc = N; with epi.tma_out the
// bf16 tile leaves through map_c, a 4-D map over the output tensor [Nimg, OH, OW, N] with box
// {64, OW, TH, TN}: the same pixel tile as the A box, so rows past the image are clipped by the TMA unit.
template <int BLOCK_N, bool B_MN>
__global__ void __launch_bounds__(kThreadsPersistent, 1)
conv_gemm_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                 const __grid_constant__ CUtensorMap map_c, float* __restrict__ Cout, const __grid_constant__ ConvShape cs, int N, int ldc, int tiles_n,
                 int b_rows_per_tap, Epilogue epi) {
  constexpr int kStages = persistent_stages<BLOCK_N>();
  constexpr int kBlockK = 64;                                   // bf16 elements of K per stage = one channel block
  constexpr int kABytes = kBlockM * kBlockKBytes, kBBytes = BLOCK_N * kBlockKBytes;
  constexpr int kChunkBytes = kBlockK * 128;
  constexpr uint32_t kStepB = B_MN ? (16 * 128) >> 4 : 32 >> 4;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + kStages * kABytes;
  constexpr int kOutTileBytes = (BLOCK_N / 64) * kOutHalfBytes;
  uint8_t* out_stage = smem + kStages * (kABytes + kBBytes);
  float* stage_base = reinterpret_cast<float*>(out_stage);             // fallback epilogue only
  uint64_t* full = reinterpret_cast<uint64_t*>(out_stage + 2 * kOutTileBytes);
  uint64_t* empty = full + kStages;
  uint64_t* tmem_full = empty + kStages;
  uint64_t* tmem_empty = tmem_full + 2;
  uint64_t* img_full = tmem_empty + 2;       // [2] output tile image staged (column-sum warps may read it)
  uint64_t* img_free = img_full + 2;         // [2] column sums of the image taken
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(img_free + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int total_tiles = cs.m_tiles * tiles_n;
  const int cblocks = cs.C / 64;
  const int k_blocks = (cs.ntaps > 0 ? cs.ntaps : cs.R * cs.S) * cblocks;
  const int box_rows = cs.TN * cs.TH * cs.OW;                   // rows one A box writes (<= 128)

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
    if (lane == 0) {
      uint32_t it = 0;
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
        const int n0 = (t % tiles_n) * BLOCK_N, mt = t / tiles_n;      // n fastest: the CTAs sharing an A box run together
        int img0, h0;
        conv_tile_origin(cs, mt, &img0, &h0);
        for (int kb = 0; kb < k_blocks; ++kb, ++it) {
          const int s = it % kStages;
          const uint32_t phase = (it / kStages) & 1;
          const int tap = kb / cblocks, cb = kb - tap * cblocks;
          const int r = tap / cs.S, sx = tap - r * cs.S;
          mbar_wait(&empty[s], phase ^ 1);
          mbar_expect_tx(&full[s], box_rows * 128 + kBBytes);
          int ch = cs.flip ? cs.pad - r : r - cs.pad, cw = cs.flip ? cs.pad - sx : sx - cs.pad, wtap = tap;
          if (cs.ntaps > 0) { ch = cs.tap_dh[tap]; cw = cs.tap_dw[tap]; wtap = cs.tap_w[tap]; }
          tma_load_4d(smem_a + s * kABytes, &map_a, &full[s], cb * 64, cw, h0 * cs.stride + ch, img0);
          const int krow = wtap * b_rows_per_tap;               // first row of this tap in W viewed [R*S*C_w, Kout]
          if (B_MN) {                                           // rows = K (channel block of the tap), columns = n
#pragma unroll
            for (int c = 0; c < BLOCK_N / 64; ++c)
              tma_load_2d(smem_b + s * kBBytes + c * kChunkBytes, &map_b, &full[s], n0 + c * 64, krow + cb * 64);
          } else {                                              // rows = n (input channel c of the tap), columns = K
            tma_load_2d(smem_b + s * kBBytes, &map_b, &full[s], cb * 64, krow + n0);
          }
        }
      }
    }
  } else if (warp == 1) {
    // converged warp, one elected lane issues (gemm_tc.cuh, "lean issue path")
    constexpr uint32_t idesc = make_idesc(1, BLOCK_N, false, B_MN);
    constexpr uint32_t hi = desc_hi();
    const uint32_t a0 = smem_u32(smem_a), b0 = smem_u32(smem_b);
    const uint32_t full0 = smem_u32(full), empty0 = smem_u32(empty);
    const uint32_t tmem_full0 = smem_u32(tmem_full), tmem_empty0 = smem_u32(tmem_empty);
    uint32_t st = 0, ph = 0, acc = 0, acc_ph = 0;
    for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
      mbar_wait_addr(tmem_empty0 + acc * 8, acc_ph ^ 1);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const uint32_t d_tmem = tmem_base + acc * BLOCK_N;
      for (int kb = 0; kb < k_blocks; ++kb) {
        mbar_wait_addr(full0 + st * 8, ph);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint32_t a_lo = desc_lo(a0 + st * kABytes, 16), b_lo = desc_lo(b0 + st * kBBytes, B_MN ? kChunkBytes : 16);
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < 4; ++k) umma2<0>(d_tmem, a_lo + 2 * k, hi, b_lo + kStepB * k, hi, idesc, kb > 0 || k > 0);
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
    uint32_t tcnt = 0;
    for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, ++tcnt) {
      const int n0 = (t % tiles_n) * BLOCK_N, mt = t / tiles_n;      // n fastest: the CTAs sharing an A box run together
      int img0, h0;
      conv_tile_origin(cs, mt, &img0, &h0);
      const long long m_start = ((long long)img0 * cs.OH + h0) * cs.OW;
      const int valid = (cs.TN > 1 ? min(cs.TN, cs.Nimg - img0) * cs.OH : min(cs.TH, cs.OH - h0)) * cs.OW;
      const uint32_t acc = tcnt & 1, acc_phase = (tcnt >> 1) & 1;
      if (epi.tma_out) {
        uint8_t* img = out_stage + acc * kOutTileBytes;
        if (warp == 2 && lane == 0) bulk_wait_read<1>();    // the store of two tiles ago has read this image
        if (epi.col_stats && tcnt >= 2) mbar_wait(&img_free[acc], ((tcnt >> 1) - 1) & 1);   // ... and so have the column sums
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
            if (n0 + h * 64 < N)
              tma_store_4d(&map_c, img + h * kOutHalfBytes, n0 + h * 64, cs.out_w0, h0 * cs.out_stride + cs.out_h0, img0,
                           epi.tma_out == 2);
          bulk_commit();
          if (epi.col_stats) mbar_arrive(&img_full[acc]);
        }
        continue;
      }
      mbar_wait(&tmem_full[acc], acc_phase);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      // rows of the tile are rows m_start .. m_start + valid of the [pixels, N] output
      float* cb = Cout ? Cout + m_start * ldc : nullptr;
      Epilogue e = epi;
      if (e.out_bf16) e.out_bf16 = static_cast<__nv_bfloat16*>(e.out_bf16) + m_start * ldc;
      drain_tile<BLOCK_N>(tmem_base + acc * BLOCK_N, true, stage_base, warp, lane, cb, valid, N, ldc, 0, n0, 0, e);
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty[acc]);
    }
    if (epi.tma_out && warp == 2 && lane == 0) bulk_wait_all();   // shared memory must outlive the last store
  } else if (epi.tma_out && epi.col_stats) {
    // ===== column-sum warps: batch-norm statistics of the staged tiles, off the epilogue's critical path =====
    uint32_t tcnt = 0;
    ColStats sums;
    const int et = threadIdx.x - kThreads;
    for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, ++tcnt) {
      const int mt = t / tiles_n;
      int img0, h0;
      conv_tile_origin(cs, mt, &img0, &h0);
      const int valid = (cs.TN > 1 ? min(cs.TN, cs.Nimg - img0) * cs.OH : min(cs.TH, cs.OH - h0)) * cs.OW;
      const uint32_t acc = tcnt & 1;
      mbar_wait(&img_full[acc], (tcnt >> 1) & 1);
      tile_col_stats<BLOCK_N>(out_stage + acc * kOutTileBytes, et, valid, sums);
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

// ---- 3x3 stride-1 convolutions with the input tile resident in shared memory -----------------------
// The kernel above fetches the A operand of every tap through L2 again: nine times the same pixels,
// and L2 -> SM bandwidth is what bounds it.  Here a tile is TH full image rows; the producer loads,
// per 64-channel block, ONE slab {64 ch, PW = OW + 2 columns from -1, TH + 2 rows from h0 - 1}
// (borders zero-filled by the TMA unit): (TH + 2) * PW positions of 128 B, contiguous, position
// q = hh * PW + ww.  Output row m of the MMA stands for (oh, ow') = (m / PW, m % PW); its input at
// tap (r, s) is slab position m + r * PW + s, so the A operand of a tap is the SAME slab read from a
// start address (r * PW + s) * 128 B further on -- a shifted K-major 128B-swizzled matrix (the swizzle
// phase of the new start goes into the descriptor's base-offset field).  Rows with ow' >= OW or
// beyond the tile's rows are not outputs: the epilogue stages zeros for them and the TMA store's box
// {64, PW, TH} clips them at the tensor edge.  dgrad (flip): the taps are mirrored.
struct HaloShape {
  int Nimg, OH, OW, PW, TH, h_tiles, m_tiles, C, flip;
  int MT;                    // M tiles of 128 positions per CTA (host-side choice, selects the kernel instance)
};

// MT = M tiles (of 128 positions) per CTA.  MT = 2: a tile is 256 consecutive positions, the two
// accumulators share every weight tile (the weight tiles dominate L2 -> SM traffic once the input is
// resident: 72 -> 42 B/clk/SM at full MMA rate, against an L2 limit of ~42), one 256-row output image.
constexpr int kSlabStages = 2;
template <int MT> constexpr int halo_slab_bytes() {          // >= (128 MT + 2 PW + 2) * 128
  return MT == 1 ? 32 * 1024 : 47 * 1024;                    // PW <= 62 / PW <= 58
}
template <int MT> constexpr int halo_max_pw() { return MT == 1 ? 62 : 58; }
template <int BLOCK_N, int MT> constexpr int halo_b_stages() { return MT == 1 ? (BLOCK_N == 64 ? 8 : 6) : (BLOCK_N == 64 ? 8 : 4); }

template <int BLOCK_N, int MT>
constexpr size_t smem_bytes_halo() {
  return 1024 + kSlabStages * halo_slab_bytes<MT>() +
         static_cast<size_t>(halo_b_stages<BLOCK_N, MT>()) * BLOCK_N * kBlockKBytes + 2 * (BLOCK_N / 64) * kOutHalfBytes +
         (2 * kSlabStages + 2 * halo_b_stages<BLOCK_N, MT>() + 8) * sizeof(uint64_t) + 16;
}
static_assert(smem_bytes_halo<128, 1>() <= 232448 && smem_bytes_halo<64, 1>() <= 232448 &&
              smem_bytes_halo<128, 2>() <= 232448 && smem_bytes_halo<64, 2>() <= 232448, "shared memory budget");

// map_a: 4-D map over the gathered activation [Nimg, OH, OW, C], box {64, PW, TH + 2, 1}.
// map_b: as conv_gemm_kernel.  map_c: 4-D map over the output [Nimg, OH, OW, N], box {64, PW, TH, 1}.
template <int BLOCK_N, bool B_MN, int MT>
__global__ void __launch_bounds__(kThreadsPersistent, 1)
conv3x3_halo_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                    const __grid_constant__ CUtensorMap map_c, const __grid_constant__ HaloShape hs, int N, int tiles_n,
                    int b_rows_per_tap, int base_offset_mode, Epilogue epi) {
  constexpr int kBStages = halo_b_stages<BLOCK_N, MT>();
  constexpr int kSlabBytes = halo_slab_bytes<MT>();
  constexpr int kRows = kBlockM * MT;                     // positions per tile
  constexpr int kHalfBytes = kRows * 128;                 // one 64-column half of the output image
  constexpr int kImages = MT == 1 ? 2 : 1;                // output images in the staging area
  constexpr int kAccCols = MT * BLOCK_N;                  // TMEM columns of one accumulator set
  constexpr int kBBytes = BLOCK_N * kBlockKBytes;
  constexpr int kChunkBytes = 64 * 128;
  constexpr uint32_t kStepB = B_MN ? (16 * 128) >> 4 : 32 >> 4;
  constexpr int kOutTileBytes = (BLOCK_N / 64) * kHalfBytes;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* slab = smem;
  uint8_t* smem_b = slab + kSlabStages * kSlabBytes;
  uint8_t* out_stage = smem_b + kBStages * kBBytes;
  uint64_t* slab_full = reinterpret_cast<uint64_t*>(out_stage + kImages * kOutTileBytes);
  uint64_t* slab_empty = slab_full + kSlabStages;
  uint64_t* b_full = slab_empty + kSlabStages;
  uint64_t* b_empty = b_full + kBStages;
  uint64_t* tmem_full = b_empty + kBStages;
  uint64_t* tmem_empty = tmem_full + 2;
  uint64_t* img_full = tmem_empty + 2;
  uint64_t* img_free = img_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(img_free + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int total_tiles = hs.m_tiles * tiles_n;
  const int cblocks = hs.C / 64;
  const uint32_t slab_tx = static_cast<uint32_t>((hs.TH + 2) * hs.PW * 128);

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" :: "l"(reinterpret_cast<uint64_t>(&map_a)) : "memory");
    asm volatile("prefetch.tensormap [%0];" :: "l"(reinterpret_cast<uint64_t>(&map_b)) : "memory");
    asm volatile("prefetch.tensormap [%0];" :: "l"(reinterpret_cast<uint64_t>(&map_c)) : "memory");
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int s = 0; s < kSlabStages; ++s) { mbar_init(&slab_full[s], 1); mbar_init(&slab_empty[s], 1); }
      for (int s = 0; s < kBStages; ++s) { mbar_init(&b_full[s], 1); mbar_init(&b_empty[s], 1); }
      for (int a = 0; a < 2; ++a) {
        mbar_init(&tmem_full[a], 1); mbar_init(&tmem_empty[a], 4);
        mbar_init(&img_full[a], 1); mbar_init(&img_free[a], kStatsThreads / 32);
      }
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                 :: "r"(smem_u32(tmem_slot)), "n"(2 * kAccCols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===== TMA producer: per channel block one slab, then the nine taps' weight tiles =====
    if (lane == 0) {
      uint32_t sit = 0, bit = 0;
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
        const int n0 = (t % tiles_n) * BLOCK_N, mt = t / tiles_n;
        const int img0 = mt / hs.h_tiles, h0 = (mt % hs.h_tiles) * hs.TH;
        for (int cb = 0; cb < cblocks; ++cb, ++sit) {
          const int ss = sit % kSlabStages;
          mbar_wait(&slab_empty[ss], ((sit / kSlabStages) & 1) ^ 1);
          mbar_expect_tx(&slab_full[ss], slab_tx);
          tma_load_4d(slab + ss * kSlabBytes, &map_a, &slab_full[ss], cb * 64, -1, h0 - 1, img0);
          for (int tap = 0; tap < 9; ++tap, ++bit) {
            const int s = bit % kBStages;
            mbar_wait(&b_empty[s], ((bit / kBStages) & 1) ^ 1);
            mbar_expect_tx(&b_full[s], kBBytes);
            const int krow = tap * b_rows_per_tap;
            if (B_MN) {
#pragma unroll
              for (int c = 0; c < BLOCK_N / 64; ++c)
                tma_load_2d(smem_b + s * kBBytes + c * kChunkBytes, &map_b, &b_full[s], n0 + c * 64, krow + cb * 64);
            } else {
              tma_load_2d(smem_b + s * kBBytes, &map_b, &b_full[s], cb * 64, krow + n0);
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: the warp stays converged, one elected lane issues (see gemm_tc.cuh) =====
    constexpr uint32_t idesc = make_idesc(1, BLOCK_N, false, B_MN);
    constexpr uint32_t hi = desc_hi();
    const uint32_t slab0 = smem_u32(slab), b0 = smem_u32(smem_b);
    const uint32_t slab_full0 = smem_u32(slab_full), slab_empty0 = smem_u32(slab_empty);
    const uint32_t b_full0 = smem_u32(b_full), b_empty0 = smem_u32(b_empty);
    const uint32_t tmem_full0 = smem_u32(tmem_full), tmem_empty0 = smem_u32(tmem_empty);
    uint32_t shift8[9];                                   // tap offset in 16-byte units of the descriptor address
#pragma unroll
    for (int tap = 0; tap < 9; ++tap) {
      const int r = tap / 3, sx = tap % 3;
      shift8[tap] = static_cast<uint32_t>(hs.flip ? (2 - r) * hs.PW + (2 - sx) : r * hs.PW + sx) * 8u;
    }
    uint32_t ss = 0, sph = 0, bs = 0, bph = 0, acc = 0, acc_ph = 0;
    for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
      mbar_wait_addr(tmem_empty0 + acc * 8, acc_ph ^ 1);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const uint32_t d_tmem = tmem_base + acc * kAccCols;
      for (int cb = 0; cb < cblocks; ++cb) {
        mbar_wait_addr(slab_full0 + ss * 8, sph);
        const uint32_t a_base = desc_lo(slab0 + ss * kSlabBytes, 16);
#pragma unroll
        for (int tap = 0; tap < 9; ++tap) {
          mbar_wait_addr(b_full0 + bs * 8, bph);
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          const uint32_t a_lo = a_base + shift8[tap];
          const uint32_t b_lo = desc_lo(b0 + bs * kBBytes, B_MN ? kChunkBytes : 16);
          if (elect_one()) {
#pragma unroll
            for (int h = 0; h < MT; ++h)                  // M tile h: 128 positions (16 KB) further on in the slab
#pragma unroll
              for (int k = 0; k < 4; ++k)
                umma2<0>(d_tmem + h * BLOCK_N, a_lo + h * 1024 + 2 * k, hi, b_lo + kStepB * k, hi, idesc,
                         cb > 0 || tap > 0 || k > 0);
            umma_commit_addr(b_empty0 + bs * 8);
          }
          __syncwarp();
          if (++bs == kBStages) { bs = 0; bph ^= 1; }
        }
        if (elect_one()) umma_commit_addr(slab_empty0 + ss * 8);
        __syncwarp();
        if (++ss == kSlabStages) { ss = 0; sph ^= 1; }
      }
      if (elect_one()) umma_commit_addr(tmem_full0 + acc * 8);
      __syncwarp();
      acc ^= 1;
      if (acc == 0) acc_ph ^= 1;
    }
  } else if (warp < 6) {
    // ===== epilogue warps =====
    uint32_t tcnt = 0;
    for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, ++tcnt) {
      const int n0 = (t % tiles_n) * BLOCK_N, mt = t / tiles_n;
      const int img0 = mt / hs.h_tiles, h0 = (mt % hs.h_tiles) * hs.TH;
      const uint32_t acc = tcnt & 1, acc_phase = (tcnt >> 1) & 1;
      const uint32_t ib = MT == 1 ? acc : 0;              // output image buffer
      uint8_t* img = out_stage + ib * kOutTileBytes;
      if (warp == 2 && lane == 0) {                       // the previous user of this image has been read by its store
        if (MT == 1) bulk_wait_read<1>(); else bulk_wait_read<0>();
      }
      if (epi.col_stats) {                                // ... and by the column-sum warps
        if (MT == 1) { if (tcnt >= 2) mbar_wait(&img_free[ib], ((tcnt >> 1) - 1) & 1); }
        else if (tcnt >= 1) mbar_wait(&img_free[ib], (tcnt - 1) & 1);
      }
      epi_barrier();
      mbar_wait(&tmem_full[acc], acc_phase);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
      for (int h = 0; h < MT; ++h) {
        const int m = h * kBlockM + (warp & 3) * 32 + lane;   // this thread's position in the tile
        const int oh = m / hs.PW, owp = m - oh * hs.PW;
        const bool keep = owp < hs.OW && oh < hs.TH && h0 + oh < hs.OH;
        drain_tile_staged<BLOCK_N>(tmem_base + acc * kAccCols + h * BLOCK_N, img, warp, lane, n0, epi, keep, h * kBlockM,
                                   kHalfBytes);
      }
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty[acc]);
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      epi_barrier();
      if (warp == 2 && lane == 0) {
#pragma unroll
        for (int h = 0; h < BLOCK_N / 64; ++h)
          if (n0 + h * 64 < N) tma_store_4d(&map_c, img + h * kHalfBytes, n0 + h * 64, 0, h0, img0, epi.tma_out == 2);
        bulk_commit();
        if (epi.col_stats) mbar_arrive(&img_full[ib]);
      }
    }
    if (warp == 2 && lane == 0) bulk_wait_all();
  } else if (epi.col_stats) {
    // ===== column-sum warps (rows that are not outputs were staged as zeros) =====
    uint32_t tcnt = 0;
    ColStats sums;
    const int et = threadIdx.x - kThreads;
    for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, ++tcnt) {
      const uint32_t ib = MT == 1 ? (tcnt & 1) : 0;
      mbar_wait(&img_full[ib], MT == 1 ? (tcnt >> 1) & 1 : tcnt & 1);
      tile_col_stats<BLOCK_N, kRows>(out_stage + ib * kOutTileBytes, et, kRows, sums);
      __syncwarp();
      if (lane == 0) mbar_arrive(&img_free[ib]);
    }
    store_col_stats<BLOCK_N>(sums, et, (blockIdx.x % tiles_n) * BLOCK_N, N, blockIdx.x / tiles_n, epi.col_stats);
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem_base), "n"(2 * kAccCols) : "memory");
  }
}

// ---- weight gradient ---------------------------------------------------------------------------------
// dW[(r,s,c), k] = sum over pixels.  The pixel tiles are the K dimension: both operands are MN-major
// boxes of P = TN*TH*OW pixel rows (P a multiple of 16, rows past the image zero-filled):
//   A chunk j (64 rows of dW = 64 channels of one tap): box of x at the tap's offset,
//   B chunk j (64 output channels):                      box of dz.
// Tiles = (128 rows of dW) x (BLOCK_N output channels) x split z over the pixel tiles; fp32 partials
// go to C + z*M*ldc and are reduced in order by splitk_reduce_kernel.
constexpr int kWgradStages = 3;

template <int BLOCK_N>
constexpr size_t smem_bytes_wgrad() {
  return 1024 + static_cast<size_t>(kWgradStages) * (2 * 128 * 128 + (BLOCK_N / 64) * 128 * 128) +
         (2 * kWgradStages + 4) * sizeof(uint64_t) + 16 + 4 * 32 * 33 * sizeof(float);
}

template <int BLOCK_N>
__global__ void __launch_bounds__(kThreads, 1)
conv_wgrad_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_dz,
                  float* __restrict__ Cout, const __grid_constant__ ConvShape cs, int M, int N, int ldc, int tiles_m,
                  int tiles_n, int splits, int k_tiles, int k_tiles_per_split) {
  constexpr int kStages = kWgradStages;
  constexpr int kAStage = 2 * 128 * 128, kBStage = (BLOCK_N / 64) * 128 * 128;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + kStages * kAStage;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + kStages * (kAStage + kBStage));
  uint64_t* empty = full + kStages;
  uint64_t* tmem_full = empty + kStages;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);
  float* stage_base = reinterpret_cast<float*>(tmem_slot + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int total_tiles = tiles_m * tiles_n * splits;
  const int P = cs.TN * cs.TH * cs.OW;                 // pixel rows per box: the K extent of one stage
  const int chunk_bytes = P * 128;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" :: "l"(reinterpret_cast<uint64_t>(&map_x)) : "memory");
    asm volatile("prefetch.tensormap [%0];" :: "l"(reinterpret_cast<uint64_t>(&map_dz)) : "memory");
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int s = 0; s < kStages; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
      for (int a = 0; a < 2; ++a) { mbar_init(&tmem_full[a], 1); mbar_init(&tmem_empty[a], 4); }
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
    if (lane == 0) {
      uint32_t it = 0;
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
        const int m0 = (t % tiles_m) * 128, n0 = ((t / tiles_m) % tiles_n) * BLOCK_N, z = t / (tiles_m * tiles_n);
        const int kt_begin = z * k_tiles_per_split, kt_end = min(k_tiles, kt_begin + k_tiles_per_split);
        const int a_chunks = (m0 + 64 < M) ? 2 : 1;          // the last row tile of dW may hold one chunk only
        for (int kt = kt_begin; kt < kt_end; ++kt, ++it) {
          const int s = it % kStages;
          const uint32_t phase = (it / kStages) & 1;
          int img0, h0;
          conv_tile_origin(cs, kt, &img0, &h0);
          mbar_wait(&empty[s], phase ^ 1);
          mbar_expect_tx(&full[s], (a_chunks + BLOCK_N / 64) * chunk_bytes);
          for (int c = 0; c < a_chunks; ++c) {
            const int row = m0 + c * 64;                       // flattened (tap, channel) row of dW
            const int tap = row / cs.C, c0 = row - tap * cs.C;
            const int r = tap / cs.S, sx = tap - r * cs.S;
            tma_load_4d(smem_a + s * kAStage + c * chunk_bytes, &map_x, &full[s], c0, sx - cs.pad,
                        h0 * cs.stride + r - cs.pad, img0);
          }
#pragma unroll
          for (int c = 0; c < BLOCK_N / 64; ++c)
            tma_load_4d(smem_b + s * kBStage + c * chunk_bytes, &map_dz, &full[s], n0 + c * 64, 0, h0, img0);
        }
      }
    }
  } else if (warp == 1) {
    // converged warp, one elected lane issues (gemm_tc.cuh, "lean issue path")
    constexpr uint32_t idesc = make_idesc(1, BLOCK_N, true, true);
    constexpr uint32_t hi = desc_hi();
    const uint32_t a0 = smem_u32(smem_a), b0 = smem_u32(smem_b);
    const uint32_t full0 = smem_u32(full), empty0 = smem_u32(empty);
    const uint32_t tmem_full0 = smem_u32(tmem_full), tmem_empty0 = smem_u32(tmem_empty);
    const int k_steps = P / 16;
    uint32_t st = 0, ph = 0, acc = 0, acc_ph = 0;
    for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
      const int z = t / (tiles_m * tiles_n);
      const int kt_begin = z * k_tiles_per_split, kt_end = min(k_tiles, kt_begin + k_tiles_per_split);
      mbar_wait_addr(tmem_empty0 + acc * 8, acc_ph ^ 1);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const uint32_t d_tmem = tmem_base + acc * BLOCK_N;
      for (int kt = kt_begin; kt < kt_end; ++kt) {
        mbar_wait_addr(full0 + st * 8, ph);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint32_t a_lo = desc_lo(a0 + st * kAStage, chunk_bytes), b_lo = desc_lo(b0 + st * kBStage, chunk_bytes);
        if (elect_one()) {
          for (int k = 0; k < k_steps; ++k)                   // 16 pixel rows = 2048 B per instruction
            umma2<0>(d_tmem, a_lo + 128u * k, hi, b_lo + 128u * k, hi, idesc, kt > kt_begin || k > 0);
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
  } else {
    uint32_t tcnt = 0;
    const Epilogue e = {nullptr, nullptr, 0, nullptr, 1};
    for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, ++tcnt) {
      const int m0 = (t % tiles_m) * 128, n0 = ((t / tiles_m) % tiles_n) * BLOCK_N, z = t / (tiles_m * tiles_n);
      const uint32_t acc = tcnt & 1, acc_phase = (tcnt >> 1) & 1;
      mbar_wait(&tmem_full[acc], acc_phase);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      drain_tile<BLOCK_N>(tmem_base + acc * BLOCK_N, true, stage_base, warp, lane, Cout, M, N, ldc, m0, n0, z, e);
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty[acc]);
    }
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
