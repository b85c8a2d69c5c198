// Host side of the implicit-GEMM convolutions (conv_tc.cuh): 4-D tensor maps, tile shapes, launches.
#include <cudaTypedefs.h>

#include <algorithm>

#include "conv_host.cuh"

namespace tfg {

static PFN_cuTensorMapEncodeTiled encode_fn() {
  static PFN_cuTensorMapEncodeTiled fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult st;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &st) == cudaSuccess &&
        st == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled>(p);
  }
  return fn;
}

// bf16 NHWC tensor [N, H, W, C]; box = {64 channels, bw columns, bh rows, bn images} visited with
// element strides {1, stride, stride, 1} (so the box spans bw*stride x bh*stride input positions).
int make_map_4d(CUtensorMap* map, const void* ptr, int N, int H, int W, int C, int bw, int bh, int bn, int stride) {
  PFN_cuTensorMapEncodeTiled enc = encode_fn();
  if (!enc) return fail(nullptr, TFG_ECUDA, "cuTensorMapEncodeTiled is not available from this driver");
  if ((reinterpret_cast<uintptr_t>(ptr) & 15) != 0 || (C * 2) % 16 != 0)
    return fail(nullptr, TFG_EINVAL, "activation tensor must be 16-byte aligned with C a multiple of 8");
  cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
  cuuint64_t strides[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
  cuuint32_t box[4] = {64, (cuuint32_t)(bw * stride), (cuuint32_t)(bh * stride), (cuuint32_t)bn};
  cuuint32_t estr[4] = {1, (cuuint32_t)stride, (cuuint32_t)stride, 1};
  if (box[1] > 256 || box[2] > 256 || box[3] > 256) return fail(nullptr, TFG_ELIMIT, "TMA box dimension above 256");
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(ptr), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return fail(nullptr, TFG_ECUDA, "cuTensorMapEncodeTiled(4d) failed with CUresult " + std::to_string((int)r));
  return TFG_OK;
}

int make_map_4d_box(CUtensorMap* map, const void* ptr, int N, int H, int W, int C, int bw, int bh) {
  PFN_cuTensorMapEncodeTiled enc = encode_fn();
  if (!enc) return fail(nullptr, TFG_ECUDA, "cuTensorMapEncodeTiled is not available from this driver");
  if ((reinterpret_cast<uintptr_t>(ptr) & 15) != 0 || (C * 2) % 16 != 0)
    return fail(nullptr, TFG_EINVAL, "activation tensor must be 16-byte aligned with C a multiple of 8");
  cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
  cuuint64_t strides[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
  cuuint32_t box[4] = {64, (cuuint32_t)bw, (cuuint32_t)bh, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  if (bw > 256 || bh > 256) return fail(nullptr, TFG_ELIMIT, "TMA box dimension above 256");
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(ptr), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return fail(nullptr, TFG_ECUDA, "cuTensorMapEncodeTiled(4d box) failed with CUresult " + std::to_string((int)r));
  return TFG_OK;
}

// conv3x3_halo_kernel: 3x3, stride 1, 64-channel blocks, at least two image rows per 128-position tile
// and a slab that fits its stage
bool halo_fits(int OH, int OW, int C, int R, int stride) {
  const int PW = OW + 2;
  return R == 3 && stride == 1 && C % 64 == 0 && OW >= 12 && PW <= tc::halo_max_pw<1>() && OH >= 2;
}

// M tiles per CTA: two (256 positions, the accumulators share every weight tile) when the wider slab
// fits its stage and the coarser tiles do not cost more in wave quantisation than the sharing returns
// (ResNet stage 3: 512 double tiles are 4 waves of 2 units against 7 waves of 1: stay with one)
int halo_mt(int Nimg, int OH, int OW, int tiles_n) {
  const int PW = OW + 2;
  if (PW > tc::halo_max_pw<2>()) return 1;
  auto waves = [&](int mt) {
    const int th = std::max(1, std::min(OH, 128 * mt / PW));
    const long long tiles = (long long)Nimg * ((OH + th - 1) / th) * tiles_n;
    return (tiles + kNumSMs - 1) / kNumSMs;
  };
  return 2.0 * (double)waves(2) <= 1.1 * (double)waves(1) ? 2 : 1;
}

void fill_halo_shape(tc::HaloShape* hs, int Nimg, int OH, int OW, int C, int flip, int tiles_n) {
  hs->Nimg = Nimg; hs->OH = OH; hs->OW = OW; hs->PW = OW + 2;
  hs->MT = halo_mt(Nimg, OH, OW, tiles_n);
  hs->TH = std::max(1, std::min(OH, 128 * hs->MT / hs->PW));
  hs->h_tiles = (OH + hs->TH - 1) / hs->TH;
  hs->m_tiles = Nimg * hs->h_tiles;
  hs->C = C; hs->flip = flip;
}

int halo_grid_rows(const tc::HaloShape& hs, int tiles_n) {
  return tc::persistent_grid((long long)hs.m_tiles * tiles_n, tiles_n, true, kNumSMs) / tiles_n;
}

template <int BN, bool B_MN, int MT>
static cudaError_t launch_halo(const HaloConv& g, cudaStream_t s) {
  static bool attr_set = false;
  constexpr size_t smem = tc::smem_bytes_halo<BN, MT>();
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(tc::conv3x3_halo_kernel<BN, B_MN, MT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)smem);
    if (e != cudaSuccess) return e;
    attr_set = true;
  }
  // Descriptor base-offset field for the shifted slab views: measured on B200, the MMA unit applies
  // the 128-byte swizzle to absolute shared-memory address bits, so a start address that is a multiple
  // of 128 B inside a 1024-B-aligned slab needs NO base offset (setting it gives wrong results).
  const int base_mode = 0;
  const int tiles_n = (g.N + BN - 1) / BN;
  const long long total = (long long)g.hs.m_tiles * tiles_n;
  tc::Epilogue epi = g.epi;
  epi.tma_out = g.tma_out;
  const int grid = tc::persistent_grid(total, tiles_n, epi.col_stats != nullptr, kNumSMs);
  tc::conv3x3_halo_kernel<BN, B_MN, MT><<<grid, tc::kThreadsPersistent, smem, s>>>(g.a, g.b, g.c, g.hs, g.N, tiles_n,
                                                                                  g.b_rows_per_tap, base_mode, epi);
  return cudaGetLastError();
}

template <int MT>
static cudaError_t launch_halo_mt(const HaloConv& g, cudaStream_t s) {
  if (g.b_mn) return g.block_n == 64 ? launch_halo<64, true, MT>(g, s) : launch_halo<128, true, MT>(g, s);
  return g.block_n == 64 ? launch_halo<64, false, MT>(g, s) : launch_halo<128, false, MT>(g, s);
}

int launch_conv_halo(tfg_job* j, const HaloConv& g, cudaStream_t s) {
  const cudaError_t e = g.hs.MT == 2 ? launch_halo_mt<2>(g, s) : launch_halo_mt<1>(g, s);
  if (j) prof_mark(j, g.name);
  if (e != cudaSuccess) return fail_cuda(j, e, g.name);
  return TFG_OK;
}

// Tile of the pixel space: TN images x TH rows x OW columns, at most 128 pixels.  rows_mult16: the
// pixel count must be a multiple of 16 (wgrad: pixels are the K dimension); rows past the image are
// then allowed (the TMA unit zero-fills them).
void pick_conv_tile(int Nimg, int OH, int OW, bool rows_mult16, int* TN, int* TH) {
  if (!rows_mult16) {
    if (OH * OW <= 64) { *TH = OH; *TN = std::max(1, std::min(Nimg, 128 / (OH * OW))); }
    else { *TN = 1; *TH = std::max(1, std::min(OH, 128 / OW)); }
    return;
  }
  int best = 0, btn = 1, bth = 1;
  for (int tn = 1; tn <= 64; ++tn)
    for (int th = 1; th <= 128; ++th) {
      if (tn > 1 && th < OH) continue;                 // several images only when a box holds whole images
      if (tn == 1 && th > OH + 15) continue;
      const int p = tn * th * OW;
      if (p > 128 || p % 16) continue;
      // prefer more valid pixels per box, then fewer wasted rows
      const int valid = tn * std::min(th, OH) * OW;
      if (valid > best) { best = valid; btn = tn; bth = th; }
    }
  *TN = btn;
  *TH = bth;
}

void fill_conv_shape(tc::ConvShape* cs, int Nimg, int OH, int OW, int TN, int TH, int R, int S, int stride, int pad,
                     int C, int flip) {
  cs->Nimg = Nimg; cs->OH = OH; cs->OW = OW; cs->TN = TN; cs->TH = TH;
  cs->h_tiles = (OH + TH - 1) / TH;
  cs->m_tiles = ((Nimg + TN - 1) / TN) * cs->h_tiles;
  cs->R = R; cs->S = S; cs->stride = stride; cs->pad = pad; cs->C = C; cs->flip = flip;
  cs->ntaps = 0; cs->out_stride = 1; cs->out_h0 = 0; cs->out_w0 = 0;
  for (int t = 0; t < 9; ++t) cs->tap_dh[t] = cs->tap_dw[t] = cs->tap_w[t] = 0;
}

// Parity class (ph, pw) of the data gradient of an R x R (R = 1 or 3), stride-2, pad R/2 convolution:
// dx[2i + ph, 2j + pw] = sum over the taps (r, s) with (ph + pad - r), (pw + pad - s) even of
// dz[i + (ph + pad - r) / 2, j + (pw + pad - s) / 2] * W[r, s].  Returns the number of taps (0: the
// class receives no gradient from this convolution).
int fill_dgrad_s2_class(tc::ConvShape* cs, int R, int ph, int pw) {
  const int pad = R / 2;
  int n = 0;
  for (int r = 0; r < R; ++r)
    for (int s = 0; s < R; ++s) {
      const int a = ph + pad - r, b = pw + pad - s;
      if ((a & 1) || (b & 1)) continue;
      cs->tap_dh[n] = (signed char)(a / 2);     // a in {-1..2}: exact for even a
      cs->tap_dw[n] = (signed char)(b / 2);
      cs->tap_w[n] = (signed char)(r * R + s);
      ++n;
    }
  cs->ntaps = n;
  cs->out_stride = 2; cs->out_h0 = ph; cs->out_w0 = pw;
  return n;
}

__global__ void __launch_bounds__(256)
conv_splitk_reduce_kernel(const float* __restrict__ partial, float* __restrict__ out, int splits, long long mn) {
  const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= mn) return;
  float s = 0.f;
  for (int z = 0; z < splits; ++z) s += partial[z * mn + i];   // fixed order: deterministic
  out[i] = s;
}

template <int BN, bool B_MN>
static cudaError_t launch_cg(const ConvGemm& g, cudaStream_t s) {
  static bool attr_set = false;
  constexpr size_t smem = tc::smem_bytes_conv<BN>();
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(tc::conv_gemm_kernel<BN, B_MN>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)smem);
    if (e != cudaSuccess) return e;
    attr_set = true;
  }
  const int tiles_n = (g.N + BN - 1) / BN;
  const long long total = (long long)g.cs.m_tiles * tiles_n;
  tc::Epilogue epi = g.epi;
  epi.direct = 1;
  epi.tma_out = (g.tma_out && epi.out_bf16 && !epi.mask) ? g.tma_out : 0;
  if (!epi.tma_out) epi.col_stats = nullptr;
  const int grid = tc::persistent_grid(total, tiles_n, epi.col_stats != nullptr, kNumSMs);
  tc::conv_gemm_kernel<BN, B_MN><<<grid, tc::kThreadsPersistent, smem, s>>>(g.a, g.b, epi.tma_out ? g.c : g.a, g.C, g.cs, g.N,
                                                                  g.ldc, tiles_n, g.b_rows_per_tap, epi);
  return cudaGetLastError();
}

int launch_conv_gemm(tfg_job* j, const ConvGemm& g, cudaStream_t s) {
  cudaError_t e;
  if (g.b_mn) e = g.block_n == 64 ? launch_cg<64, true>(g, s) : launch_cg<128, true>(g, s);
  else e = g.block_n == 64 ? launch_cg<64, false>(g, s) : launch_cg<128, false>(g, s);
  if (j) prof_mark(j, g.name);
  if (e != cudaSuccess) return fail_cuda(j, e, g.name);
  return TFG_OK;
}

template <int BN>
static cudaError_t launch_cw(const ConvWgrad& g, float* dst, int splits, int k_tiles, int ktps, cudaStream_t s) {
  static bool attr_set = false;
  constexpr size_t smem = tc::smem_bytes_wgrad<BN>();
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(tc::conv_wgrad_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    attr_set = true;
  }
  const int tiles_m = (g.M + 127) / 128, tiles_n = (g.N + BN - 1) / BN;
  const long long total = (long long)tiles_m * tiles_n * splits;
  const int grid = (int)std::min<long long>(total, kNumSMs);
  tc::conv_wgrad_kernel<BN><<<grid, tc::kThreads, smem, s>>>(g.x, g.dz, dst, g.cs, g.M, g.N, g.N, tiles_m, tiles_n, splits,
                                                             k_tiles, ktps);
  return cudaGetLastError();
}

int launch_conv_wgrad(tfg_job* j, const ConvWgrad& g, cudaStream_t s) {
  const int k_tiles = g.cs.m_tiles;
  int splits = std::max(1, std::min(g.splits, k_tiles));
  const int ktps = (k_tiles + splits - 1) / splits;
  splits = (k_tiles + ktps - 1) / ktps;
  float* dst = splits > 1 ? g.partial : g.C;
  cudaError_t e = g.block_n == 64 ? launch_cw<64>(g, dst, splits, k_tiles, ktps, s)
                                  : launch_cw<128>(g, dst, splits, k_tiles, ktps, s);
  if (j) prof_mark(j, g.name);
  if (e != cudaSuccess) return fail_cuda(j, e, g.name);
  if (splits > 1) {
    const long long mn = (long long)g.M * g.N;
    conv_splitk_reduce_kernel<<<(int)((mn + 255) / 256), 256, 0, s>>>(g.partial, g.C, splits, mn);
    if (j) prof_mark(j, "split-K reduce");
    if ((e = cudaGetLastError()) != cudaSuccess) return fail_cuda(j, e, "conv_splitk_reduce_kernel");
  }
  return TFG_OK;
}

}  // namespace tfg
