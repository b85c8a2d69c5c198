// Host side of the tensor path: TMA tensor maps, launches of the tcgen05 GEMM
// (gemm_tc.cuh), the operand-preparation kernels (casts, transposes, im2col in operand
// precision) and the model's contractions in math modes bf16 / tf32.
#include <cuda_bf16.h>
#include <cudaTypedefs.h>

#include <algorithm>

#include "common.cuh"
#include "gemm_tc.cuh"
#include "conv_x3.cuh"
#include "gemm_x3.cuh"
#include "model_mnist.cuh"
#include "tc_host.cuh"

namespace tfg {

// ---- tensor maps ---------------------------------------------------------------------------
static PFN_cuTensorMapEncodeTiled get_encode() {
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

// Operand with `rows` (M or N) and `cols` (K).  K-major: stored [rows, K], pitch ld, box = 128 B
// of K x box_rows.  MN-major: stored [K, rows], pitch ld, box = 128 B of rows x one stage of K.
int make_map(CUtensorMap* map, int kind, const void* ptr, int64_t rows, int64_t cols, int64_t ld,
             int box_rows, bool mn_major) {
  PFN_cuTensorMapEncodeTiled enc = get_encode();
  if (!enc) return fail(nullptr, TFG_ECUDA, "cuTensorMapEncodeTiled is not available from this driver");
  const int esize = kind == TFG_MATH_BF16 ? 2 : 4;
  if ((reinterpret_cast<uintptr_t>(ptr) & 15) != 0 || ((ld * esize) & 15) != 0)
    return fail(nullptr, TFG_EINVAL, "tensor-path operand must be 16-byte aligned with a 16-byte row pitch");
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)(ld * esize)};
  cuuint32_t box[2] = {(cuuint32_t)(tc::kBlockKBytes / esize), (cuuint32_t)box_rows};
  if (mn_major) {
    dims[0] = (cuuint64_t)rows; dims[1] = (cuuint64_t)cols;
    box[0] = (cuuint32_t)(128 / esize); box[1] = (cuuint32_t)(tc::kBlockKBytes / esize);
  }
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(map, kind == TFG_MATH_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32,
                   2, const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   (mn_major && kind == TFG_MATH_TF32) ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B
                                                       : CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return fail(nullptr, TFG_ECUDA, "cuTensorMapEncodeTiled failed with CUresult " + std::to_string((int)r));
  return TFG_OK;
}

// fp32 NHWC tensor [N, H, W, C]: box {32 channels (one 128-byte line), bw columns, bh rows, 1 image}.
// mn_major: the box feeds an MN-major tf32 operand (pixels = K): 128-byte swizzle with 32-byte atoms.
int make_map_4d_f32(CUtensorMap* map, const void* ptr, int N, int H, int W, int C, int bw, int bh, bool mn_major) {
  PFN_cuTensorMapEncodeTiled enc = get_encode();
  if (!enc) return fail(nullptr, TFG_ECUDA, "cuTensorMapEncodeTiled is not available from this driver");
  if ((reinterpret_cast<uintptr_t>(ptr) & 15) != 0 || C % 32 != 0)
    return fail(nullptr, TFG_EINVAL, "activation tensor must be 16-byte aligned with C a multiple of 32");
  cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
  cuuint64_t strides[3] = {(cuuint64_t)C * 4, (cuuint64_t)W * C * 4, (cuuint64_t)H * W * C * 4};
  cuuint32_t box[4] = {32, (cuuint32_t)bw, (cuuint32_t)bh, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<void*>(ptr), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, mn_major ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return fail(nullptr, TFG_ECUDA, "cuTensorMapEncodeTiled(4d fp32) failed with CUresult " + std::to_string((int)r));
  return TFG_OK;
}

__global__ void __launch_bounds__(256)
splitk_reduce_kernel(const float* __restrict__ partial, float* __restrict__ out, int splits, long long mn,
                     const float* __restrict__ bias, int N, int relu) {
  const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= mn) return;
  float s = 0.f;
  for (int z = 0; z < splits; ++z) s += partial[z * mn + i];   // fixed order: deterministic
  if (bias) s += bias[i % N];
  if (relu) s = fmaxf(s, 0.f);
  out[i] = s;
}

template <int BN, int KIND, bool A_MN, bool B_MN, int STAGES>
static cudaError_t launch_stage(const TcGemm& g, float* dst, int ldc, int k_blocks, int kps, int splits,
                                tc::Epilogue epi, cudaStream_t s) {
  static bool attr_set = false;
  constexpr size_t smem = tc::smem_bytes<BN, STAGES>();
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(tc::gemm_tc_kernel<BN, KIND, A_MN, B_MN, STAGES>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    attr_set = true;
  }
  dim3 grid((g.N + BN - 1) / BN, (g.M + tc::kBlockM - 1) / tc::kBlockM, splits);
  tc::gemm_tc_kernel<BN, KIND, A_MN, B_MN, STAGES><<<grid, tc::kThreads, smem, s>>>(g.a, g.b, dst, g.M, g.N, ldc,
                                                                                      k_blocks, kps, epi);
  return cudaGetLastError();
}

// 3xTF32 split product (gemm_x3.cuh): stages of 48 KB (BLOCK_N = 64) / 64 KB (128), one tile per CTA
template <int BN, bool A_MN, bool B_MN, int STAGES>
static cudaError_t launch_x3_stage(const TcGemm& g, float* dst, int ldc, int k_blocks, int kps, int splits,
                                   tc::Epilogue epi, cudaStream_t s) {
  static bool attr_set = false;
  constexpr size_t smem = tc::smem_bytes_x3<BN, STAGES>();
  static_assert(smem <= 232448, "shared memory budget");
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(tc::gemm_x3_kernel<BN, A_MN, B_MN, STAGES>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    attr_set = true;
  }
  dim3 grid((g.N + BN - 1) / BN, (g.M + tc::kBlockM - 1) / tc::kBlockM, splits);
  tc::gemm_x3_kernel<BN, A_MN, B_MN, STAGES><<<grid, tc::kThreads, smem, s>>>(
      g.a, g.b, g.a_lo, g.b_lo, dst, g.M, g.N, ldc, k_blocks, kps, g.chunk_kb > 0 ? g.chunk_kb : 2, epi);
  return cudaGetLastError();
}

template <int BN, bool A_MN, bool B_MN>
static cudaError_t launch_x3(const TcGemm& g, float* dst, int ldc, int k_blocks, int kps, int splits,
                             tc::Epilogue epi, cudaStream_t s) {
  const long long ctas = (long long)((g.N + BN - 1) / BN) * ((g.M + tc::kBlockM - 1) / tc::kBlockM) * splits;
  if (kps <= 1) return launch_x3_stage<BN, A_MN, B_MN, 1>(g, dst, ldc, k_blocks, kps, splits, epi, s);
  if (kps <= 3 || ctas >= 2ll * kNumSMs) return launch_x3_stage<BN, A_MN, B_MN, 2>(g, dst, ldc, k_blocks, kps, splits, epi, s);
  return launch_x3_stage<BN, A_MN, B_MN, 3>(g, dst, ldc, k_blocks, kps, splits, epi, s);
}

template <int BN>
static cudaError_t launch_x3_major(const TcGemm& g, float* dst, int ldc, int k_blocks, int kps, int splits,
                                   tc::Epilogue epi, cudaStream_t s) {
  if (!g.a_mn && !g.b_mn) return launch_x3<BN, false, false>(g, dst, ldc, k_blocks, kps, splits, epi, s);
  if (!g.a_mn && g.b_mn) return launch_x3<BN, false, true>(g, dst, ldc, k_blocks, kps, splits, epi, s);
  if (g.a_mn && g.b_mn) return launch_x3<BN, true, true>(g, dst, ldc, k_blocks, kps, splits, epi, s);
  return launch_x3<BN, true, false>(g, dst, ldc, k_blocks, kps, splits, epi, s);
}

template <int BN, int KIND, bool A_MN, bool B_MN>
static cudaError_t launch_persistent(const TcGemm& g, float* dst, int ldc, int k_blocks, int kps, int splits,
                                     tc::Epilogue epi, cudaStream_t s) {
  static bool attr_set = false;
  constexpr size_t smem = tc::smem_bytes_persistent<BN>();
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(tc::gemm_tc_persistent_kernel<BN, KIND, A_MN, B_MN>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    attr_set = true;
  }
  const int tiles_m = (g.M + tc::kBlockM - 1) / tc::kBlockM, tiles_n = (g.N + BN - 1) / BN;
  const long long total = (long long)tiles_m * tiles_n * splits;
  const int grid = tc::persistent_grid(total, tiles_n, epi.tma_out && epi.col_stats, kNumSMs);
  tc::gemm_tc_persistent_kernel<BN, KIND, A_MN, B_MN><<<grid, tc::kThreadsPersistent, smem, s>>>(
      g.a, g.b, epi.tma_out ? g.c : g.a, dst, g.M, g.N, ldc, k_blocks, kps, tiles_m, tiles_n, splits, epi);
  return cudaGetLastError();
}

template <int BN, int KIND, bool A_MN, bool B_MN>
static cudaError_t launch_inst(const TcGemm& g, float* dst, int ldc, int k_blocks, int kps, int splits,
                               tc::Epilogue epi, cudaStream_t s) {
  // Ring depth.  A shallow ring leaves shared memory for more CTAs per SM, whose prologues and
  // epilogues then overlap each other's loads: that wins whenever the grid is several waves deep
  // (measured on the ResNet-50 layers: +14 % end to end); a grid of about one wave needs the deep
  // ring to overlap its own loads with its MMAs (the MNIST GEMMs).
  const long long ctas = (long long)((g.N + BN - 1) / BN) * ((g.M + tc::kBlockM - 1) / tc::kBlockM) * splits;
  if (ctas >= 2ll * kNumSMs || epi.tma_out)   // at least two waves of tiles: one persistent CTA per SM
    return launch_persistent<BN, KIND, A_MN, B_MN>(g, dst, ldc, k_blocks, kps, splits, epi, s);
  int stages = kps <= 1 ? 1 : kps <= 3 ? 2 : 4;
  if (ctas >= 4ll * kNumSMs) stages = 1;
  else if (ctas >= 2ll * kNumSMs) stages = std::min(stages, 2);
  // one wave of CTAs with a long K loop each (split-K weight gradients: K = all pixels): bandwidth is
  // bytes in flight over the 2 us TMA round trip, so take the deepest ring that fits
  if (stages == 4 && kps >= 48 && ctas <= kNumSMs) stages = 6;
  if (stages == 1) return launch_stage<BN, KIND, A_MN, B_MN, 1>(g, dst, ldc, k_blocks, kps, splits, epi, s);
  if (stages == 2) return launch_stage<BN, KIND, A_MN, B_MN, 2>(g, dst, ldc, k_blocks, kps, splits, epi, s);
  if (stages == 6) return launch_stage<BN, KIND, A_MN, B_MN, 6>(g, dst, ldc, k_blocks, kps, splits, epi, s);
  return launch_stage<BN, KIND, A_MN, B_MN, 4>(g, dst, ldc, k_blocks, kps, splits, epi, s);
}

template <int BN, int KIND>
static cudaError_t launch_major(const TcGemm& g, float* dst, int ldc, int k_blocks, int kps, int splits,
                                tc::Epilogue epi, cudaStream_t s) {
  if (!g.a_mn && !g.b_mn) return launch_inst<BN, KIND, false, false>(g, dst, ldc, k_blocks, kps, splits, epi, s);
  if (!g.a_mn && g.b_mn) return launch_inst<BN, KIND, false, true>(g, dst, ldc, k_blocks, kps, splits, epi, s);
  if (g.a_mn && g.b_mn) return launch_inst<BN, KIND, true, true>(g, dst, ldc, k_blocks, kps, splits, epi, s);
  return launch_inst<BN, KIND, true, false>(g, dst, ldc, k_blocks, kps, splits, epi, s);
}

// j may be NULL (stateless test entry point)
int launch_tc(tfg_job* j, const TcGemm& g, int kind, cudaStream_t s) {
  if (g.x3 && (kind != TFG_MATH_TF32 || g.tma_out)) return fail(j, TFG_EINVAL, "the split product takes fp32 operands and a plain store");
  const int block_k = tc::kBlockKBytes / (kind == TFG_MATH_BF16 ? 2 : 4);
  const int k_blocks = (g.K + block_k - 1) / block_k;
  int splits = std::max(1, std::min(g.splits, k_blocks));
  const int kps = (k_blocks + splits - 1) / splits;
  splits = (k_blocks + kps - 1) / kps;                 // no empty split
  float* dst = splits > 1 ? g.partial : g.C;
  const int ldc = splits > 1 ? g.N : g.ldc;
  tc::Epilogue epi = splits > 1 ? tc::Epilogue{nullptr, nullptr, 0, nullptr, 0} : g.epi;
  epi.direct = 1;   // register-direct stores (measured on the ResNet-50 layers: 6100 vs 5230 images/s against the
                    // smem-transposed stores, which remain the path for ragged tails)
  epi.tma_out = 0;
  if (g.tma_out) {
    if (splits > 1 || !g.epi.out_bf16 || g.epi.mask || kind != TFG_MATH_BF16)
      return fail(j, TFG_EINVAL, "the TMA output path takes a bf16 output, no mask and no split-K");
    epi.tma_out = g.tma_out;
  }
  if (splits > 1 && g.epi.out_bf16) return fail(j, TFG_EINVAL, "split-K writes fp32 only");
  if (splits > 1 && (!g.partial || g.ldc != g.N))
    return fail(j, TFG_EINVAL, "split-K needs a partial buffer and a dense output");
  cudaError_t e;
  if (g.x3)
    e = g.block_n == 64 ? launch_x3_major<64>(g, dst, ldc, k_blocks, kps, splits, epi, s)
                        : launch_x3_major<128>(g, dst, ldc, k_blocks, kps, splits, epi, s);
  else if (kind == TFG_MATH_BF16)
    e = g.block_n == 64 ? launch_major<64, 0>(g, dst, ldc, k_blocks, kps, splits, epi, s)
                        : launch_major<128, 0>(g, dst, ldc, k_blocks, kps, splits, epi, s);
  else
    e = g.block_n == 64 ? launch_major<64, 1>(g, dst, ldc, k_blocks, kps, splits, epi, s)
                        : launch_major<128, 1>(g, dst, ldc, k_blocks, kps, splits, epi, s);
  if (j) prof_mark(j, g.name);
  if (e != cudaSuccess) return fail_cuda(j, e, g.name);
  if (splits > 1 && !g.reduce_later) {
    const long long mn = (long long)g.M * g.N;
    if (g.epi.mask) return fail(j, TFG_EINVAL, "split-K cannot apply a mask epilogue");
    splitk_reduce_kernel<<<(int)((mn + 255) / 256), 256, 0, s>>>(g.partial, g.C, splits, mn, g.epi.bias, g.N,
                                                                 g.epi.relu);
    if (j) prof_mark(j, "split-K reduce");
    if ((e = cudaGetLastError()) != cudaSuccess) return fail_cuda(j, e, "splitk_reduce_kernel");
  }
  return TFG_OK;
}

int tc_splits(const TcGemm& g, int kind) {
  const int block_k = tc::kBlockKBytes / (kind == TFG_MATH_BF16 ? 2 : 4);
  const int k_blocks = (g.K + block_k - 1) / block_k;
  const int splits = std::max(1, std::min(g.splits, k_blocks));
  const int kps = (k_blocks + splits - 1) / splits;
  return (k_blocks + kps - 1) / kps;
}

// ---- operand preparation ---------------------------------------------------------------------
template <typename T> __device__ __forceinline__ float to_f32(T v);
template <> __device__ __forceinline__ float to_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f32<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }

// row-major im2col of a1[B,14,14,32] in operand precision: col[m, (kh*5+kw)*32 + c]; a thread moves
// 16 bytes of col (4 fp32 / 8 bf16 channels)
template <typename T>
__global__ void __launch_bounds__(256)
im2col_lp_kernel(const float* __restrict__ a1, T* __restrict__ col, int B) {
  constexpr int kPer = 16 / sizeof(T);          // channels per thread
  constexpr int kGroups = 32 / kPer;            // threads per tap
  const long long total = (long long)B * 196 * 25 * kGroups;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int cg = (int)(i % kGroups);
    const int kk = (int)((i / kGroups) % 25);
    const long long m = (i / kGroups) / 25;
    const int ow = (int)(m % 14), oh = (int)((m / 14) % 14);
    const long long b = m / 196;
    const int ih = oh + kk / 5 - 2, iw = ow + kk % 5 - 2;
    const bool ok = ih >= 0 && ih < 14 && iw >= 0 && iw < 14;
    const float* src = a1 + ((b * 14 + ih) * 14 + iw) * 32 + cg * kPer;
    T* dst = col + m * 800 + kk * 32 + cg * kPer;
#pragma unroll
    for (int q = 0; q < kPer / 4; ++q) {
      const float4 v = ok ? *reinterpret_cast<const float4*>(src + 4 * q) : make_float4(0.f, 0.f, 0.f, 0.f);
      store4(dst + 4 * q, v);
    }
  }
}

// dense un-pooled gradient in operand precision: gy[m, c] = argmax position ? masked grad : 0
template <typename T>
__global__ void __launch_bounds__(256)
unpool_lp_kernel(const float* __restrict__ gp, const float* __restrict__ a, const uint8_t* __restrict__ idx,
                 T* __restrict__ gy, float* __restrict__ gy_lo, int B, int HO, int C) {
  const long long total = (long long)B * HO * HO * C;
  const int HI = HO * 2;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    const int pw = (int)((i / C) % HO), ph = (int)((i / C / HO) % HO);
    const long long b = i / C / HO / HO;
    const float g = a[i] > 0.f ? gp[i] : 0.f;
    const int arg = idx[i];
#pragma unroll
    for (int dy = 0; dy < 2; ++dy)
#pragma unroll
      for (int dx = 0; dx < 2; ++dx) {
        const long long m = (b * HI + 2 * ph + dy) * HI + 2 * pw + dx;
        gy[m * C + c] = cvt<T>((dy * 2 + dx == arg) ? g : 0.f);
        if (gy_lo) gy_lo[m * C + c] = (dy * 2 + dx == arg) ? tc::tf32_lo(g) : 0.f;
      }
  }
}

// ---- 3xTF32 path: the gradient of the 2x2 max-pool + relu, four channels per thread -----------------------
// (the forward pooling lives in the epilogue of the conv2 forward kernel, conv_x3.cuh)
// gp[B,7,7,64] (gradient at the pooled, post-relu map) -> the dense gradient
// (gp = the pooled gradient da2[B,7,7,64], here still in split-K partial sums)
// gy[B,14,14,64] at the convolution output (+ low-order plane): the masked value at the argmax position, 0 at
// the other three.  The masked pooled gradient is also what the bias gradient sums: every thread keeps the sum
// of its four channels over its positions, a block adds its 16 position lanes in lane order, and
// bias_partial[block][64] goes to reduce_rows_kernel -- the separate column-sum pass over da2 is gone.
constexpr int kUnpoolBlocks = 2 * kNumSMs;
constexpr int kMaxUnpoolSplits = 4;   // K splits of the fc1 data gradient that unpool_x3_kernel adds itself (it uses 3)
__global__ void __launch_bounds__(256)
unpool_x3_kernel(const float* __restrict__ gp_partial, int splits, const float* __restrict__ a, const uint8_t* __restrict__ idx,
                 float* __restrict__ gy, float* __restrict__ gy_lo, float* __restrict__ bias_partial, int B) {
  __shared__ float4 red[16][16];
  const long long total = (long long)B * 49 * 16;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int q = (int)(i & 15);                      // == threadIdx.x & 15: the stride is a multiple of 16
    const long long pos = i >> 4;
    const int pw = (int)(pos % 7), ph = (int)((pos / 7) % 7);
    const long long b = pos / 49;
    // the pooled gradient arrives as the K-split partial sums of the fc1 data-gradient GEMM: added here, in split order
    float4 pv[kMaxUnpoolSplits];
#pragma unroll
    for (int z = 0; z < kMaxUnpoolSplits; ++z)       // all loads first, then the sum in split order
      if (z < splits) pv[z] = *reinterpret_cast<const float4*>(gp_partial + (z * total + i) * 4);
    float4 g4 = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int z = 0; z < kMaxUnpoolSplits; ++z)
      if (z < splits) { g4.x += pv[z].x; g4.y += pv[z].y; g4.z += pv[z].z; g4.w += pv[z].w; }
    const float4 a4 = *reinterpret_cast<const float4*>(a + i * 4);
    const uchar4 k4 = *reinterpret_cast<const uchar4*>(idx + i * 4);
    const float g[4] = {a4.x > 0.f ? g4.x : 0.f, a4.y > 0.f ? g4.y : 0.f, a4.z > 0.f ? g4.z : 0.f, a4.w > 0.f ? g4.w : 0.f};
    const int k[4] = {k4.x, k4.y, k4.z, k4.w};
    acc.x += g[0]; acc.y += g[1]; acc.z += g[2]; acc.w += g[3];
#pragma unroll
    for (int dy = 0; dy < 2; ++dy)
#pragma unroll
      for (int dx = 0; dx < 2; ++dx) {
        const int w = dy * 2 + dx;
        const long long m = ((b * 14 + 2 * ph + dy) * 14 + 2 * pw + dx) * 64 + q * 4;
        float v[4], l[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) { v[u] = k[u] == w ? g[u] : 0.f; l[u] = k[u] == w ? tc::tf32_lo(g[u]) : 0.f; }
        *reinterpret_cast<float4*>(gy + m) = make_float4(v[0], v[1], v[2], v[3]);
        *reinterpret_cast<float4*>(gy_lo + m) = make_float4(l[0], l[1], l[2], l[3]);
      }
  }
  red[threadIdx.x >> 4][threadIdx.x & 15] = acc;
  __syncthreads();
  if (threadIdx.x < 16) {
    float4 s = red[0][threadIdx.x];
#pragma unroll
    for (int r = 1; r < 16; ++r) {
      const float4 v = red[r][threadIdx.x];
      s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
    }
    *reinterpret_cast<float4*>(bias_partial + blockIdx.x * 64 + threadIdx.x * 4) = s;
  }
}

// ---- model state of the tensor path --------------------------------------------------------
// Operands are the row-major buffers themselves, cast to operand precision; which index is
// contiguous decides K-major vs MN-major per GEMM, so nothing is ever transposed:
//   conv2 fwd   y2[M2,64]     = col[M2,800] (K)   x W2[800,64]   (MN)
//   fc1   fwd   h1[B,1024]    = a2[B,3136]  (K)   x W3[3136,1024] (MN)
//   fc1   wgrad dW3[3136,1024]= a2[B,3136]  (MN)  x dh1[B,1024]  (MN)      K = B
//   fc1   dgrad da2[B,3136]   = dh1[B,1024] (K)   x W3[3136,1024] (K)
//   conv2 wgrad dW2[800,64]   = col[M2,800] (MN)  x dy2[M2,64]   (MN)      K = M2, split-K
//   conv2 dgrad dcol[M2,800]  = dy2[M2,64]  (K)   x W2[800,64]   (K)
// Weights: bf16 math reads the job's bf16 shadow of the parameter buffer, which the apply
// kernel (and tfg_var_set) keeps current; tf32 math reads the fp32 parameters themselves.
// Activations: a2 / dh1 get their bf16 copies from the kernels that produce them
// (pool_relu_kernel, fc2_bwd_kernel); tf32 reads the fp32 tensors in place.

struct TcSlot {            // per replica slot: operand copies of activations + GEMM descriptors
  void *col_lp = nullptr, *a2_lp = nullptr, *dh1_lp = nullptr, *dy2_lp = nullptr;   // a2/dh1: bf16 math only
  float *a1_lo = nullptr, *a2_lo = nullptr, *dh1_lo = nullptr, *dy2_lo = nullptr;   // 3xTF32: low-order planes
  float* partial = nullptr;
  TcGemm conv2_fwd, fc1_fwd, fc1_wgrad, fc1_dgrad, conv2_wgrad, conv2_dgrad;
  // 3xTF32: conv2 as implicit GEMMs (conv_x3.cuh): maps of the slabs (x = a1 / dy2 and their low-order
  // planes) and of W2 viewed [800, 64], per kernel
  CUtensorMap c2f_x, c2f_x_lo, c2f_w, c2f_w_lo;       // forward: a1 slabs, W2 MN-major boxes
  CUtensorMap c2d_x, c2d_x_lo, c2d_w, c2d_w_lo;       // data gradient: dy2 slabs, W2 K-major boxes
  CUtensorMap c2w_x, c2w_x_lo, c2w_dy, c2w_dy_lo;     // weight gradient: a1 slab and dy2 boxes, MN-major
  tc::WgradPlan wplan;
  float* wpartial = nullptr;                          // [nz][800][64]
};

constexpr int kSplitW2Tc = 21;   // 7 M-tiles x 21 K-splits = 147 CTAs on 148 SMs
constexpr int kSplitFc1Fwd = 9;  // B = 128: 16 N-tiles x 9 K-splits = 144 CTAs on 148 SMs
constexpr int kSplitFc1Dgrad = 3;  // 49 N-tiles x 3 = 147 CTAs

static inline int blocks_for(long long n, int per = 256) {
  long long b = (n + per - 1) / per;
  if (b > kNumSMs * 16) b = kNumSMs * 16;
  return (int)std::max<long long>(b, 1);
}

#define TC_LAUNCH_OK(j, what)                                     \
  do {                                                            \
    prof_mark((j), what);                                         \
    cudaError_t e__ = cudaGetLastError();                         \
    if (e__ != cudaSuccess) return fail_cuda((j), e__, what);     \
  } while (0)

void mnist_tc_destroy(tfg_job* j) {
  MnistModel* m = j->model;
  if (!m) return;
  for (auto& sl : m->slots) {
    if (!sl.tc) continue;
    TcSlot* t = static_cast<TcSlot*>(sl.tc);
    void* ps[] = {t->col_lp, t->a2_lp, t->dh1_lp, t->dy2_lp, t->partial, t->a1_lo, t->a2_lo, t->dh1_lo, t->dy2_lo, t->wpartial};
    for (void* p : ps) cudaFree(p);
    delete t;
    sl.tc = nullptr;
  }
}

void* mnist_tc_dh1_bf16(tfg_job* j, int slot) {
  MnistModel* m = j->model;
  if (!m || m->math != TFG_MATH_BF16 || !m->slots[slot].tc) return nullptr;
  return static_cast<TcSlot*>(m->slots[slot].tc)->dh1_lp;
}
float* mnist_tc_dh1_lo(tfg_job* j, int slot) {
  MnistModel* m = j->model;
  if (!m || m->math != TFG_MATH_TF32X3 || !m->slots[slot].tc) return nullptr;
  return static_cast<TcSlot*>(m->slots[slot].tc)->dh1_lo;
}

// Allocations only; tensor maps need the flat parameter/gradient buffers, which exist after
// tfg_vars_finalize, so they are encoded lazily on the first forward (mnist_tc_prepare).
int mnist_tc_build(tfg_job* j) {
  MnistModel* m = j->model;
  const bool bf = m->math == TFG_MATH_BF16;
  const size_t es = bf ? 2 : 4;
  const long long B = m->B, M2 = B * 196;
  cudaError_t e = cudaSuccess;
  auto alloc = [&](void** p, size_t bytes) { if (e == cudaSuccess) e = cudaMalloc(p, bytes ? bytes : 16); };
  const bool x3 = m->math == TFG_MATH_TF32X3;
  for (auto& sl : m->slots) {
    TcSlot* t = new TcSlot();
    sl.tc = t;
    if (!x3) alloc(&t->col_lp, (size_t)M2 * 800 * es);      // 3xTF32 runs conv2 without an im2col matrix
    alloc(&t->dy2_lp, (size_t)M2 * 64 * es);
    if (bf) {
      alloc(&t->a2_lp, (size_t)B * 3136 * es);
      alloc(&t->dh1_lp, (size_t)B * 1024 * es);
    }
    if (x3) {
      // CTAs per M-tile group of the conv2 weight gradient: groups 0..2 hold two M tiles, group 3 one
      const int tiles = (int)B * tc::kWTilesPerImage;
      const int nz2 = std::min(42, tiles), nz1 = std::min(kNumSMs - 3 * 42, tiles);
      t->wplan.cta0[0] = 0;
      for (int g = 0; g < 4; ++g) t->wplan.cta0[g + 1] = t->wplan.cta0[g] + (g < 3 ? nz2 : nz1);
      alloc((void**)&t->wpartial, (size_t)nz2 * 800 * 64 * 4);
      alloc((void**)&t->a1_lo, (size_t)B * 6272 * 4);
      alloc((void**)&t->dy2_lo, (size_t)M2 * 64 * 4);
      alloc((void**)&t->a2_lo, (size_t)B * 3136 * 4);
      alloc((void**)&t->dh1_lo, (size_t)B * 1024 * 4);
    }
    alloc((void**)&t->partial, sizeof(float) * std::max<size_t>({(size_t)kSplitW2Tc * 800 * 64,
                                                                  (size_t)kSplitFc1Fwd * B * 1024,
                                                                  (size_t)kSplitFc1Dgrad * B * 3136}));
  }
  if (e != cudaSuccess) return fail_cuda(j, e, "cudaMalloc(tensor-path operands)");
  return TFG_OK;
}

static int mnist_tc_prepare(tfg_job* j, int slot) {
  MnistModel* m = j->model;
  MnistSlot& s = m->slots[slot];
  TcSlot* t = static_cast<TcSlot*>(s.tc);
  if (t->fc1_fwd.C) return TFG_OK;
  const bool x3 = m->math == TFG_MATH_TF32X3;
  const int kind = x3 ? TFG_MATH_TF32 : m->math;      // element kind of the operands
  const bool bf = kind == TFG_MATH_BF16;
  if ((bf || x3) && !j->params_lp) return fail(j, TFG_ESTATE, "this math mode needs the parameter shadow");
  auto wgt_lo = [&](int v) -> const void* { return static_cast<const float*>(j->params_lp) + j->vars[m->var[v]].offset; };
  auto wgt = [&](int v) -> const void* {   // weight operand of variable v
    const int64_t off = j->vars[m->var[v]].offset;
    return bf ? static_cast<const void*>(static_cast<const char*>(j->params_lp) + 2 * off)
              : static_cast<const void*>(j->params + off);
  };
  const void* w2 = wgt(2);
  const void* w3 = wgt(4);
  const void* a2 = bf ? t->a2_lp : static_cast<const void*>(s.a2);
  const void* dh1 = bf ? t->dh1_lp : static_cast<const void*>(s.dh1);
  const int B = m->B, M2 = B * 196;
  const float* P = j->params;
  float* G = j->grads[slot];
  auto par = [&](int v) { return P + j->vars[m->var[v]].offset; };
  auto grd = [&](int v) { return G + j->vars[m->var[v]].offset; };
  int rc;
  // operand = (pointer, row-major pitch, mn_major); rows/K follow from the GEMM shape
  auto mk = [&](TcGemm& g, const char* name, int M, int N, int K, const void* A, long long lda, bool a_mn,
                const void* Bp, long long ldb, bool b_mn, float* C, int block_n, int splits,
                tc::Epilogue epi, const void* A_lo = nullptr, const void* B_lo = nullptr) -> int {
    g.name = name; g.M = M; g.N = N; g.K = K; g.C = C; g.ldc = N;
    g.block_n = block_n; g.splits = splits; g.epi = epi; g.partial = t->partial;
    g.a_mn = a_mn; g.b_mn = b_mn;
    g.x3 = x3;
    if (x3) {
      if ((rc = make_map(&g.a_lo, kind, A_lo, M, K, lda, tc::kBlockM, a_mn))) return rc;
      if ((rc = make_map(&g.b_lo, kind, B_lo, N, K, ldb, block_n, b_mn))) return rc;
    }
    if ((rc = make_map(&g.a, kind, A, M, K, lda, tc::kBlockM, a_mn))) return rc;
    return make_map(&g.b, kind, Bp, N, K, ldb, block_n, b_mn);
  };
  const void *w2l = x3 ? wgt_lo(2) : nullptr, *w3l = x3 ? wgt_lo(4) : nullptr;
  if (x3) {
    // conv2 forward / data gradient / weight gradient as implicit GEMMs over shared-memory slabs
    if ((rc = make_map_4d_f32(&t->c2f_x, s.a1, B, 14, 14, 32, 18, 18, false))) return rc;
    if ((rc = make_map_4d_f32(&t->c2f_x_lo, t->a1_lo, B, 14, 14, 32, 18, 18, false))) return rc;
    if ((rc = make_map(&t->c2f_w, kind, w2, 64, 800, 64, 64, true))) return rc;          // boxes {32 n, 32 k lines}
    if ((rc = make_map(&t->c2f_w_lo, kind, w2l, 64, 800, 64, 64, true))) return rc;
    if ((rc = make_map_4d_f32(&t->c2d_x, t->dy2_lp, B, 14, 14, 64, 18, 18, false))) return rc;
    if ((rc = make_map_4d_f32(&t->c2d_x_lo, t->dy2_lo, B, 14, 14, 64, 18, 18, false))) return rc;
    if ((rc = make_map(&t->c2d_w, kind, w2, 800, 64, 64, 32, false))) return rc;         // boxes {32 k, 32 rows}
    if ((rc = make_map(&t->c2d_w_lo, kind, w2l, 800, 64, 64, 32, false))) return rc;
    if ((rc = make_map_4d_f32(&t->c2w_x, s.a1, B, 14, 14, 32, 18, tc::kWTileRows + 4, true))) return rc;
    if ((rc = make_map_4d_f32(&t->c2w_x_lo, t->a1_lo, B, 14, 14, 32, 18, tc::kWTileRows + 4, true))) return rc;
    if ((rc = make_map_4d_f32(&t->c2w_dy, t->dy2_lp, B, 14, 14, 64, 18, tc::kWTileRows, true))) return rc;
    if ((rc = make_map_4d_f32(&t->c2w_dy_lo, t->dy2_lo, B, 14, 14, 64, 18, tc::kWTileRows, true))) return rc;
  } else if ((rc = mk(t->conv2_fwd, "tc gemm conv2 fwd", M2, 64, 800, t->col_lp, 800, false, w2, 64, true, s.y2, 64, 1,
               tc::Epilogue{par(3), nullptr, 0, nullptr, 0}))) return rc;
  if ((rc = mk(t->fc1_fwd, "tc gemm fc1 fwd", B, 1024, 3136, a2, 3136, false, w3, 1024, true, s.h1, 64,
               kSplitFc1Fwd, tc::Epilogue{par(5), nullptr, 1, nullptr, 0}, t->a2_lo, w3l))) return rc;
  if ((rc = mk(t->fc1_wgrad, "tc gemm fc1 wgrad", 3136, 1024, B, a2, 3136, true, dh1, 1024, true, grd(4),
               // 3xTF32: 400 tiles of 128 x 64, two CTAs per SM (measured 20 vs 27 us for 200 tiles of 128 x 128)
               x3 ? 64 : 128, 1, tc::Epilogue{nullptr, nullptr, 0, nullptr, 0}, t->a2_lo, t->dh1_lo))) return rc;
  if ((rc = mk(t->fc1_dgrad, "tc gemm fc1 dgrad", B, 3136, 1024, dh1, 1024, false, w3, 1024, false, s.da2,
               64, kSplitFc1Dgrad, tc::Epilogue{nullptr, nullptr, 0, nullptr, 0}, t->dh1_lo, w3l))) return rc;
  if (x3) {   // their consumers (head_fwd_kernel, unpool_x3_kernel) add the K-split partial sums themselves
    t->fc1_fwd.reduce_later = true;
    t->fc1_dgrad.reduce_later = true;
  }
  if (!x3) {
    if ((rc = mk(t->conv2_wgrad, "tc gemm conv2 wgrad", 800, 64, M2, t->col_lp, 800, true, t->dy2_lp, 64, true, grd(2),
                 64, kSplitW2Tc, tc::Epilogue{nullptr, nullptr, 0, nullptr, 0}))) return rc;
    if ((rc = mk(t->conv2_dgrad, "tc gemm conv2 dgrad", M2, 800, 64, t->dy2_lp, 64, false, w2, 64, false, s.dcol,
                 128, 1, tc::Epilogue{nullptr, nullptr, 0, nullptr, 0}))) return rc;
  }
  if (bf) {   // dcol is the step's largest tensor: keep it in bf16, written through the TMA store epilogue
    t->conv2_dgrad.epi.out_bf16 = s.dcol;
    t->conv2_dgrad.tma_out = 1;
    if ((rc = make_map(&t->conv2_dgrad.c, kind, s.dcol, M2, 800, 800, tc::kBlockM, false))) return rc;
  }
  return TFG_OK;
}

template <bool DGRAD>
static int launch_conv2_x3(tfg_job* j, const char* name, const CUtensorMap& x, const CUtensorMap& x_lo, const CUtensorMap& w,
                           const CUtensorMap& w_lo, float* out, const float* bias, const float* mask, int B, cudaStream_t st,
                           float* pool_a = nullptr, float* pool_lo = nullptr, uint8_t* pool_idx = nullptr) {
  static bool attr_set = false;
  constexpr size_t smem = tc::smem_bytes_c2<DGRAD>();
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(tc::conv2_x3_kernel<DGRAD>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return fail_cuda(j, e, name);
    attr_set = true;
  }
  // a chunk of the main product: three taps of 32 input channels forward (12 accumulating MMAs; the [main |
  // small] accumulator is twice as wide to read back, so the chunk is longer than the GEMM's), one tap of 64
  // channels backward (8)
  tc::conv2_x3_kernel<DGRAD><<<std::min(B, kNumSMs), tc::kC2Threads, smem, st>>>(x, x_lo, w, w_lo, out, bias, mask, B,
                                                                                  DGRAD ? 1 : 3, pool_a, pool_lo, pool_idx);
  TC_LAUNCH_OK(j, name);
  return TFG_OK;
}

static int launch_conv2_wgrad_x3(tfg_job* j, TcSlot* t, float* dw, int B, cudaStream_t st) {
  static bool attr_set = false;
  constexpr size_t smem = tc::smem_bytes_c2_wgrad();
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(tc::conv2_wgrad_x3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return fail_cuda(j, e, "conv2 wgrad x3");
    attr_set = true;
  }
  tc::conv2_wgrad_x3_kernel<<<t->wplan.cta0[4], tc::kC2Threads, smem, st>>>(t->c2w_x, t->c2w_x_lo, t->c2w_dy, t->c2w_dy_lo,
                                                                            t->wpartial, B, t->wplan);
  TC_LAUNCH_OK(j, "conv2 wgrad x3");
  tc::conv2_wgrad_reduce_kernel<<<(800 * 16 * 4 + 255) / 256, 256, 0, st>>>(t->wpartial, dw, t->wplan);
  TC_LAUNCH_OK(j, "conv2 wgrad reduce");
  return TFG_OK;
}

template <typename T>
static int tc_forward_t(tfg_job* j, int slot) {
  MnistModel* m = j->model;
  MnistSlot& s = m->slots[slot];
  TcSlot* t = static_cast<TcSlot*>(s.tc);
  cudaStream_t st = j->stream;
  const int B = m->B, M2 = B * 196;
  int rc = mnist_tc_prepare(j, slot);
  if (rc) return rc;
  const int kind = sizeof(T) == 2 ? TFG_MATH_BF16 : TFG_MATH_TF32;
  if (m->math == TFG_MATH_TF32X3) {
    const float* b2 = j->params + j->vars[m->var[3]].offset;
    // with the 2x2 max-pool + relu in its epilogue: a2, its low-order plane and the arg-max codes; y2 is never written
    if ((rc = launch_conv2_x3<false>(j, "conv2 fwd x3", t->c2f_x, t->c2f_x_lo, t->c2f_w, t->c2f_w_lo, nullptr, b2, nullptr, B, st,
                                     s.a2, t->a2_lo, s.idx2)))
      return rc;
  } else {
    // conv2 as im2col GEMM
    im2col_lp_kernel<T><<<blocks_for((long long)M2 * 25 * (32 / (16 / (int)sizeof(T)))), 256, 0, st>>>(s.a1, (T*)t->col_lp, B);
    TC_LAUNCH_OK(j, "im2col_lp_kernel");
    if ((rc = launch_tc(j, t->conv2_fwd, kind, st))) return rc;
  }
  if (m->math != TFG_MATH_TF32X3 && (rc = mnist_pool2(j, slot, t->a2_lp, t->a2_lo))) return rc;   // also drops the operand copy of a2
  return launch_tc(j, t->fc1_fwd, kind, st);
}

template <typename T>
static int tc_backward_t(tfg_job* j, int slot) {
  MnistModel* m = j->model;
  MnistSlot& s = m->slots[slot];
  TcSlot* t = static_cast<TcSlot*>(s.tc);
  cudaStream_t st = j->stream;
  const int B = m->B;
  int rc;
  const int kind = sizeof(T) == 2 ? TFG_MATH_BF16 : TFG_MATH_TF32;
  const bool x3 = m->math == TFG_MATH_TF32X3;
  // Weight gradients are leaves of the step.  In the 3xTF32 path they run on the side branch (joined at the
  // end of mnist_forward_backward) while the data-gradient chain goes on: their CTAs fill the SMs the
  // chain's one-wave kernels leave idle and their heads and tails overlap the chain's.
  cudaStream_t leaf = x3 ? side_fork(j) : st;                   // ordered after fc2_bwd_kernel (dh1)
  if ((rc = launch_tc(j, t->fc1_wgrad, kind, leaf))) return rc;   // dh1 (bf16 / low-order plane) came from fc2_bwd_kernel
  if ((rc = launch_tc(j, t->fc1_dgrad, kind, st))) return rc;
  if (x3) {
    // fc1_dgrad left its K-split partial sums in t->partial (reduce_later); the bias partials go to the da2 buffer,
    // which nothing else uses in this path
    if (tc_splits(t->fc1_dgrad, kind) > kMaxUnpoolSplits) return fail(j, TFG_ESTATE, "fc1 data gradient uses more K splits than unpool_x3_kernel adds");
    unpool_x3_kernel<<<kUnpoolBlocks, 256, 0, st>>>(t->partial, tc_splits(t->fc1_dgrad, kind), s.a2, s.idx2, (float*)t->dy2_lp,
                                                    t->dy2_lo, s.da2, B);
    TC_LAUNCH_OK(j, "unpool_x3_kernel");
    leaf = side_fork(j);                                        // ordered after the un-pooling (dy2 and its bias partials)
    if ((rc = mnist_reduce_rows(j, s.da2, kUnpoolBlocks, 64, j->grads[slot] + j->vars[m->var[3]].offset, leaf))) return rc;   // db2
  } else {
    unpool_lp_kernel<T><<<blocks_for((long long)B * 3136), 256, 0, st>>>(s.da2, s.a2, s.idx2, (T*)t->dy2_lp, t->dy2_lo, B, 7, 64);
    TC_LAUNCH_OK(j, "unpool_lp_kernel");
  }
  if (x3) {
    if ((rc = launch_conv2_wgrad_x3(j, t, j->grads[slot] + j->vars[m->var[2]].offset, B, leaf))) return rc;
    // dx straight into g1m, masked by a1 > 0: no dcol, no col2im
    return launch_conv2_x3<true>(j, "conv2 dgrad x3", t->c2d_x, t->c2d_x_lo, t->c2d_w, t->c2d_w_lo, s.g1m, nullptr, s.a1, B, st);
  }
  if ((rc = launch_tc(j, t->conv2_wgrad, kind, st))) return rc;
  return launch_tc(j, t->conv2_dgrad, kind, st);
}

const float* mnist_tc_fc1_partial(tfg_job* j, int slot, int* splits) {
  MnistModel* m = j->model;
  if (!m || m->math != TFG_MATH_TF32X3 || !m->slots[slot].tc) return nullptr;
  TcSlot* t = static_cast<TcSlot*>(m->slots[slot].tc);
  *splits = tc_splits(t->fc1_fwd, TFG_MATH_TF32);
  return *splits > 1 ? t->partial : nullptr;
}

float* mnist_tc_a1_lo(tfg_job* j, int slot) {
  MnistModel* m = j->model;
  if (!m || m->math != TFG_MATH_TF32X3 || !m->slots[slot].tc) return nullptr;
  return static_cast<TcSlot*>(m->slots[slot].tc)->a1_lo;
}

int mnist_tc_forward(tfg_job* j, int slot) {
  return j->model->math == TFG_MATH_BF16 ? tc_forward_t<__nv_bfloat16>(j, slot) : tc_forward_t<float>(j, slot);
}
int mnist_tc_backward(tfg_job* j, int slot) {
  return j->model->math == TFG_MATH_BF16 ? tc_backward_t<__nv_bfloat16>(j, slot) : tc_backward_t<float>(j, slot);
}

}  // namespace tfg

// ---- stateless GEMM entry points (parity tests of the tensor kernel alone) -------------------
static int gemm_tc_entry(int math, const void* a_dev, const void* a_lo, int a_mn_major, const void* b_dev,
                         const void* b_lo, int b_mn_major, float* c_dev, int M, int N, int K, int64_t lda, int64_t ldb,
                         int block_n, int splits, float* partial_dev, const float* bias_dev, int relu, void* stream,
                         int chunk_kb = 0) {
  using namespace tfg;
  if (!a_dev || !b_dev || !c_dev || M < 1 || N < 1 || K < 1) return fail(nullptr, TFG_EINVAL, "bad argument");
  if (block_n != 64 && block_n != 128) return fail(nullptr, TFG_EINVAL, "block_n must be 64 or 128");
  TcGemm g;
  g.name = "tc gemm";
  g.M = M; g.N = N; g.K = K; g.C = c_dev; g.ldc = N; g.block_n = block_n; g.splits = splits < 1 ? 1 : splits;
  g.partial = partial_dev;
  g.a_mn = a_mn_major != 0; g.b_mn = b_mn_major != 0;
  g.epi = tc::Epilogue{bias_dev, nullptr, relu, nullptr, 0};
  int rc = make_map(&g.a, math, a_dev, M, K, lda, tc::kBlockM, g.a_mn);
  if (rc) return rc;
  if ((rc = make_map(&g.b, math, b_dev, N, K, ldb, block_n, g.b_mn))) return rc;
  if (a_lo) {
    g.x3 = true;
    g.chunk_kb = chunk_kb;
    if ((rc = make_map(&g.a_lo, math, a_lo, M, K, lda, tc::kBlockM, g.a_mn))) return rc;
    if ((rc = make_map(&g.b_lo, math, b_lo, N, K, ldb, block_n, g.b_mn))) return rc;
  }
  return launch_tc(nullptr, g, math, (cudaStream_t)stream);
}

extern "C" int tfg_gemm_tc(int math, const void* a_dev, int a_mn_major, const void* b_dev, int b_mn_major,
                           float* c_dev, int M, int N, int K, int64_t lda, int64_t ldb, int block_n, int splits,
                           float* partial_dev, const float* bias_dev, int relu, void* stream) {
  using namespace tfg;
  if (math != TFG_MATH_BF16 && math != TFG_MATH_TF32) return fail(nullptr, TFG_EINVAL, "math must be bf16 or tf32");
  return gemm_tc_entry(math, a_dev, nullptr, a_mn_major, b_dev, nullptr, b_mn_major, c_dev, M, N, K, lda, ldb, block_n,
                       splits, partial_dev, bias_dev, relu, stream);
}

extern "C" int tfg_gemm_tc_x3(const float* a_dev, const float* a_lo_dev, int a_mn_major, const float* b_dev,
                              const float* b_lo_dev, int b_mn_major, float* c_dev, int M, int N, int K, int64_t lda,
                              int64_t ldb, int block_n, int splits, int chunk_kb, float* partial_dev, const float* bias_dev,
                              int relu, void* stream) {
  using namespace tfg;
  if (!a_lo_dev || !b_lo_dev) return fail(nullptr, TFG_EINVAL, "the split product needs both low-order planes");
  if (chunk_kb < 0) return fail(nullptr, TFG_EINVAL, "chunk_kb must be >= 0");
  return gemm_tc_entry(TFG_MATH_TF32, a_dev, a_lo_dev, a_mn_major, b_dev, b_lo_dev, b_mn_major, c_dev, M, N, K, lda, ldb,
                       block_n, splits, partial_dev, bias_dev, relu, stream, chunk_kb);
}

extern "C" int tfg_tf32_split_lo(const float* src_dev, float* lo_dev, int64_t n, void* stream) {
  using namespace tfg;
  if (!src_dev || !lo_dev || n < 0) return fail(nullptr, TFG_EINVAL, "bad argument");
  return launch_cast_lp(src_dev, lo_dev, n, 2, (cudaStream_t)stream) ? fail(nullptr, TFG_ECUDA, "split kernel launch failed") : TFG_OK;
}

#ifdef TFG_TRACE
// development aid: the per-CTA wait counters of the conv2 forward (which = 0) / data gradient (1) kernel
extern "C" int tfg_debug_c2_trace(int which, long long* out /* [160][16] */) {
  using namespace tfg;
  cudaError_t e = cudaMemcpyFromSymbol(out, tc::g_c2_trace, sizeof(long long) * 160 * 16,
                                       (size_t)which * 160 * 16 * sizeof(long long));
  return e == cudaSuccess ? TFG_OK : fail_cuda(nullptr, e, "cudaMemcpyFromSymbol(trace)");
}
#endif
