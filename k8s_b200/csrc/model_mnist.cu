// The sample model of the sync-PS step: the canonical mnist_deep
// CNN, forward and backward, one local replica at a time.
//
//   x[B,28,28,1] -conv5x5(1->32,SAME)+b-> pool2x2 -> relu  = a1[B,14,14,32]
//                -conv5x5(32->64,SAME)+b-> pool2x2 -> relu = a2[B,7,7,64]
//                -fc 3136->1024 +b, relu-> h1 -fc 1024->10 +b-> logits -> mean softmax xent
//
// NHWC activations, HWIO filters (TensorFlow's layouts).  relu(maxpool(z)) is computed as
// maxpool-then-relu: both are monotone, so values and gradients equal TF's
// conv->relu->pool order, and only the pooled map plus a 2-bit argmax is stored.
//
// Contractions (conv2 as im2col GEMM, fc1, their dgrad/wgrad) run on the tensor path when
// the job's math mode asks for it (gemm_tc.cu); everything else -- conv1 (K = 25, far too
// thin for a 128-wide MMA tile), pooling, bias reductions, softmax/loss -- is HBM-bound
// work in plain CUDA-core / warp-shuffle kernels here.
#include <algorithm>

#include "common.cuh"
#include "gemm_tc.cuh"
#include "model_mnist.cuh"

namespace tfg {

// ===== conv1 forward, fused with bias + 2x2 max-pool + relu ===============================
// kConv1Split blocks per image (each stages the whole 3-KB image and takes a quarter of the 14x14x32
// outputs: one block per image left 20 of the 148 SMs idle and 8 warps per SM to hide the
// shared-memory latency).  Lane = output channel (32), so input reads broadcast and filter reads are
// conflict-free.
constexpr int kConv1Split = 4;
constexpr int kConv1Part = 14 * 14 * 32 / kConv1Split;     // 1568 outputs per block

__global__ void __launch_bounds__(256)
conv1_pool_relu_kernel(const float* __restrict__ x, const float* __restrict__ w1,
                       const float* __restrict__ b1, float* __restrict__ a1,
                       uint8_t* __restrict__ idx1, float* __restrict__ a1_lo /* 3xTF32: low-order plane of a1, or NULL */) {
  __shared__ float xs[32][33];   // image with a 2-pixel zero border
  __shared__ float ws[25][32];
  __shared__ float bs[32];
  const int b = blockIdx.x / kConv1Split, part = blockIdx.x % kConv1Split, t = threadIdx.x;
  for (int i = t; i < 32 * 32; i += 256) {
    int r = i >> 5, c = i & 31;
    int ir = r - 2, ic = c - 2;
    xs[r][c] = (ir >= 0 && ir < 28 && ic >= 0 && ic < 28) ? x[(size_t)b * 784 + ir * 28 + ic] : 0.f;
  }
  for (int i = t; i < 800; i += 256) ws[i >> 5][i & 31] = w1[i];
  if (t < 32) bs[t] = b1[t];
  __syncthreads();
  for (int o = part * kConv1Part + t; o < (part + 1) * kConv1Part; o += 256) {
    const int c = o & 31, pw = (o >> 5) % 14, ph = (o >> 5) / 14;
    float patch[6][6];
#pragma unroll
    for (int r = 0; r < 6; ++r)
#pragma unroll
      for (int q = 0; q < 6; ++q) patch[r][q] = xs[2 * ph + r][2 * pw + q];
    float best = 0.f;
    int arg = 0;
#pragma unroll
    for (int dy = 0; dy < 2; ++dy)
#pragma unroll
      for (int dx = 0; dx < 2; ++dx) {
        float acc = 0.f;
#pragma unroll
        for (int kh = 0; kh < 5; ++kh)
#pragma unroll
          for (int kw = 0; kw < 5; ++kw) acc = fmaf(patch[dy + kh][dx + kw], ws[kh * 5 + kw][c], acc);
        acc += bs[c];
        if ((dy | dx) == 0 || acc > best) { best = acc; arg = dy * 2 + dx; }
      }
    const size_t oi = (size_t)b * (14 * 14 * 32) + o;
    a1[oi] = fmaxf(best, 0.f);
    if (a1_lo) a1_lo[oi] = tc::tf32_lo(fmaxf(best, 0.f));
    idx1[oi] = (uint8_t)arg;
  }
}

// ===== conv1 backward: dW1[25,32], db1[32] partials per image ============================
// g1m = gradient at the pooled pre-relu map (already masked by a1 > 0).
__global__ void __launch_bounds__(256)
conv1_bwd_kernel(const float* __restrict__ x, const float* __restrict__ g1m,
                 const uint8_t* __restrict__ idx1, float* __restrict__ partial /*[kConv1Split * B][832]*/) {
  __shared__ float xs[32][33];
  __shared__ float red[8][26][32];
  const int b = blockIdx.x / kConv1Split, part = blockIdx.x % kConv1Split, t = threadIdx.x, w = t >> 5, c = t & 31;
  // this thread's gradients and arg-max codes, fetched while the image is staged (the loop below
  // used to pay one global-load latency per output)
  constexpr int kIters = (kConv1Part + 255) / 256;
  float gs[kIters];
  int as[kIters];
#pragma unroll
  for (int u = 0; u < kIters; ++u) {
    const int o = part * kConv1Part + t + 256 * u;
    if (o < (part + 1) * kConv1Part) {
      const size_t oi = (size_t)b * (14 * 14 * 32) + o;
      gs[u] = g1m[oi];
      as[u] = idx1[oi];
    }
  }
  for (int i = t; i < 32 * 32; i += 256) {
    int r = i >> 5, q = i & 31;
    int ir = r - 2, ic = q - 2;
    xs[r][q] = (ir >= 0 && ir < 28 && ic >= 0 && ic < 28) ? x[(size_t)b * 784 + ir * 28 + ic] : 0.f;
  }
  __syncthreads();
  float acc[25];
  float accb = 0.f;
#pragma unroll
  for (int k = 0; k < 25; ++k) acc[k] = 0.f;
  // this block's quarter of the pooled map
#pragma unroll
  for (int u = 0; u < kIters; ++u) {
    const int o = part * kConv1Part + t + 256 * u;   // o & 31 == c
    if (o >= (part + 1) * kConv1Part) break;
    const int pw = (o >> 5) % 14, ph = (o >> 5) / 14;
    const float g = gs[u];
    const int a = as[u];
    const int r0 = 2 * ph + (a >> 1), c0 = 2 * pw + (a & 1);
    accb += g;
#pragma unroll
    for (int kh = 0; kh < 5; ++kh)
#pragma unroll
      for (int kw = 0; kw < 5; ++kw) acc[kh * 5 + kw] = fmaf(xs[r0 + kh][c0 + kw], g, acc[kh * 5 + kw]);
  }
#pragma unroll
  for (int k = 0; k < 25; ++k) red[w][k][c] = acc[k];
  red[w][25][c] = accb;
  __syncthreads();
  for (int i = t; i < 26 * 32; i += 256) {
    const int k = i >> 5, cc = i & 31;
    float s = red[0][k][cc];
#pragma unroll
    for (int ww = 1; ww < 8; ++ww) s += red[ww][k][cc];
    partial[(size_t)blockIdx.x * 832 + i] = s;   // [25*32 dW | 32 db], kConv1Split rows per image
  }
}

// ===== im2col for conv2: a1[B,14,14,32] -> col[B*196, 800], k = (kh*5+kw)*32 + c ===========
template <typename T>
__global__ void __launch_bounds__(256)
im2col_kernel(const float* __restrict__ a1, T* __restrict__ col, T* __restrict__ col_t, int B) {
  // one thread per (row m, tap kk, 4 channels)
  const long long total = (long long)B * 196 * 25 * 8;
  const long long M = (long long)B * 196;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c4 = (int)(i & 7);
    const int kk = (int)((i >> 3) % 25);
    const long long m = (i >> 3) / 25;
    const int ow = (int)(m % 14), oh = (int)((m / 14) % 14);
    const long long b = m / 196;
    const int ih = oh + kk / 5 - 2, iw = ow + kk % 5 - 2;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (ih >= 0 && ih < 14 && iw >= 0 && iw < 14)
      v = *reinterpret_cast<const float4*>(a1 + ((b * 14 + ih) * 14 + iw) * 32 + c4 * 4);
    const int k = kk * 32 + c4 * 4;
    store4(col + m * 800 + k, v);
    if (col_t) {
      col_t[(long long)(k + 0) * M + m] = cvt<T>(v.x);
      col_t[(long long)(k + 1) * M + m] = cvt<T>(v.y);
      col_t[(long long)(k + 2) * M + m] = cvt<T>(v.z);
      col_t[(long long)(k + 3) * M + m] = cvt<T>(v.w);
    }
  }
}

// ===== pool2 forward: y2[B,14,14,64] (bias included) -> a2[B,7,7,64] + argmax =============
__global__ void __launch_bounds__(256)
pool_relu_kernel(const float* __restrict__ y, float* __restrict__ a, uint8_t* __restrict__ idx,
                 __nv_bfloat16* __restrict__ a_bf /* optional operand-precision copy */,
                 float* __restrict__ a_lo /* optional low-order plane of the 3xTF32 split */, int B, int HO, int C) {
  const long long total = (long long)B * HO * HO * C;
  const int HI = HO * 2;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    const int pw = (int)((i / C) % HO), ph = (int)((i / C / HO) % HO);
    const long long b = i / C / HO / HO;
    float best = 0.f;
    int arg = 0;
#pragma unroll
    for (int dy = 0; dy < 2; ++dy)
#pragma unroll
      for (int dx = 0; dx < 2; ++dx) {
        float v = y[((b * HI + 2 * ph + dy) * HI + 2 * pw + dx) * C + c];
        if ((dy | dx) == 0 || v > best) { best = v; arg = dy * 2 + dx; }
      }
    a[i] = fmaxf(best, 0.f);
    if (a_bf) a_bf[i] = __float2bfloat16_rn(fmaxf(best, 0.f));
    if (a_lo) a_lo[i] = tc::tf32_lo(fmaxf(best, 0.f));
    idx[i] = (uint8_t)arg;
  }
}

// ===== un-pool backward: gp[B,HO,HO,C] (grad at pooled post-relu map) -> dense gy =========
template <typename T>
__global__ void __launch_bounds__(256)
unpool_kernel(const float* __restrict__ gp, const float* __restrict__ a, const uint8_t* __restrict__ idx,
              float* __restrict__ gy, T* __restrict__ gy_lp, T* __restrict__ gy_t, int B, int HO, int C) {
  const long long total = (long long)B * HO * HO * C;
  const int HI = HO * 2;
  const long long M = (long long)B * HI * HI;
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
        const float v = (dy * 2 + dx == arg) ? g : 0.f;
        const long long m = (b * HI + 2 * ph + dy) * HI + 2 * pw + dx;
        if (gy) gy[m * C + c] = v;
        if (gy_lp) gy_lp[m * C + c] = cvt<T>(v);
        if (gy_t) gy_t[(long long)c * M + m] = cvt<T>(v);
      }
  }
}

// masked pooled gradient (for the bias reduction): out = a > 0 ? g : 0
__global__ void __launch_bounds__(256)
relu_mask_kernel(const float* __restrict__ g, const float* __restrict__ a, float* __restrict__ out,
                 long long n) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x)
    out[i] = a[i] > 0.f ? g[i] : 0.f;
}

// ===== col2im: dcol[B*196, 800] -> g1m[B,14,14,32], masked by a1 > 0 ======================
// T = float, or bf16 when the tensor path's data-gradient GEMM stored dcol in bf16 (half the bytes
// of the step's largest tensor)
__device__ __forceinline__ float4 load4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ float4 load4(const __nv_bfloat16* p) {
  const uint2 u = *reinterpret_cast<const uint2*>(p);
  const float2 a = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&u.x));
  const float2 b = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&u.y));
  return make_float4(a.x, a.y, b.x, b.y);
}

template <typename T>
__global__ void __launch_bounds__(256)
col2im_mask_kernel(const T* __restrict__ dcol, const float* __restrict__ a1,
                   float* __restrict__ g1m, int B) {
  const long long total = (long long)B * 196 * 8;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c4 = (int)(i & 7);
    const long long pix = i >> 3;
    const int w = (int)(pix % 14), h = (int)((pix / 14) % 14);
    const long long b = pix / 196;
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int kh = 0; kh < 5; ++kh) {
      const int oh = h - kh + 2;
      if (oh < 0 || oh >= 14) continue;
#pragma unroll
      for (int kw = 0; kw < 5; ++kw) {
        const int ow = w - kw + 2;
        if (ow < 0 || ow >= 14) continue;
        const float4 v = load4(dcol + ((b * 14 + oh) * 14 + ow) * 800 + (kh * 5 + kw) * 32 + c4 * 4);
        s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
      }
    }
    const float4 av = *reinterpret_cast<const float4*>(a1 + pix * 32 + c4 * 4);
    s.x = av.x > 0.f ? s.x : 0.f; s.y = av.y > 0.f ? s.y : 0.f;
    s.z = av.z > 0.f ? s.z : 0.f; s.w = av.w > 0.f ? s.w : 0.f;
    *reinterpret_cast<float4*>(g1m + pix * 32 + c4 * 4) = s;
  }
}

// ===== CUDA-core SGEMM with strided operands (fp32 math mode and the thin GEMMs) ==========
// C[M,N] = A(m,k) * B(k,n) (+bias[n]) (relu) (mask by other > 0); A(m,k) = A[m*sam + k*sak],
// B(k,n) = B[k*sbk + n*sbn].  gridDim.z > 1 = split-K: partial products go to
// C + z*M*N and are summed by colsum_kernel-style reduce (deterministic, no atomics).
struct GemmEpi {
  const float* bias;    // [N] or NULL
  const float* mask;    // [M,N] or NULL: keep where mask > 0
  int relu;
};

template <bool A_KCONTIG, bool B_NCONTIG>
__global__ void __launch_bounds__(256)
sgemm_kernel(const float* __restrict__ A, const float* __restrict__ Bm, float* __restrict__ C,
             int M, int N, int K, long long sam, long long sak, long long sbk, long long sbn,
             int k_per_split, GemmEpi epi) {
  __shared__ float As[16][64 + 4];
  __shared__ float Bs[16][64 + 4];
  const int t = threadIdx.x, tx = t & 15, ty = t >> 4;
  const int m0 = blockIdx.y * 64, n0 = blockIdx.x * 64;
  const int kbeg = blockIdx.z * k_per_split;
  const int kend = min(K, kbeg + k_per_split);
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int jn = 0; jn < 4; ++jn) acc[i][jn] = 0.f;
  for (int k0 = kbeg; k0 < kend; k0 += 16) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      int m, k;
      if (A_KCONTIG) { k = t & 15; m = (t >> 4) + 16 * i; } else { m = t & 63; k = (t >> 6) + 4 * i; }
      const int gm = m0 + m, gk = k0 + k;
      As[k][m] = (gm < M && gk < kend) ? A[gm * sam + gk * sak] : 0.f;
      int n, kb;
      if (B_NCONTIG) { n = t & 63; kb = (t >> 6) + 4 * i; } else { kb = t & 15; n = (t >> 4) + 16 * i; }
      const int gn = n0 + n, gkb = k0 + kb;
      Bs[kb][n] = (gn < N && gkb < kend) ? Bm[gkb * sbk + gn * sbn] : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 16; ++k) {
      const float4 a = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
      const float4 b = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
      const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int jn = 0; jn < 4; ++jn) acc[i][jn] = fmaf(av[i], bv[jn], acc[i][jn]);
    }
    __syncthreads();
  }
  float* Cz = C + (size_t)blockIdx.z * M * N;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int gm = m0 + ty * 4 + i;
    if (gm >= M) continue;
#pragma unroll
    for (int jn = 0; jn < 4; ++jn) {
      const int gn = n0 + tx * 4 + jn;
      if (gn >= N) continue;
      float v = acc[i][jn];
      if (epi.bias) v += epi.bias[gn];
      if (epi.relu) v = fmaxf(v, 0.f);
      if (epi.mask) v = epi.mask[(size_t)gm * N + gn] > 0.f ? v : 0.f;
      Cz[(size_t)gm * N + gn] = v;
    }
  }
}

// ===== deterministic column sums: in[R,C] -> out[C] =======================================
// stage 1: partial[chunk][C] over 64-row chunks; stage 2: sum the chunks in order.
__global__ void __launch_bounds__(256)
colsum_stage1_kernel(const float* __restrict__ in, float* __restrict__ partial, int R, int C,
                     int rows_per_chunk) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const int r0 = blockIdx.y * rows_per_chunk, r1 = min(R, r0 + rows_per_chunk);
  float s = 0.f;
  for (int r = r0; r < r1; ++r) s += in[(size_t)r * C + c];
  partial[(size_t)blockIdx.y * C + c] = s;
}

__global__ void __launch_bounds__(256)
colsum_stage2_kernel(const float* __restrict__ partial, float* __restrict__ out, int chunks, int C,
                     long long stride) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  float s = 0.f;
  for (int k = 0; k < chunks; ++k) s += partial[(size_t)k * stride + c];
  out[c] = s;
}

// ===== softmax cross-entropy forward + backward, one warp per example ====================
__global__ void __launch_bounds__(256)
softmax_xent_kernel(const float* __restrict__ logits, const int32_t* __restrict__ labels,
                    float* __restrict__ dlogits, float* __restrict__ loss_vec, int B) {
  const int lane = threadIdx.x & 31;
  const int ex = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (ex >= B) return;
  const float z = lane < 10 ? logits[ex * 10 + lane] : -INFINITY;
  float mx = z;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  const float e = lane < 10 ? expf(z - mx) : 0.f;
  float sum = e;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  const int lab = labels[ex];
  const float zl = __shfl_sync(0xffffffffu, z, lab & 31);
  if (lane < 10) dlogits[ex * 10 + lane] = (e / sum - (lane == lab ? 1.f : 0.f)) / (float)B;
  if (lane == 0) loss_vec[ex] = logf(sum) - (zl - mx);
}


// ===== classifier head, forward: logits = h1 * W4 + b4, softmax cross-entropy, dlogits ======
// One block of 8 warps per example: warp w covers k in [128w, 128w+128), partial sums are
// combined by xor-shuffles and a fixed-order pass over the 8 warps.
constexpr int kMaxFusedSplits = 12;   // K splits whose partial sums head_fwd_kernel can add itself (fc1 forward uses 9)

__global__ void __launch_bounds__(256)
head_fwd_kernel(float* __restrict__ h1, const float* __restrict__ h1_partial, int splits, const float* __restrict__ b3,
                const float* __restrict__ w4, const float* __restrict__ b4,
                const int32_t* __restrict__ labels, float* __restrict__ logits, float* __restrict__ dlogits,
                float* __restrict__ loss_vec, int B, unsigned int* __restrict__ done, float* __restrict__ loss_out) {
  __shared__ float part[8][10];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int ex = blockIdx.x;
  float acc[10];
#pragma unroll
  for (int n = 0; n < 10; ++n) acc[n] = 0.f;
  float* h = h1 + (size_t)ex * 1024;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int k = w * 128 + i * 32 + lane;
    float hv;
    if (h1_partial) {   // h1 = relu(b3 + sum of fc1's K-split partial sums), added here in split order and kept for the backward pass
      float pv[kMaxFusedSplits];
#pragma unroll
      for (int z = 0; z < kMaxFusedSplits; ++z)      // all loads first, then the sum in split order
        if (z < splits) pv[z] = h1_partial[((size_t)z * B + ex) * 1024 + k];
      hv = 0.f;
#pragma unroll
      for (int z = 0; z < kMaxFusedSplits; ++z)
        if (z < splits) hv += pv[z];
      hv = fmaxf(hv + b3[k], 0.f);
      h[k] = hv;
    } else {
      hv = h[k];
    }
    const float2* wr = reinterpret_cast<const float2*>(w4 + k * 10);   // rows are 40 B: 8-B aligned
#pragma unroll
    for (int n2 = 0; n2 < 5; ++n2) {
      const float2 wv = wr[n2];
      acc[2 * n2] = fmaf(hv, wv.x, acc[2 * n2]);
      acc[2 * n2 + 1] = fmaf(hv, wv.y, acc[2 * n2 + 1]);
    }
  }
#pragma unroll
  for (int n = 0; n < 10; ++n) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc[n] += __shfl_xor_sync(0xffffffffu, acc[n], o);
    if (lane == 0) part[w][n] = acc[n];
  }
  __syncthreads();
  if (w != 0) return;
  float z = -INFINITY;
  if (lane < 10) {
    z = b4[lane];
#pragma unroll
    for (int ww = 0; ww < 8; ++ww) z += part[ww][lane];
  }
  float mx = z;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  const float e = lane < 10 ? expf(z - mx) : 0.f;
  float sum = e;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  const int lab = labels[ex];
  const float zl = __shfl_sync(0xffffffffu, z, lab & 31);
  if (lane < 10) {
    logits[ex * 10 + lane] = z;
    dlogits[ex * 10 + lane] = (e / sum - (lane == lab ? 1.f : 0.f)) / (float)B;
  }
  if (lane == 0) loss_vec[ex] = logf(sum) - (zl - mx);
  // the block that finishes last folds the per-example losses into the replica's mean, always in example
  // order (a launch of its own used to do this)
  __shared__ bool last;
  if (lane == 0) {
    __threadfence();
    last = atomicAdd(done, 1u) == gridDim.x - 1;
  }
  __syncwarp();
  if (last) {
    __threadfence();
    float s = 0.f;
    for (int i = lane; i < B; i += 32) s += __ldcg(loss_vec + i);   // lane l: examples l, l + 32, ... in order
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) { *loss_out = s / (float)B; *done = 0u; }
  }
}

// ===== classifier head, backward =========================================================
//   dW4[k,:] = sum_b h1[b,k] * dlogits[b,:]        dh1[b,k] = (h1[b,k] > 0) * sum_n dlogits[b,n] W4[k,n]
//   db3[k]   = sum_b dh1[b,k]                      db4[n]   = sum_b dlogits[b,n]
// One thread per hidden unit k and per chunk of 16 examples (blockIdx.y); chunk partials
// [chunk][1024*10 | 1024] are summed in chunk order by fc2_bwd_reduce_kernel: no atomics.
constexpr int kFc2Chunk = 16;
constexpr int kFc2PartialStride = 1024 * 11;

__global__ void __launch_bounds__(128)
fc2_bwd_kernel(const float* __restrict__ h1, const float* __restrict__ dlogits, const float* __restrict__ w4,
               float* __restrict__ partial, float* __restrict__ dh1, __nv_bfloat16* __restrict__ dh1_bf,
               float* __restrict__ dh1_lo, int B) {
  __shared__ float dl[kFc2Chunk][10];
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  const int b0 = blockIdx.y * kFc2Chunk;
  const int nb = min(kFc2Chunk, B - b0);
  for (int i = threadIdx.x; i < nb * 10; i += blockDim.x) dl[i / 10][i % 10] = dlogits[(size_t)b0 * 10 + i];
  float w[10], dw[10];
#pragma unroll
  for (int n = 0; n < 10; ++n) { w[n] = w4[k * 10 + n]; dw[n] = 0.f; }
  __syncthreads();
  float bsum = 0.f;
  float hvs[kFc2Chunk];
#pragma unroll
  for (int b = 0; b < kFc2Chunk; ++b)            // the chunk's activations in flight together
    if (b < nb) hvs[b] = h1[(size_t)(b0 + b) * 1024 + k];
#pragma unroll
  for (int b = 0; b < kFc2Chunk; ++b) {
    if (b >= nb) break;
    const float hv = hvs[b];
    float v = 0.f;
#pragma unroll
    for (int n = 0; n < 10; ++n) {
      const float d = dl[b][n];
      dw[n] = fmaf(hv, d, dw[n]);
      v = fmaf(d, w[n], v);
    }
    v = hv > 0.f ? v : 0.f;
    dh1[(size_t)(b0 + b) * 1024 + k] = v;
    if (dh1_bf) dh1_bf[(size_t)(b0 + b) * 1024 + k] = __float2bfloat16_rn(v);
    if (dh1_lo) dh1_lo[(size_t)(b0 + b) * 1024 + k] = tc::tf32_lo(v);
    bsum += v;
  }
  float* p = partial + (size_t)blockIdx.y * kFc2PartialStride;
#pragma unroll
  for (int n = 0; n < 10; ++n) p[k * 10 + n] = dw[n];
  p[10240 + k] = bsum;
}

__global__ void __launch_bounds__(256)
fc2_bwd_reduce_kernel(const float* __restrict__ partial, const float* __restrict__ dlogits,
                      float* __restrict__ dw4, float* __restrict__ db3, float* __restrict__ db4, int chunks,
                      int B) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < kFc2PartialStride) {
    float s = 0.f;
    for (int c = 0; c < chunks; ++c) s += partial[(size_t)c * kFc2PartialStride + i];
    if (i < 10240) dw4[i] = s; else db3[i - 10240] = s;
  } else if (i < kFc2PartialStride + 10) {
    const int n = i - kFc2PartialStride;
    float s = 0.f;
    for (int b = 0; b < B; ++b) s += dlogits[b * 10 + n];
    db4[n] = s;
  }
}

// ===== masked column sums with enough blocks to fill the chip ==============================
// in[R,C] (optionally zeroed where mask[r,c] <= 0) -> partial[chunk][C]; block = 64 columns x 4 row lanes
__global__ void __launch_bounds__(256)
colsum_wide_kernel(const float* __restrict__ in, const float* __restrict__ mask, float* __restrict__ partial,
                   int R, int C, int rows_per_chunk) {
  __shared__ float red[4][64];
  const int cx = threadIdx.x & 63, ry = threadIdx.x >> 6;
  const int c = blockIdx.x * 64 + cx;
  const int r0 = blockIdx.y * rows_per_chunk, r1 = min(R, r0 + rows_per_chunk);
  float s = 0.f;
  if (c < C)
    for (int r = r0 + ry; r < r1; r += 4) {
      const float v = in[(size_t)r * C + c];
      s += (!mask || mask[(size_t)r * C + c] > 0.f) ? v : 0.f;
    }
  red[ry][cx] = s;
  __syncthreads();
  if (ry == 0 && c < C) partial[(size_t)blockIdx.y * C + c] = ((red[0][cx] + red[1][cx]) + red[2][cx]) + red[3][cx];
}

// out[c] = sum over rows r of in[r*stride + c], rows added in a fixed order (row lane ty takes rows
// ty, ty+32, ...; the 32 lane sums are then added 0..31).  Columns below split_c go to out0, the
// rest to out1.  Block = 32 columns x 32 row lanes, so reads stay 128-byte coalesced and a thread's
// chain is R / 32 loads long (these reductions are latency-bound: few columns, a few hundred rows).
__global__ void __launch_bounds__(1024)
reduce_rows_kernel(const float* __restrict__ in, int R, int C, long long stride, float* __restrict__ out0,
                   int split_c, float* __restrict__ out1) {
  __shared__ float red[32][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + tx;
  float s = 0.f;
  if (c < C)
    for (int r = ty; r < R; r += 32) s += in[(size_t)r * stride + c];
  red[ty][tx] = s;
  __syncthreads();
  if (ty == 0 && c < C) {
    float t = red[0][tx];
#pragma unroll
    for (int k = 1; k < 32; ++k) t += red[k][tx];
    if (c < split_c) out0[c] = t; else out1[c - split_c] = t;
  }
}

// ---------------------------------------------------------------------------------------
static inline int blocks_for(long long n, int per = 256) {
  long long b = (n + per - 1) / per;
  if (b > kNumSMs * 16) b = kNumSMs * 16;
  return (int)std::max<long long>(b, 1);
}

#define LAUNCH_OK(j, what)                                          \
  do {                                                              \
    prof_mark((j), what);                                           \
    cudaError_t e__ = cudaGetLastError();                           \
    if (e__ != cudaSuccess) return fail_cuda((j), e__, what);       \
  } while (0)

// C = A*B through the CUDA-core kernel.  splits > 1 needs `partial` of splits*M*N floats.
static int sgemm(tfg_job* j, const char* name, const float* A, bool a_kcontig, long long sam, long long sak,
                 const float* Bm, bool b_ncontig, long long sbk, long long sbn, float* C, int M, int N,
                 int K, GemmEpi epi, int splits, float* partial) {
  dim3 grid((N + 63) / 64, (M + 63) / 64, splits);
  int kps = ((K + splits - 1) / splits + 15) / 16 * 16;
  float* dst = splits > 1 ? partial : C;
  GemmEpi e = splits > 1 ? GemmEpi{nullptr, nullptr, 0} : epi;
  cudaStream_t s = j->stream;
  if (a_kcontig && b_ncontig)
    sgemm_kernel<true, true><<<grid, 256, 0, s>>>(A, Bm, dst, M, N, K, sam, sak, sbk, sbn, kps, e);
  else if (a_kcontig && !b_ncontig)
    sgemm_kernel<true, false><<<grid, 256, 0, s>>>(A, Bm, dst, M, N, K, sam, sak, sbk, sbn, kps, e);
  else if (!a_kcontig && b_ncontig)
    sgemm_kernel<false, true><<<grid, 256, 0, s>>>(A, Bm, dst, M, N, K, sam, sak, sbk, sbn, kps, e);
  else
    sgemm_kernel<false, false><<<grid, 256, 0, s>>>(A, Bm, dst, M, N, K, sam, sak, sbk, sbn, kps, e);
  LAUNCH_OK(j, name);
  if (splits > 1) {
    const int MN = M * N;
    colsum_stage2_kernel<<<(MN + 255) / 256, 256, 0, s>>>(partial, C, splits, MN, MN);
    LAUNCH_OK(j, "split-K reduce");
  }
  return TFG_OK;
}

static int colsum(tfg_job* j, const float* in, const float* mask, float* out, int R, int C, float* partial) {
  const int col_blocks = (C + 63) / 64;
  int chunks = std::max(1, std::min((R + 31) / 32, 128));
  const int rows_per_chunk = (R + chunks - 1) / chunks;
  chunks = (R + rows_per_chunk - 1) / rows_per_chunk;
  cudaStream_t s = j->stream;
  colsum_wide_kernel<<<dim3(col_blocks, chunks), 256, 0, s>>>(in, mask, partial, R, C, rows_per_chunk);
  LAUNCH_OK(j, "colsum_wide_kernel");
  reduce_rows_kernel<<<(C + 31) / 32, 1024, 0, s>>>(partial, chunks, C, C, out, C, nullptr);
  LAUNCH_OK(j, "reduce_rows_kernel");
  return TFG_OK;
}

// ---------------------------------------------------------------------------------------
static const char* kVarNames[8] = {"W_conv1", "b_conv1", "W_conv2", "b_conv2",
                                   "W_fc1",   "b_fc1",   "W_fc2",   "b_fc2"};

int mnist_build(tfg_job* j, int batch, int math) {
  MnistModel* m = new MnistModel();
  m->B = batch;
  m->math = math;
  const int64_t shapes[8][4] = {{5, 5, 1, 32}, {32}, {5, 5, 32, 64}, {64},
                                {3136, 1024},  {1024}, {1024, 10}, {10}};
  const int ranks[8] = {4, 1, 4, 1, 2, 1, 2, 1};
  for (int v = 0; v < 8; ++v) {
    int rc = tfg_var_register(j, kVarNames[v], "VariableV2", TFG_DT_F32, shapes[v], ranks[v], 1,
                              &m->var[v], nullptr);
    if (rc) { delete m; return rc; }
  }
  int gs = 0;
  int rc = tfg_var_register(j, "global_step", "VariableV2", TFG_DT_I64, nullptr, 0, 0, &gs, nullptr);
  if (rc) { delete m; return rc; }
  const size_t B = (size_t)batch;
  const size_t nslots = j->local_replicas.size();
  m->slots.resize(nslots);
  auto alloc = [&](void** p, size_t bytes) -> cudaError_t { return cudaMalloc(p, bytes ? bytes : 16); };
  cudaError_t e = cudaSuccess;
  for (size_t s = 0; s < nslots && e == cudaSuccess; ++s) {
    MnistSlot& sl = m->slots[s];
#define A_(field, count, type) if (e == cudaSuccess) e = alloc((void**)&sl.field, sizeof(type) * (count))
    A_(x_stage, B * 784, float);
    A_(lab_stage, B, int32_t);
    A_(a1, B * 6272, float);
    A_(idx1, B * 6272, uint8_t);
    A_(col, math == TFG_MATH_FP32 ? B * 196 * 800 : 4, float);            // the im2col matrix: CUDA-core path only
    A_(y2, B * 196 * 64, float);
    A_(a2, B * 3136, float);
    A_(idx2, B * 3136, uint8_t);
    A_(h1, B * 1024, float);
    A_(logits, B * 10, float);
    A_(dlogits, B * 10, float);
    A_(loss_vec, B, float);
    A_(head_done, 1, unsigned int);
    if (e == cudaSuccess) e = cudaMemset(sl.head_done, 0, sizeof(unsigned int));
    A_(dh1, B * 1024, float);
    A_(da2, B * 3136, float);
    A_(da2m, B * 3136, float);
    A_(dy2, B * 196 * 64, float);
    A_(dcol, math == TFG_MATH_TF32X3 ? 4 : B * 196 * 800, float);         // 3xTF32 writes the input gradient directly
    A_(g1m, B * 6272, float);
    // scratch for deterministic reductions: split-K partials of dW2 (32 x 800 x 64), the
    // per-image conv1 partials (B x 832) and column-sum chunks
    size_t sc = std::max<size_t>({(size_t)kSplitW2 * 800 * 64, kConv1Split * B * 832, ((B + 15) / 16) * (size_t)(1024 * 11),
                                  ((B * 49 + 63) / 64) * 1024, (size_t)3136 * 1024});
    sl.scratch_floats = sc;
    A_(scratch, sc, float);
    A_(scratch_c1, (size_t)kConv1Split * B * 832, float);
#undef A_
  }
  if (e != cudaSuccess) {
    j->model = m;
    mnist_destroy(j);
    return fail_cuda(j, e, "cudaMalloc(model activations)");
  }
  j->model = m;
  if (math != TFG_MATH_FP32) {
    rc = mnist_tc_build(j);
    if (rc) { mnist_destroy(j); return rc; }
  }
  return TFG_OK;
}

void mnist_destroy(tfg_job* j) {
  MnistModel* m = j->model;
  if (!m) return;
  mnist_tc_destroy(j);
  for (auto& sl : m->slots) {
    void* ps[] = {sl.head_done, sl.x_stage, sl.lab_stage, sl.a1, sl.idx1, sl.col, sl.y2, sl.a2, sl.idx2, sl.h1,
                  sl.logits, sl.dlogits, sl.loss_vec, sl.dh1, sl.da2, sl.da2m, sl.dy2, sl.dcol,
                  sl.g1m, sl.scratch, sl.scratch_c1};
    for (void* p : ps) cudaFree(p);
  }
  delete m;
  j->model = nullptr;
}

int mnist_batch(const tfg_job* j) { return j->model ? j->model->B : 0; }
int mnist_math(const tfg_job* j) { return j->model ? j->model->math : TFG_MATH_FP32; }
float* mnist_stage_images(tfg_job* j, int slot) { return j->model->slots[slot].x_stage; }
int32_t* mnist_stage_labels(tfg_job* j, int slot) { return j->model->slots[slot].lab_stage; }

int mnist_logits(tfg_job* j, int slot, float* host, size_t bytes) {
  MnistModel* m = j->model;
  if (bytes != sizeof(float) * 10 * (size_t)m->B) return fail(j, TFG_EINVAL, "logits byte count mismatch");
  TFG_CUDA(j, cudaMemcpyAsync(host, m->slots[slot].logits, bytes, cudaMemcpyDeviceToHost, j->stream));
  TFG_CUDA(j, cudaStreamSynchronize(j->stream));
  return TFG_OK;
}

int mnist_reduce_rows(tfg_job* j, const float* in, int R, int C, float* out, cudaStream_t st) {
  reduce_rows_kernel<<<(C + 31) / 32, 1024, 0, st>>>(in, R, C, C, out, C, nullptr);
  LAUNCH_OK(j, "reduce_rows_kernel");
  return TFG_OK;
}

int mnist_pool2(tfg_job* j, int slot, void* a2_bf16, float* a2_lo) {
  MnistModel* m = j->model;
  MnistSlot& s = m->slots[slot];
  pool_relu_kernel<<<blocks_for((long long)m->B * 3136), 256, 0, j->stream>>>(
      s.y2, s.a2, s.idx2, static_cast<__nv_bfloat16*>(a2_bf16), a2_lo, m->B, 7, 64);
  LAUNCH_OK(j, "pool_relu_kernel");
  return TFG_OK;
}

int mnist_forward_backward(tfg_job* j, int slot, const float* x, const int32_t* labels) {
  MnistModel* m = j->model;
  MnistSlot& s = m->slots[slot];
  const int B = m->B;
  cudaStream_t st = j->stream;
  const float* P = j->params;
  float* G = j->grads[slot];
  auto par = [&](int v) { return P + j->vars[m->var[v]].offset; };
  auto grd = [&](int v) { return G + j->vars[m->var[v]].offset; };
  const int M2 = B * 196;
  int rc;
  const bool tc = m->math != TFG_MATH_FP32;

  // ---------------- forward ----------------
  {
  NvtxRange nvtx_fwd("mnist forward");
  const bool derive_lo = lp_derived_per_step(j);
  if (derive_lo) {   // low-order plane of the weights from this rank's fp32 copy, beside conv1 (which does not read it)
    if (launch_cast_lp(j->params, j->params_lp, j->flat_count, j->lp_mode, side_fork(j)))
      return fail(j, TFG_ECUDA, "split_lo4_kernel launch failed");
    prof_mark(j, "weights low-order plane");
  }
  conv1_pool_relu_kernel<<<B * kConv1Split, 256, 0, st>>>(x, par(0), par(1), s.a1, s.idx1, mnist_tc_a1_lo(j, slot));
  LAUNCH_OK(j, "conv1_pool_relu_kernel");
  if (derive_lo) side_join(j);
  if (tc) {
    if ((rc = mnist_tc_forward(j, slot))) return rc;   // conv2 + fc1 on the tensor path
  } else {
    im2col_kernel<float><<<blocks_for((long long)M2 * 200), 256, 0, st>>>(s.a1, s.col, nullptr, B);
    LAUNCH_OK(j, "im2col_kernel");
    // y2[M2,64] = col[M2,800] * W2[800,64] + b2
    if ((rc = sgemm(j, "gemm conv2 fwd", s.col, true, 800, 1, par(2), true, 64, 1, s.y2, M2, 64, 800,
                    GemmEpi{par(3), nullptr, 0}, 1, nullptr))) return rc;
    if ((rc = mnist_pool2(j, slot, nullptr, nullptr))) return rc;
    // h1[B,1024] = relu(a2[B,3136] * W3[3136,1024] + b3)
    if ((rc = sgemm(j, "gemm fc1 fwd", s.a2, true, 3136, 1, par(4), true, 1024, 1, s.h1, B, 1024, 3136,
                    GemmEpi{par(5), nullptr, 1}, 1, nullptr))) return rc;
  }
  // logits = h1 * W4 + b4, softmax cross-entropy and dlogits in one warp-shuffle kernel
  int fc1_splits = 1;
  const float* fc1_partial = tc ? mnist_tc_fc1_partial(j, slot, &fc1_splits) : nullptr;
  if (fc1_partial && fc1_splits > kMaxFusedSplits) return fail(j, TFG_ESTATE, "fc1 forward uses more K splits than head_fwd_kernel adds");
  head_fwd_kernel<<<B, 256, 0, st>>>(s.h1, fc1_partial, fc1_splits, par(5), par(6), par(7), labels, s.logits, s.dlogits,
                                     s.loss_vec, B, s.head_done, j->loss + j->local_replicas[slot]);
  LAUNCH_OK(j, "head_fwd_kernel");

  }
  // ---------------- backward ----------------
  NvtxRange nvtx_bwd("mnist backward");
  // dW4, db4, dh1 (masked by h1 > 0) and db3 in one pass over the batch
  {
    const int chunks = (B + kFc2Chunk - 1) / kFc2Chunk;
    fc2_bwd_kernel<<<dim3(1024 / 128, chunks), 128, 0, st>>>(s.h1, s.dlogits, par(6), s.scratch, s.dh1,
                                                             static_cast<__nv_bfloat16*>(mnist_tc_dh1_bf16(j, slot)),
                                                             mnist_tc_dh1_lo(j, slot), B);
    LAUNCH_OK(j, "fc2_bwd_kernel");
    // the gradients of W4, b3 and b4 are leaves of the step: side branch (joined before the function returns)
    fc2_bwd_reduce_kernel<<<(kFc2PartialStride + 10 + 255) / 256, 256, 0, side_fork(j)>>>(s.scratch, s.dlogits, grd(6), grd(5),
                                                                                         grd(7), chunks, B);
    LAUNCH_OK(j, "fc2_bwd_reduce_kernel");
  }
  if (tc) {
    if ((rc = mnist_tc_backward(j, slot))) return rc;  // dW3, da2, dcol, dW2 on the tensor path
  } else {
    // dW3[3136,1024] = a2^T * dh1
    if ((rc = sgemm(j, "gemm fc1 wgrad", s.a2, false, 1, 3136, s.dh1, true, 1024, 1, grd(4), 3136, 1024, B,
                    GemmEpi{nullptr, nullptr, 0}, 1, nullptr))) return rc;
    // da2[B,3136] = dh1 * W3^T
    if ((rc = sgemm(j, "gemm fc1 dgrad", s.dh1, true, 1024, 1, par(4), false, 1, 1024, s.da2, B, 3136, 1024,
                    GemmEpi{nullptr, nullptr, 0}, 1, nullptr))) return rc;
    unpool_kernel<float><<<blocks_for((long long)B * 3136), 256, 0, st>>>(
        s.da2, s.a2, s.idx2, s.dy2, nullptr, nullptr, B, 7, 64);
    LAUNCH_OK(j, "unpool_kernel");
    // dW2[800,64] = col^T * dy2, split-K over M2
    if ((rc = sgemm(j, "gemm conv2 wgrad", s.col, false, 1, 800, s.dy2, true, 64, 1, grd(2), 800, 64, M2,
                    GemmEpi{nullptr, nullptr, 0}, kSplitW2, s.scratch))) return rc;
    // dcol[M2,800] = dy2 * W2^T
    if ((rc = sgemm(j, "gemm conv2 dgrad", s.dy2, true, 64, 1, par(2), false, 1, 64, s.dcol, M2, 800, 64,
                    GemmEpi{nullptr, nullptr, 0}, 1, nullptr))) return rc;
  }
  // db2 = sum over pooled positions of the masked pooled gradient (3xTF32: taken by the un-pooling kernel)
  if (m->math != TFG_MATH_TF32X3 && (rc = colsum(j, s.da2, s.a2, grd(3), B * 49, 64, s.scratch))) return rc;
  if (m->math == TFG_MATH_TF32X3) {
    // the implicit data-gradient kernel wrote g1m itself (conv_x3.cuh)
  } else if (m->math == TFG_MATH_BF16) {   // the tensor path's dgrad GEMM stored dcol as bf16 (mnist_tc_prepare)
    col2im_mask_kernel<__nv_bfloat16><<<blocks_for((long long)B * 196 * 8), 256, 0, st>>>(
        reinterpret_cast<const __nv_bfloat16*>(s.dcol), s.a1, s.g1m, B);
    LAUNCH_OK(j, "col2im_mask_kernel");
  } else {
    col2im_mask_kernel<float><<<blocks_for((long long)B * 196 * 8), 256, 0, st>>>(s.dcol, s.a1, s.g1m, B);
    LAUNCH_OK(j, "col2im_mask_kernel");
  }
  // own scratch: s.scratch may still be read by the side branch (fc2_bwd_reduce_kernel)
  conv1_bwd_kernel<<<kConv1Split * B, 256, 0, st>>>(x, s.g1m, s.idx1, s.scratch_c1);
  LAUNCH_OK(j, "conv1_bwd_kernel");
  // [B][832] per-image partials -> dW1 (first 800 columns) and db1 (last 32), image order
  reduce_rows_kernel<<<(832 + 31) / 32, 1024, 0, st>>>(s.scratch_c1, kConv1Split * B, 832, 832, grd(0), 800, grd(1));
  LAUNCH_OK(j, "reduce_rows_kernel");
  side_join(j);
  return TFG_OK;
}

}  // namespace tfg
