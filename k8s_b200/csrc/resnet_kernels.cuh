// HBM-bound kernels around the contractions of the ResNet-50 step (BASELINE config 4):
// im2col / col2im for NHWC bf16 activations, batch-norm statistics and apply (forward and
// backward), 3x3/2 max-pool, global average pool, 1000-way softmax cross-entropy.
// Activations and their gradients are bf16; statistics, reductions and parameters fp32.
// Every reduction runs in a fixed order (two-stage, no floating-point atomics).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace tfg {
namespace rn {

typedef __nv_bfloat16 bf16;

struct ConvGeom {
  int N, H, W, C;        // input
  int R, S, stride, pad;
  int OH, OW;
  int Kp;                // R*S*C rounded up to a multiple of 8 (16-byte row pitch for TMA)
};

__device__ __forceinline__ float b2f(bf16 v) { return __bfloat162float(v); }
__device__ __forceinline__ bf16 f2b(float v) { return __float2bfloat16_rn(v); }

// 8 bf16 <-> 8 floats through one 16-byte access
__device__ __forceinline__ void ld8(const bf16* p, float (&v)[8]) {
  const uint4 u = *reinterpret_cast<const uint4*>(p);
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
  for (int k = 0; k < 4; ++k) { const float2 f = __bfloat1622float2(h[k]); v[2 * k] = f.x; v[2 * k + 1] = f.y; }
}
// the same in two steps, so that a kernel can issue several loads before it touches the first value
__device__ __forceinline__ uint4 ldraw(const bf16* p) { return __ldg(reinterpret_cast<const uint4*>(p)); }
__device__ __forceinline__ void unpack8(const uint4& u, float (&v)[8]) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
  for (int k = 0; k < 4; ++k) { const float2 f = __bfloat1622float2(h[k]); v[2 * k] = f.x; v[2 * k + 1] = f.y; }
}
__device__ __forceinline__ void st8(bf16* p, const float (&v)[8]) {
  uint4 u;
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
  for (int k = 0; k < 4; ++k) h[k] = __floats2bfloat162_rn(v[2 * k], v[2 * k + 1]);
  *reinterpret_cast<uint4*>(p) = u;
}
__device__ __forceinline__ void ldf8(const float* p, float (&v)[8]) {
  const float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}

// images fp32 NHWC -> bf16
__global__ void __launch_bounds__(256) to_bf16_kernel(const float* __restrict__ src, bf16* __restrict__ dst, long long n) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    dst[i] = f2b(src[i]);
}

// ---- im2col: in[N,H,W,C] -> col[N*OH*OW, Kp], k = (r*S + s)*C + c ---------------------------
// VEC = 8 channels (16 B) per thread when C % 8 == 0, else 1.
template <int VEC>
__global__ void __launch_bounds__(256)
im2col_kernel(const bf16* __restrict__ in, bf16* __restrict__ col, ConvGeom g) {
  const int cv = g.C / VEC;
  const long long M = (long long)g.N * g.OH * g.OW;
  const long long total = M * g.R * g.S * cv;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % cv) * VEC;
    long long t = i / cv;
    const int s = (int)(t % g.S); t /= g.S;
    const int r = (int)(t % g.R); t /= g.R;
    const long long m = t;
    const int ow = (int)(m % g.OW), oh = (int)((m / g.OW) % g.OH);
    const long long n = m / ((long long)g.OW * g.OH);
    const int ih = oh * g.stride + r - g.pad, iw = ow * g.stride + s - g.pad;
    const bool ok = ih >= 0 && ih < g.H && iw >= 0 && iw < g.W;
    bf16* dst = col + m * g.Kp + (r * g.S + s) * g.C + c;
    const bf16* src = in + ((n * g.H + ih) * g.W + iw) * g.C + c;
    if (VEC == 8) {
      *reinterpret_cast<uint4*>(dst) = ok ? *reinterpret_cast<const uint4*>(src) : make_uint4(0, 0, 0, 0);
    } else {
      dst[0] = ok ? src[0] : f2b(0.f);
    }
  }
}

// Thin-channel variant (the 7x7x3 stem): a thread builds 8 consecutive elements (one 16-byte
// store) of a col row, consecutive threads consecutive pieces, so the 1 GB of stores is fully
// coalesced (pad columns [R*S*C, Kp) included).  Within one filter row the S*C source elements
// are contiguous in the image, so the 2-byte gathered reads stay in a few L1 lines.
__global__ void __launch_bounds__(256)
im2col_rows_kernel(const bf16* __restrict__ in, bf16* __restrict__ col, ConvGeom g) {
  const long long M = (long long)g.N * g.OH * g.OW;
  const int kv = g.Kp >> 3;
  const long long total = M * kv;
  const int K = g.R * g.S * g.C, SC = g.S * g.C;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int piece = (int)(i % kv);
    const long long m = i / kv;
    const int ow = (int)(m % g.OW), oh = (int)((m / g.OW) % g.OH);
    const long long n = m / ((long long)g.OW * g.OH);
    const int ih0 = oh * g.stride - g.pad, iw0 = ow * g.stride - g.pad;
    const int lo = max(0, -iw0 * g.C), hi = min(SC, (g.W - iw0) * g.C);   // in-image part of a filter row
    const bf16* base = in + ((n * g.H + ih0) * g.W + iw0) * g.C;
    int k = piece << 3;
    int r = k / SC, sc = k - r * SC;
    uint4 pk;
    unsigned short* h = reinterpret_cast<unsigned short*>(&pk);
#pragma unroll
    for (int u = 0; u < 8; ++u, ++k) {
      unsigned short x = 0;
      const int ih = ih0 + r;
      if (k < K && ih >= 0 && ih < g.H && sc >= lo && sc < hi)
        x = __ldg(reinterpret_cast<const unsigned short*>(base) + (long long)r * g.W * g.C + sc);
      h[u] = x;
      if (++sc == SC) { sc = 0; ++r; }
    }
    *reinterpret_cast<uint4*>(col + m * g.Kp + (piece << 3)) = pk;
  }
}

// zero the pad columns [R*S*C, Kp) of col (only the stem has any)
__global__ void __launch_bounds__(256) zero_pad_cols_kernel(bf16* __restrict__ col, long long M, int K, int Kp) {
  const int padw = Kp - K;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < M * padw; i += (long long)gridDim.x * blockDim.x)
    col[(i / padw) * Kp + K + (i % padw)] = f2b(0.f);
}

// ---- col2im (gather form): din[n,h,w,c] (+)= sum over taps of dcol[(n,oh,ow), (r*S+s)*C + c] ---
template <int VEC>
__global__ void __launch_bounds__(256)
col2im_kernel(const bf16* __restrict__ dcol, bf16* __restrict__ din, ConvGeom g, int accumulate) {
  const int cv = g.C / VEC;
  const long long total = (long long)g.N * g.H * g.W * cv;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % cv) * VEC;
    const long long pix = i / cv;
    const int w = (int)(pix % g.W), h = (int)((pix / g.W) % g.H);
    const long long n = pix / ((long long)g.W * g.H);
    float acc[VEC];
#pragma unroll
    for (int u = 0; u < VEC; ++u) acc[u] = 0.f;
    for (int r = 0; r < g.R; ++r) {
      const int th = h + g.pad - r;
      if (th < 0 || th % g.stride) continue;
      const int oh = th / g.stride;
      if (oh >= g.OH) continue;
      for (int s = 0; s < g.S; ++s) {
        const int tw = w + g.pad - s;
        if (tw < 0 || tw % g.stride) continue;
        const int ow = tw / g.stride;
        if (ow >= g.OW) continue;
        const bf16* src = dcol + ((n * g.OH + oh) * g.OW + ow) * g.Kp + (r * g.S + s) * g.C + c;
        if (VEC == 8) {
          const uint4 v = *reinterpret_cast<const uint4*>(src);
          const bf16* e = reinterpret_cast<const bf16*>(&v);
#pragma unroll
          for (int u = 0; u < VEC; ++u) acc[u] += b2f(e[u]);
        } else {
          acc[0] += b2f(src[0]);
        }
      }
    }
    bf16* dst = din + pix * g.C + c;
    if (VEC == 8) {
      float o[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) o[u] = acc[u % VEC];
      if (accumulate) {
        float prev[8];
        ld8(dst, prev);
#pragma unroll
        for (int u = 0; u < 8; ++u) o[u] += prev[u];
      }
      st8(dst, o);
    } else {
      if (accumulate) acc[0] += b2f(dst[0]);
      dst[0] = f2b(acc[0]);
    }
  }
}

// dst (+)= src, bf16
__global__ void __launch_bounds__(256)
add_bf16_kernel(const bf16* __restrict__ src, bf16* __restrict__ dst, long long n, int accumulate) {
  const long long n8 = n >> 3;   // every activation tensor of the net has a multiple of 8 elements
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n8; i += (long long)gridDim.x * blockDim.x) {
    float a[8], b[8];
    ld8(src + (i << 3), a);
    if (accumulate) {
      ld8(dst + (i << 3), b);
#pragma unroll
      for (int k = 0; k < 8; ++k) a[k] += b[k];
    }
    st8(dst + (i << 3), a);
  }
}

// ---- batch norm ----------------------------------------------------------------------------
// Per-channel sums over rows run in two stages inside ONE launch: every block stores the sums
// of its row chunk to partial[chunk][2][C]; the block that finishes last for a group of 64
// channels (a counter in `done`, which it resets) folds the chunks in a fixed order and writes
// the statistics, so the result does not depend on which block came last and no float atomics
// are involved.  block = 8 channel groups (8 channels = 16 B each) x 32 row lanes; C is a
// multiple of 64 for every layer of the net.
//
// These kernels move 2-byte elements at HBM speed, which leaves about ten instructions per
// element and stream on an SM; hence the relu mask of the forward pass is kept as one bit per
// element (written by bn_apply_kernel) instead of being re-derived from y or z in the two
// backward kernels, and per-channel constants live in registers (a thread keeps its channels).
constexpr int kBnRowsInFlight = 4;

// returns true in the block that stored the last chunk of channel group blockIdx.x; *fold_s and
// *fold_q then hold the folded sums of channel blockIdx.x * 64 + threadIdx.x in threads 0..63
__device__ __forceinline__ bool bn_block_reduce(float (&s)[8], float (&q)[8], float* partial, unsigned int* done,
                                                int chunks, int C, int cg, int rl, float* fold_s, float* fold_q) {
  __shared__ float r1[32][65], r2[32][65];
  __shared__ bool last;
#pragma unroll
  for (int k = 0; k < 8; ++k) { r1[rl][cg * 8 + k] = s[k]; r2[rl][cg * 8 + k] = q[k]; }
  __syncthreads();
  const int t = threadIdx.x;
  if (t < 64) {
    float a = 0.f, b = 0.f;
#pragma unroll 8
    for (int l = 0; l < 32; ++l) { a += r1[l][t]; b += r2[l][t]; }   // fixed order
    const int c = blockIdx.x * 64 + t;
    __stcg(&partial[((size_t)blockIdx.y * 2 + 0) * C + c], a);
    __stcg(&partial[((size_t)blockIdx.y * 2 + 1) * C + c], b);
    __threadfence();
  }
  __syncthreads();
  if (t == 0) {
    const unsigned int prev = atomicAdd(&done[blockIdx.x], 1u);
    last = prev == (unsigned int)chunks - 1u;
    if (last) done[blockIdx.x] = 0u;   // ready for the next launch on this stream
  }
  __syncthreads();
  if (!last) return false;
  __threadfence();
  // fold: 16 strands of interleaved chunks x 16 channel quads (many loads in flight: the fold is
  // the tail of every launch), then the strands in order 0..15
  const int quad = t & 15, strand = t >> 4;
  const float* base = partial + blockIdx.x * 64 + quad * 4;
  float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
#pragma unroll 8
  for (int k = strand; k < chunks; k += 16) {
    const float4 x = __ldcg(reinterpret_cast<const float4*>(base + ((size_t)k * 2 + 0) * C));
    const float4 y = __ldcg(reinterpret_cast<const float4*>(base + ((size_t)k * 2 + 1) * C));
    a.x += x.x; a.y += x.y; a.z += x.z; a.w += x.w;
    b.x += y.x; b.y += y.y; b.z += y.z; b.w += y.w;
  }
  __syncthreads();                    // r1 / r2 are reused below
  r1[strand][quad * 4 + 0] = a.x; r1[strand][quad * 4 + 1] = a.y; r1[strand][quad * 4 + 2] = a.z; r1[strand][quad * 4 + 3] = a.w;
  r2[strand][quad * 4 + 0] = b.x; r2[strand][quad * 4 + 1] = b.y; r2[strand][quad * 4 + 2] = b.z; r2[strand][quad * 4 + 3] = b.w;
  __syncthreads();
  if (t < 64) {
    float fs = 0.f, fq = 0.f;
#pragma unroll
    for (int l = 0; l < 16; ++l) { fs += r1[l][t]; fq += r2[l][t]; }
    *fold_s = fs;
    *fold_q = fq;
  }
  return true;
}

// mean, invstd (biased variance, eps) and the moving statistics update from the folded sums
// (threads 0..63 of the block that folded channel group blockIdx.x)
__device__ __forceinline__ void bn_write_stats(float fs, float fq, long long M, float eps, float decay,
                                               float* __restrict__ mean, float* __restrict__ invstd,
                                               float* __restrict__ mov_mean, float* __restrict__ mov_var) {
  if (threadIdx.x >= 64) return;
  const int c = blockIdx.x * 64 + threadIdx.x;
  const float mu = fs / (float)M;
  const float var = fmaxf(fq / (float)M - mu * mu, 0.f);
  mean[c] = mu;
  invstd[c] = rsqrtf(var + eps);
  if (mov_mean) {
    mov_mean[c] = decay * mov_mean[c] + (1.f - decay) * mu;
    mov_var[c] = decay * mov_var[c] + (1.f - decay) * var * ((float)M / (float)max(M - 1, 1ll));
  }
}

// forward statistics from z itself
__global__ void __launch_bounds__(256)
bn_stats_kernel(const bf16* __restrict__ z, float* partial, unsigned int* done, long long M, int C, int rows_per_chunk,
                float eps, float decay, float* __restrict__ mean, float* __restrict__ invstd,
                float* __restrict__ mov_mean, float* __restrict__ mov_var) {
  const int cg = threadIdx.x & 7, rl = threadIdx.x >> 3;
  const int c0 = blockIdx.x * 64 + cg * 8;
  const long long r0 = (long long)blockIdx.y * rows_per_chunk, rend = min(M, r0 + rows_per_chunk);
  float s[8], q[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) s[k] = q[k] = 0.f;
  for (long long r = r0 + rl; r < rend; r += 32 * kBnRowsInFlight) {
    uint4 raw[kBnRowsInFlight];   // all loads first (rows past the chunk re-read row r and are not added)
#pragma unroll
    for (int u = 0; u < kBnRowsInFlight; ++u) {
      const long long rr = r + 32 * u;
      raw[u] = ldraw(z + (rr < rend ? rr : r) * C + c0);
    }
#pragma unroll
    for (int u = 0; u < kBnRowsInFlight; ++u) {
      if (r + 32 * u < rend) {
        float v[8];
        unpack8(raw[u], v);
#pragma unroll
        for (int k = 0; k < 8; ++k) { s[k] += v[k]; q[k] = fmaf(v[k], v[k], q[k]); }
      }
    }
  }
  float fs = 0.f, fq = 0.f;
  if (!bn_block_reduce(s, q, partial, done, gridDim.y, C, cg, rl, &fs, &fq)) return;
  bn_write_stats(fs, fq, M, eps, decay, mean, invstd, mov_mean, mov_var);
}

// The same statistics from the column sums that the convolution's GEMM epilogue left in
// stats[record][2][C] (sum, sum of squares over the rows one CTA slice saw): z is not read again.
// At most a few hundred records, so one block of 1024 threads per 64 channels folds them in a
// single pass: 16 channel quads x 64 strands of interleaved records, then the strands in order.
__global__ void __launch_bounds__(1024)
bn_fold_kernel(const float* __restrict__ stats, int records, long long M, int C, float eps, float decay,
               float* __restrict__ mean, float* __restrict__ invstd, float* __restrict__ mov_mean,
               float* __restrict__ mov_var) {
  __shared__ float f1[64][65], f2[64][65];
  const int t = threadIdx.x, quad = t & 15, strand = t >> 4;
  const float* base = stats + blockIdx.x * 64 + quad * 4;
  float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
#pragma unroll 4
  for (int r = strand; r < records; r += 64) {
    const float4 x = __ldg(reinterpret_cast<const float4*>(base + (size_t)r * 2 * C));
    const float4 y = __ldg(reinterpret_cast<const float4*>(base + (size_t)r * 2 * C + C));
    a.x += x.x; a.y += x.y; a.z += x.z; a.w += x.w;
    b.x += y.x; b.y += y.y; b.z += y.z; b.w += y.w;
  }
  f1[strand][quad * 4 + 0] = a.x; f1[strand][quad * 4 + 1] = a.y; f1[strand][quad * 4 + 2] = a.z; f1[strand][quad * 4 + 3] = a.w;
  f2[strand][quad * 4 + 0] = b.x; f2[strand][quad * 4 + 1] = b.y; f2[strand][quad * 4 + 2] = b.z; f2[strand][quad * 4 + 3] = b.w;
  __syncthreads();
  if (t < 64) {
    float fs = 0.f, fq = 0.f;
#pragma unroll 8
    for (int l = 0; l < 64; ++l) { fs += f1[l][t]; fq += f2[l][t]; }
    bn_write_stats(fs, fq, M, eps, decay, mean, invstd, mov_mean, mov_var);
  }
}

// y = gamma * (z - mean) * invstd + beta (+ residual) (relu), bf16 in / out.  8 channels per
// thread; the grid stride is a multiple of C (C divides 2048 = 8 * blockDim), so a thread keeps
// its channels and loads the per-channel values once.  With relu, bit k of relu_bits[i] says
// whether element 8 i + k passed (value > 0): the mask the backward kernels apply.
__global__ void __launch_bounds__(256, 3)
bn_apply_kernel(const bf16* __restrict__ z, const float* __restrict__ mean, const float* __restrict__ invstd,
                const float* __restrict__ gamma, const float* __restrict__ beta, const bf16* __restrict__ residual,
                bf16* __restrict__ y, uint8_t* __restrict__ relu_bits, long long total, int C) {
  const long long total8 = total >> 3;
  const long long stride = (long long)gridDim.x * blockDim.x;
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const int c = (int)((i << 3) % C);
  float mu[8], sc[8], be[8];
  {
    float is[8], ga[8];
    ldf8(mean + c, mu); ldf8(invstd + c, is); ldf8(gamma + c, ga); ldf8(beta + c, be);
#pragma unroll
    for (int k = 0; k < 8; ++k) sc[k] = ga[k] * is[k];
  }
  for (; i < total8; i += 2 * stride) {
    const long long i2 = i + stride < total8 ? i + stride : i;   // second element (or the first again: same values rewritten)
    const uint4 z0 = ldraw(z + (i << 3)), z1 = ldraw(z + (i2 << 3));
    uint4 q0 = make_uint4(0, 0, 0, 0), q1 = q0;
    if (residual) { q0 = ldraw(residual + (i << 3)); q1 = ldraw(residual + (i2 << 3)); }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      float v[8], rs[8];
      unpack8(u ? z1 : z0, v);
      unpack8(u ? q1 : q0, rs);
      unsigned int bits = 0;
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        float o = fmaf(v[k] - mu[k], sc[k], be[k]);
        if (residual) o += rs[k];
        if (relu_bits) {
          bits |= (o > 0.f ? 1u : 0u) << k;
          o = fmaxf(o, 0.f);
        }
        v[k] = o;
      }
      const long long io = u ? i2 : i;
      st8(y + (io << 3), v);
      if (relu_bits) relu_bits[io] = (uint8_t)bits;
    }
  }
}

// backward statistics: dbeta = sum g, dgamma = sum g * xhat, g = dy under the relu mask.
// Same linear walk as bn_bwd_apply_kernel (a thread keeps its 8 channels, consecutive threads
// consecutive 16-byte pieces: whole rows, not 128-byte strips), two pieces in flight; the block
// then adds the threads that share channels (fixed order) and leaves one record
// stats[blockIdx.x][2][C] = (sum g, sum g * (z - mean)) for bn_bwd_fold_kernel.
template <bool RELU>
__global__ void __launch_bounds__(256, 3)
bn_bwd_stats_kernel(const bf16* __restrict__ dy, const uint8_t* __restrict__ relu_bits, const bf16* __restrict__ z,
                    const float* __restrict__ mean, float* __restrict__ stats, long long total, int C) {
  __shared__ float red[2][8][257];
  const long long total8 = total >> 3;
  const long long stride = (long long)gridDim.x * blockDim.x;
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const int c = (int)((i << 3) % C);
  float s[8], q[8], mu[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) s[k] = q[k] = 0.f;
  ldf8(mean + c, mu);
  for (; i < total8; i += 2 * stride) {
    const long long i2 = i + stride;
    const bool two = i2 < total8;
    const long long j2 = two ? i2 : i;
    const uint4 g0 = ldraw(dy + (i << 3)), g1 = ldraw(dy + (j2 << 3));
    const uint4 z0 = ldraw(z + (i << 3)), z1 = ldraw(z + (j2 << 3));
    unsigned int b0 = 0xffu, b1 = 0xffu;
    if (RELU) { b0 = __ldg(relu_bits + i); b1 = __ldg(relu_bits + j2); }
    if (!two) b1 = 0u;                       // the tail piece counts once
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      float g[8], zz[8];
      unpack8(u ? g1 : g0, g);
      unpack8(u ? z1 : z0, zz);
      const unsigned int bits = u ? b1 : b0;
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const float gg = ((bits >> k) & 1u) ? g[k] : 0.f;
        s[k] += gg;
        q[k] = fmaf(gg, zz[k] - mu[k], q[k]);
      }
    }
  }
  // threads t and t' hold the same channels iff t = t' (mod C / 8)
  const int t = threadIdx.x;
#pragma unroll
  for (int k = 0; k < 8; ++k) { red[0][k][t] = s[k]; red[1][k][t] = q[k]; }
  __syncthreads();
  const int groups = min(C >> 3, 256);        // distinct channel groups in the block
  for (int o = t; o < groups * 8; o += 256) {
    const int grp = o >> 3, k = o & 7;
    float a = 0.f, b = 0.f;
    for (int tt = grp; tt < 256; tt += groups) { a += red[0][k][tt]; b += red[1][k][tt]; }
    // channel of (thread grp of this block, element k): the block's first piece starts at channel
    // (blockIdx.x * 2048) % C = 0, since C divides 2048
    const int ch = grp * 8 + k;
    stats[((size_t)blockIdx.x * 2 + 0) * C + ch] = a;
    stats[((size_t)blockIdx.x * 2 + 1) * C + ch] = b;
  }
}

// dbeta = sum of the records' first halves, dgamma = invstd * sum of the second halves; one
// block of 1024 threads per 64 channels (see bn_fold_kernel)
__global__ void __launch_bounds__(1024)
bn_bwd_fold_kernel(const float* __restrict__ stats, int records, int C, const float* __restrict__ invstd,
                   float* __restrict__ dbeta, float* __restrict__ dgamma) {
  __shared__ float f1[64][65], f2[64][65];
  const int t = threadIdx.x, quad = t & 15, strand = t >> 4;
  const float* base = stats + blockIdx.x * 64 + quad * 4;
  float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
#pragma unroll 4
  for (int r = strand; r < records; r += 64) {
    const float4 x = __ldg(reinterpret_cast<const float4*>(base + (size_t)r * 2 * C));
    const float4 y = __ldg(reinterpret_cast<const float4*>(base + (size_t)r * 2 * C + C));
    a.x += x.x; a.y += x.y; a.z += x.z; a.w += x.w;
    b.x += y.x; b.y += y.y; b.z += y.z; b.w += y.w;
  }
  f1[strand][quad * 4 + 0] = a.x; f1[strand][quad * 4 + 1] = a.y; f1[strand][quad * 4 + 2] = a.z; f1[strand][quad * 4 + 3] = a.w;
  f2[strand][quad * 4 + 0] = b.x; f2[strand][quad * 4 + 1] = b.y; f2[strand][quad * 4 + 2] = b.z; f2[strand][quad * 4 + 3] = b.w;
  __syncthreads();
  if (t < 64) {
    float fs = 0.f, fq = 0.f;
#pragma unroll 8
    for (int l = 0; l < 64; ++l) { fs += f1[l][t]; fq += f2[l][t]; }
    const int ch = blockIdx.x * 64 + t;
    dbeta[ch] = fs;
    dgamma[ch] = fq * invstd[ch];
  }
}

// dz = gamma * invstd * (g - dbeta/M - xhat * dgamma/M); optionally also writes the masked g
// (the gradient that flows into the residual branch).  8 channels per thread, per-channel
// values loaded once (see bn_apply_kernel) and folded: dz = k1 * (g - db) - k2 * (z - mu) with
// k1 = gamma * invstd, db = dbeta / M, k2 = k1 * invstd * dgamma / M.
template <bool RELU>
__global__ void __launch_bounds__(256, 3)
bn_bwd_apply_kernel(const bf16* __restrict__ dy, const uint8_t* __restrict__ relu_bits, const bf16* __restrict__ z,
                    const float* __restrict__ mean, const float* __restrict__ invstd, const float* __restrict__ gamma,
                    const float* __restrict__ dbeta, const float* __restrict__ dgamma, bf16* __restrict__ dz,
                    bf16* __restrict__ g_out, int g_accumulate, long long total, int C, long long M) {
  const float inv_m = 1.f / (float)M;
  const long long total8 = total >> 3;
  const long long stride = (long long)gridDim.x * blockDim.x;
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const int c = (int)((i << 3) % C);
  float mu[8], db[8], k1[8], k2[8];
  {
    float is[8], ga[8];
    ldf8(mean + c, mu); ldf8(invstd + c, is); ldf8(gamma + c, ga);
    ldf8(dbeta + c, db); ldf8(dgamma + c, k2);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      k1[k] = ga[k] * is[k];
      db[k] *= inv_m;
      k2[k] = k1[k] * is[k] * (k2[k] * inv_m);
    }
  }
  for (; i < total8; i += 2 * stride) {
    const long long i2 = i + stride < total8 ? i + stride : i;
    const bool acc = g_out && g_accumulate;
    const uint4 g0 = ldraw(dy + (i << 3)), g1 = ldraw(dy + (i2 << 3));
    const uint4 z0 = ldraw(z + (i << 3)), z1 = ldraw(z + (i2 << 3));
    unsigned int b0 = 0xffu, b1 = 0xffu;
    if (RELU) { b0 = __ldg(relu_bits + i); b1 = __ldg(relu_bits + i2); }
    uint4 p0 = make_uint4(0, 0, 0, 0), p1 = p0;
    if (acc) {   // not __ldg: g_out is written below
      p0 = *reinterpret_cast<const uint4*>(g_out + (i << 3));
      p1 = *reinterpret_cast<const uint4*>(g_out + (i2 << 3));
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      if (u == 1 && i2 == i) break;   // odd tail: the accumulate must not run twice
      float g[8], zz[8], o[8], prev[8];
      unpack8(u ? g1 : g0, g);
      unpack8(u ? z1 : z0, zz);
      const unsigned int bits = u ? b1 : b0;
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        if (RELU && !((bits >> k) & 1u)) g[k] = 0.f;
        o[k] = k1[k] * (g[k] - db[k]) - k2[k] * (zz[k] - mu[k]);
      }
      const long long io = u ? i2 : i;
      st8(dz + (io << 3), o);
      if (g_out) {
        if (acc) {
          unpack8(u ? p1 : p0, prev);
#pragma unroll
          for (int k = 0; k < 8; ++k) g[k] += prev[k];
        }
        st8(g_out + (io << 3), g);
      }
    }
  }
}

// ---- 3x3 stride-2 pad-1 max pool -------------------------------------------------------------
__global__ void __launch_bounds__(256)
maxpool_fwd_kernel(const bf16* __restrict__ x, bf16* __restrict__ y, uint8_t* __restrict__ arg, int N, int H, int W,
                   int C, int OH, int OW) {
  const int cv = C >> 3;   // 8 channels per thread
  const long long total = (long long)N * OH * OW * cv;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % cv) << 3;
    const int ow = (int)((i / cv) % OW), oh = (int)((i / cv / OW) % OH);
    const long long n = i / cv / OW / OH;
    float best[8], v[8];
    int a[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) { best[k] = -INFINITY; a[k] = 0; }
    for (int r = 0; r < 3; ++r)
      for (int s = 0; s < 3; ++s) {
        const int ih = oh * 2 + r - 1, iw = ow * 2 + s - 1;
        if (ih < 0 || ih >= H || iw < 0 || iw >= W) continue;
        ld8(x + ((n * H + ih) * W + iw) * C + c, v);
#pragma unroll
        for (int k = 0; k < 8; ++k)
          if (v[k] > best[k]) { best[k] = v[k]; a[k] = r * 3 + s; }
      }
    const long long o = (((n * OH + oh) * OW + ow) * C) + c;
    st8(y + o, best);
    uint2 pk;
    pk.x = (uint32_t)a[0] | ((uint32_t)a[1] << 8) | ((uint32_t)a[2] << 16) | ((uint32_t)a[3] << 24);
    pk.y = (uint32_t)a[4] | ((uint32_t)a[5] << 8) | ((uint32_t)a[6] << 16) | ((uint32_t)a[7] << 24);
    *reinterpret_cast<uint2*>(arg + o) = pk;
  }
}

// gather form: dx[n,h,w,c] = sum of dy over the (up to 4) windows whose arg-max is (h,w)
__global__ void __launch_bounds__(256)
maxpool_bwd_kernel(const bf16* __restrict__ dy, const uint8_t* __restrict__ arg, bf16* __restrict__ dx, int N, int H,
                   int W, int C, int OH, int OW) {
  const int cv = C >> 3;
  const long long total = (long long)N * H * W * cv;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % cv) << 3;
    const int w = (int)((i / cv) % W), h = (int)((i / cv / W) % H);
    const long long n = i / cv / W / H;
    float acc[8], g[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] = 0.f;
    for (int r = 0; r < 3; ++r) {
      const int th = h + 1 - r;
      if (th < 0 || (th & 1)) continue;
      const int oh = th >> 1;
      if (oh >= OH) continue;
      for (int s = 0; s < 3; ++s) {
        const int tw = w + 1 - s;
        if (tw < 0 || (tw & 1)) continue;
        const int ow = tw >> 1;
        if (ow >= OW) continue;
        const long long o = ((n * OH + oh) * OW + ow) * C + c;
        const uint2 pk = *reinterpret_cast<const uint2*>(arg + o);
        ld8(dy + o, g);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const uint32_t a = ((k < 4 ? pk.x : pk.y) >> (8 * (k & 3))) & 0xffu;
          if (a == (uint32_t)(r * 3 + s)) acc[k] += g[k];
        }
      }
    }
    st8(dx + (((n * H + h) * W + w) * C) + c, acc);
  }
}

// ---- global average pool ---------------------------------------------------------------------
__global__ void __launch_bounds__(256)
avgpool_fwd_kernel(const bf16* __restrict__ x, bf16* __restrict__ p, int N, int HW, int C) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N * C) return;
  const int c = i % C, n = i / C;
  float s = 0.f;
  for (int k = 0; k < HW; ++k) s += b2f(x[((long long)n * HW + k) * C + c]);
  p[i] = f2b(s / (float)HW);
}

__global__ void __launch_bounds__(256)
avgpool_bwd_kernel(const bf16* __restrict__ dp, bf16* __restrict__ dx, int N, int HW, int C) {
  const long long total = (long long)N * HW * C;
  const float inv = 1.f / (float)HW;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    const long long n = i / ((long long)HW * C);
    dx[i] = f2b(b2f(dp[n * C + c]) * inv);
  }
}

// ---- softmax cross-entropy over `classes` logits, one warp per example ---------------------------
__global__ void __launch_bounds__(256)
softmax_xent_wide_kernel(const float* __restrict__ logits, const int32_t* __restrict__ labels,
                         bf16* __restrict__ dlogits, float* __restrict__ loss_vec, int B, int classes) {
  const int lane = threadIdx.x & 31;
  const int ex = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (ex >= B) return;
  const float* z = logits + (size_t)ex * classes;
  float mx = -INFINITY;
  for (int k = lane; k < classes; k += 32) mx = fmaxf(mx, z[k]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  float sum = 0.f;
  for (int k = lane; k < classes; k += 32) sum += expf(z[k] - mx);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  const int lab = labels[ex];
  const float inv_b = 1.f / (float)B;
  for (int k = lane; k < classes; k += 32)
    dlogits[(size_t)ex * classes + k] = f2b((expf(z[k] - mx) / sum - (k == lab ? 1.f : 0.f)) * inv_b);
  if (lane == 0) loss_vec[ex] = logf(sum) - (z[lab] - mx);
}

// column sums of a bf16 matrix [R, C] -> fp32 out[C] (R small: the FC bias gradient), fixed order
__global__ void __launch_bounds__(256)
colsum_bf16_kernel(const bf16* __restrict__ in, float* __restrict__ out, int R, int C) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  float s = 0.f;
  for (int r = 0; r < R; ++r) s += b2f(in[(size_t)r * C + c]);
  out[c] = s;
}

// mean of loss_vec[B] in a fixed order -> *out
__global__ void __launch_bounds__(256) mean_kernel(const float* __restrict__ v, float* out, int B) {
  __shared__ float part[256];
  float s = 0.f;
  for (int i = threadIdx.x; i < B; i += 256) s += v[i];
  part[threadIdx.x] = s;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o) part[threadIdx.x] += part[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) *out = part[0] / (float)B;
}

}  // namespace rn
}  // namespace tfg
