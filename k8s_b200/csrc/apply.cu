// Dense PS update: gradient push (peer loads) + mean + optimizer apply + parameter pull
// (peer stores) in one kernel, and the cross-rank flag barrier.  SURVEY.md §8a A12-A14.
//
// HBM-bound.  Algorithmic bytes per parameter on the owner: 4*n_grads read (remote over
// NVLink when n_gpus > 1) + 4 read var (+4 accum) + 4 written per pull destination
// (+4 accum).  Single GPU, SGD: 12 B/param; momentum: 20 B/param.
#include <algorithm>
#include <type_traits>

#include <cuda_bf16.h>

#include "common.cuh"
#include "gemm_tc.cuh"

namespace tfg {

constexpr int kApplyThreads = 256;
constexpr int kChunkBig = 4096;         // elements per work chunk: 4 float4 per thread (large shards: bytes in flight)
constexpr int kChunkSmall = 1024;       // 1 float4 per thread (small shards: enough CTAs to cover the chip)
// replicas whose gradients are requested before the first add.  Large shards (VEC = 4) already keep
// four 16-byte loads per thread and four CTAs per SM in flight -- 10 MB chip-wide, several times
// what NVLink's latency x bandwidth needs -- and a second group costs a CTA of occupancy (measured:
// 0.97 -> 0.88 of HBM peak on one GPU); small shards (VEC = 1) take eight.
template <int VEC> constexpr int load_group() { return VEC == 4 ? 1 : 8; }

struct ApplyArgs {
  const float* grads[TFG_MAX_REPLICAS];
  float* pull[TFG_MAX_GPUS];
  void* lp[TFG_MAX_GPUS];            // operand-precision shadows of the parameter buffers (own + peers), n_lp of them
  int lp_mode;                       // 1: bf16 copy, 2: fp32 low-order plane of the 3xTF32 split (gemm_tc.cuh:tf32_lo)
  const float* loss_src[TFG_MAX_REPLICAS];
  float* var;
  float* accum;
  float* loss_dst;
  const Segment* segs;   // NULL: the single inline segment `seg0`
  Segment seg0;
  int n_segs;
  const int* err;        // the job's error word: set by a barrier that timed out; the update is then skipped
  int n_grads, n_pull, n_lp, n_loss, kind;
  long long n_chunks;
  float divisor, lr, mu;
  // Cross-rank synchronisation folded into the kernel (sync != 0, n_ranks > 1), see push_apply_pull_kernel
  int sync, me, n_ranks;
  uint32_t* flags[TFG_MAX_GPUS];     // every rank's flag block (own one local, others peer-mapped): kFlagsReady / kFlagsPulled rows
  uint32_t* epoch;                   // device-resident step counter of this rank: a captured launch replays correctly
  unsigned int* done;                // CTAs of this launch that have finished their stores
  int* err_out;                      // error word to set when a peer does not show up
};

__device__ __forceinline__ float4 ld_stream(const float* p) {
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
  return r;
}

// One optimizer update, written with explicit round-to-nearest ops so that nvcc cannot
// contract mul+add into FMA: TF 1.3's Eigen expressions round every product.
template <int KIND>
__device__ __forceinline__ void update1(float g, float& v, float& a, float lr, float mu) {
  if (KIND == TFG_OPT_SGD) {
    v = __fsub_rn(v, __fmul_rn(g, lr));
  } else if (KIND == TFG_OPT_MOMENTUM) {
    a = __fadd_rn(__fmul_rn(a, mu), g);
    v = __fsub_rn(v, __fmul_rn(a, lr));
  } else {
    a = __fadd_rn(__fmul_rn(a, mu), g);
    v = __fsub_rn(v, __fadd_rn(__fmul_rn(g, lr), __fmul_rn(__fmul_rn(a, mu), lr)));
  }
}

// wait until every other rank has stored an epoch >= `epoch` into this rank's row `row` of the flag
// block; false after 20 s (a peer rank died: fail, do not hang)
__device__ __forceinline__ bool wait_peers(const ApplyArgs& a, int row, uint32_t epoch, int t) {
  bool ok = true;
  if (t < a.n_ranks && t != a.me) {
    const uint32_t* f = a.flags[a.me] + row * TFG_MAX_GPUS + t;
    unsigned long long t0, t1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    for (;;) {
      uint32_t seen;
      asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(seen) : "l"(f) : "memory");
      if ((int32_t)(seen - epoch) >= 0) break;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
      if (t1 - t0 > 20ull * 1000 * 1000 * 1000) { ok = false; break; }
      __nanosleep(100);
    }
  }
  return ok;
}
__device__ __forceinline__ void signal_peers(const ApplyArgs& a, int row, uint32_t epoch, int t) {
  if (t < a.n_ranks && t != a.me)
    asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(a.flags[t] + row * TFG_MAX_GPUS + a.me), "r"(epoch) : "memory");
}

// The whole exchange of one step in ONE launch per rank.  With a.sync the two cross-rank barriers
// that used to bracket it as launches of their own are part of the kernel:
//   1. CTA 0 tells every peer "my gradients (and my forward's reads of the parameters) are done"
//      (release store of the step's epoch into the peer's `ready` row), and every CTA waits until all
//      peers have said so before it loads a remote gradient or overwrites a remote parameter;
//   2. the CTA that finishes last tells every peer "my stores into your parameter buffer have landed"
//      and waits for the same word from every peer before the kernel ends, so the next forward on
//      this stream reads complete parameters.
// VEC = float4 per thread and chunk (4: large shards, bytes in flight; 1: small shards, more CTAs).
template <int KIND, int VEC, bool SYNC>
__global__ void __launch_bounds__(kApplyThreads, SYNC ? 3 : 1)
push_apply_pull_kernel(const __grid_constant__ ApplyArgs a) {
  constexpr int kChunk = VEC * 4 * kApplyThreads;
  constexpr int kLoadGroup = load_group<VEC>();
  constexpr bool sync = SYNC;
  uint32_t epoch = 0;
  if (sync) {
    __shared__ int peers_ok;
    epoch = *reinterpret_cast<volatile uint32_t*>(a.epoch) + 1;   // advanced by the last CTA, after every CTA has read it
    if (threadIdx.x == 0) peers_ok = 1;
    __syncthreads();
    if (blockIdx.x == 0) {
      __threadfence_system();
      signal_peers(a, kFlagsReady, epoch, threadIdx.x);
    }
    if (!wait_peers(a, kFlagsReady, epoch, threadIdx.x)) { peers_ok = 0; atomicExch(a.err_out, 1); }
    __syncthreads();
    // a peer rank is gone: its gradients are stale, so no parameter may change; the host reports
    // TFG_EPEER when it reads the error word
    if (!peers_ok) return;
  } else if (a.err && *reinterpret_cast<const volatile int*>(a.err) != 0) {
    return;   // the barrier launched in front of this kernel gave up
  }
  // mean loss over replicas, replica order (the step's scalar result)
  if (blockIdx.x == 0 && threadIdx.x == 0 && a.n_loss > 0) {
    float s = a.loss_src[0][0];
    for (int r = 1; r < a.n_loss; ++r) s = __fadd_rn(s, a.loss_src[r][0]);
    a.loss_dst[0] = __fdiv_rn(s, (float)a.n_loss);
  }
  for (long long c = blockIdx.x; c < a.n_chunks; c += gridDim.x) {
    Segment sg = a.seg0;
    if (a.segs) {  // few segments: binary search on the first chunk index
      int lo = 0, hi = a.n_segs - 1;
      while (lo < hi) {
        int mid = (lo + hi + 1) >> 1;
        if (a.segs[mid].chunk0 <= c) lo = mid; else hi = mid - 1;
      }
      sg = a.segs[lo];
    }
    const long long base = sg.off + (c - sg.chunk0) * (long long)kChunk;
    const long long end = sg.off + sg.cnt;
    float4 g[VEC];
    long long idx[VEC];
#pragma unroll
    for (int u = 0; u < VEC; ++u) {
      idx[u] = base + 4ll * (threadIdx.x + u * kApplyThreads);
      g[u] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    // the replicas' gradients kLoadGroup at a time: all loads of a group are in flight before the first
    // add (remote loads are NVLink round trips), and the sum still runs in replica order
    if constexpr (kLoadGroup == 1) {
#pragma unroll
      for (int u = 0; u < VEC; ++u)
        if (idx[u] < end) g[u] = ld_stream(a.grads[0] + idx[u]);
      for (int r = 1; r < a.n_grads; ++r) {
        float4 t[VEC];
#pragma unroll
        for (int u = 0; u < VEC; ++u)
          if (idx[u] < end) t[u] = ld_stream(a.grads[r] + idx[u]);
#pragma unroll
        for (int u = 0; u < VEC; ++u)
          if (idx[u] < end) {
            g[u].x = __fadd_rn(g[u].x, t[u].x); g[u].y = __fadd_rn(g[u].y, t[u].y);
            g[u].z = __fadd_rn(g[u].z, t[u].z); g[u].w = __fadd_rn(g[u].w, t[u].w);
          }
      }
    } else {
      for (int r0 = 0; r0 < a.n_grads; r0 += kLoadGroup) {
        float4 t[kLoadGroup][VEC];
  #pragma unroll
        for (int rr = 0; rr < kLoadGroup; ++rr)
          if (r0 + rr < a.n_grads) {
  #pragma unroll
            for (int u = 0; u < VEC; ++u)
              if (idx[u] < end) t[rr][u] = ld_stream(a.grads[r0 + rr] + idx[u]);
          }
  #pragma unroll
        for (int rr = 0; rr < kLoadGroup; ++rr)
          if (r0 + rr < a.n_grads) {
  #pragma unroll
            for (int u = 0; u < VEC; ++u)
              if (idx[u] < end) {
                if (r0 + rr == 0) {
                  g[u] = t[rr][u];
                } else {
                  g[u].x = __fadd_rn(g[u].x, t[rr][u].x); g[u].y = __fadd_rn(g[u].y, t[rr][u].y);
                  g[u].z = __fadd_rn(g[u].z, t[rr][u].z); g[u].w = __fadd_rn(g[u].w, t[rr][u].w);
                }
              }
          }
      }
    }
    float4 v[VEC], m[VEC];
#pragma unroll
    for (int u = 0; u < VEC; ++u)
      if (idx[u] < end) {
        v[u] = *reinterpret_cast<const float4*>(a.var + idx[u]);
        if (KIND != TFG_OPT_SGD) m[u] = *reinterpret_cast<const float4*>(a.accum + idx[u]);
      }
#pragma unroll
    for (int u = 0; u < VEC; ++u)
      if (idx[u] < end) {
        float4 gg = g[u];
        gg.x = __fdiv_rn(gg.x, a.divisor); gg.y = __fdiv_rn(gg.y, a.divisor);
        gg.z = __fdiv_rn(gg.z, a.divisor); gg.w = __fdiv_rn(gg.w, a.divisor);
        update1<KIND>(gg.x, v[u].x, m[u].x, a.lr, a.mu);
        update1<KIND>(gg.y, v[u].y, m[u].y, a.lr, a.mu);
        update1<KIND>(gg.z, v[u].z, m[u].z, a.lr, a.mu);
        update1<KIND>(gg.w, v[u].w, m[u].w, a.lr, a.mu);
        if (KIND != TFG_OPT_SGD) *reinterpret_cast<float4*>(a.accum + idx[u]) = m[u];
        if (a.var) *reinterpret_cast<float4*>(a.var + idx[u]) = v[u];
        for (int p = 0; p < a.n_pull; ++p) *reinterpret_cast<float4*>(a.pull[p] + idx[u]) = v[u];
        if (a.n_lp > 0 && a.lp_mode == 1) {   // the tensor path reads its weights from this bf16 copy: no cast kernels per step
          const __nv_bfloat162 lo = __floats2bfloat162_rn(v[u].x, v[u].y), hi = __floats2bfloat162_rn(v[u].z, v[u].w);
          uint2 pk;
          pk.x = *reinterpret_cast<const uint32_t*>(&lo);
          pk.y = *reinterpret_cast<const uint32_t*>(&hi);
          for (int p = 0; p < a.n_lp; ++p) *reinterpret_cast<uint2*>(static_cast<__nv_bfloat16*>(a.lp[p]) + idx[u]) = pk;
        } else if (a.n_lp > 0) {              // 3xTF32: the low-order plane of the new value
          const float4 lo = make_float4(tc::tf32_lo(v[u].x), tc::tf32_lo(v[u].y), tc::tf32_lo(v[u].z), tc::tf32_lo(v[u].w));
          for (int p = 0; p < a.n_lp; ++p) *reinterpret_cast<float4*>(static_cast<float*>(a.lp[p]) + idx[u]) = lo;
        }
      }
  }
  if (sync) {
    __shared__ int last;
    __threadfence_system();            // this thread's peer stores, before the ticket
    __syncthreads();
    if (threadIdx.x == 0) last = atomicAdd(a.done, 1u) == gridDim.x - 1;
    __syncthreads();
    if (last) {                        // every CTA's stores are out: tell the peers, wait for theirs
      __threadfence_system();
      signal_peers(a, kFlagsPulled, epoch, threadIdx.x);
      if (!wait_peers(a, kFlagsPulled, epoch, threadIdx.x)) atomicExch(a.err_out, 1);
      __syncthreads();
      if (threadIdx.x == 0) { *a.done = 0u; *a.epoch = epoch; }
    }
  }
}

// ---- the same exchange with the operands streamed through shared memory by the bulk-copy engine ------
// Multi-GPU ranks (measured against the register kernel above on the same buffers, 4 ranks, 256 M
// parameters: 0.88 vs 0.81 of the measured peer-copy bandwidth per direction; the register kernel stays
// for single-GPU jobs, where it holds 0.97 of HBM peak, and for jobs with too many replicas for the
// stage budget).  Here a producer
// warp requests, per 1024-element chunk, one 4-KB bulk copy (cp.async.bulk, the TMA unit's linear mode)
// per replica gradient -- local or peer memory alike -- plus the owner's parameter and slot values, into
// a ring of shared-memory stages that complete on mbarriers; up to ~200 KB per SM are in flight however
// long the link's latency is.  The 256 consumer threads read their float4 of every array from the
// stage, sum the replicas in replica order, apply the optimizer (update1: no FMA contraction, identical
// bits to the register kernel) and store the new values to the local and peer buffers.
constexpr int kTmaChunk = 1024;                       // elements per chunk and array
constexpr int kTmaThreads = kApplyThreads + 32;       // consumers + the producer warp
constexpr int kTmaSmemBudget = 200 * 1024;

__device__ __forceinline__ void bulk_load(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               :: "r"(tc::smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(tc::smem_u32(bar)) : "memory");
}

// Development aid (make EXTRA=-DTFG_TRACE, scripts/exchange_trace.py): %globaltimer stamps (ns) of the phases
// of the last launch, from CTA 0 and from the CTA that finished last.  Not part of the shipped build.
#ifdef TFG_TRACE
__device__ unsigned long long g_xchg_trace[8];
__device__ __forceinline__ void xchg_stamp(int slot) {
  unsigned long long v;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(v));
  g_xchg_trace[slot] = v;
}
#define XCHG_STAMP(cond, slot) do { if (cond) xchg_stamp(slot); } while (0)
#else
#define XCHG_STAMP(cond, slot)
#endif

template <int KIND, bool SYNC>
__global__ void __launch_bounds__(kTmaThreads, 1)
push_apply_pull_tma_kernel(const __grid_constant__ ApplyArgs a, int stages) {
  extern __shared__ __align__(128) uint8_t smem_dyn[];
  const int n_arr = a.n_grads + 1 + (KIND != TFG_OPT_SGD ? 1 : 0);   // gradients, var, (accum)
  const int stage_floats = n_arr * kTmaChunk;
  float* ring = reinterpret_cast<float*>(smem_dyn);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem_dyn + static_cast<size_t>(stages) * stage_floats * 4);
  uint64_t* empty = full + stages;
  const int t = threadIdx.x;
  uint32_t epoch = 0;
  XCHG_STAMP(blockIdx.x == 0 && t == 0, 0);                      // kernel entry
  if (t == 0) {
    for (int s = 0; s < stages; ++s) { tc::mbar_init(&full[s], 1); tc::mbar_init(&empty[s], kApplyThreads / 32); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (SYNC) {
    __shared__ int peers_ok;
    epoch = *reinterpret_cast<volatile uint32_t*>(a.epoch) + 1;
    if (t == 0) peers_ok = 1;
    __syncthreads();
    if (blockIdx.x == 0) {
      __threadfence_system();
      signal_peers(a, kFlagsReady, epoch, t);
    }
    if (!wait_peers(a, kFlagsReady, epoch, t)) { peers_ok = 0; atomicExch(a.err_out, 1); }
    __syncthreads();
    if (!peers_ok) return;
    XCHG_STAMP(blockIdx.x == 0 && t == 0, 1);                    // every peer is ready
  } else {
    __syncthreads();
    if (a.err && *reinterpret_cast<const volatile int*>(a.err) != 0) return;
  }
  if (blockIdx.x == 0 && t == 0 && a.n_loss > 0) {
    float s = a.loss_src[0][0];
    for (int r = 1; r < a.n_loss; ++r) s = __fadd_rn(s, a.loss_src[r][0]);
    a.loss_dst[0] = __fdiv_rn(s, (float)a.n_loss);
  }
  auto chunk_range = [&](long long c, long long* base, long long* end) {
    Segment sg = a.seg0;
    if (a.segs) {
      int lo = 0, hi = a.n_segs - 1;
      while (lo < hi) {
        int mid = (lo + hi + 1) >> 1;
        if (a.segs[mid].chunk0 <= c) lo = mid; else hi = mid - 1;
      }
      sg = a.segs[lo];
    }
    *base = sg.off + (c - sg.chunk0) * (long long)kTmaChunk;
    *end = sg.off + sg.cnt;
  };
  if (t >= kApplyThreads) {
    // ===== producer warp: one lane requests the chunk's arrays =====
    if (t == kApplyThreads) {
      uint32_t it = 0;
      for (long long c = blockIdx.x; c < a.n_chunks; c += gridDim.x, ++it) {
        const int s = it % stages;
        tc::mbar_wait(&empty[s], ((it / stages) & 1) ^ 1);
        long long base, end;
        chunk_range(c, &base, &end);
        const uint32_t bytes = static_cast<uint32_t>(min((long long)kTmaChunk, end - base)) * 4u;
        tc::mbar_expect_tx(&full[s], bytes * n_arr);
        float* dst = ring + static_cast<size_t>(s) * stage_floats;
        for (int r = 0; r < a.n_grads; ++r) bulk_load(dst + r * kTmaChunk, a.grads[r] + base, bytes, &full[s]);
        bulk_load(dst + a.n_grads * kTmaChunk, a.var + base, bytes, &full[s]);
        if (KIND != TFG_OPT_SGD) bulk_load(dst + (a.n_grads + 1) * kTmaChunk, a.accum + base, bytes, &full[s]);
      }
    }
  } else {
    // ===== consumers =====
    uint32_t it = 0;
    for (long long c = blockIdx.x; c < a.n_chunks; c += gridDim.x, ++it) {
      const int s = it % stages;
      long long base, end;
      chunk_range(c, &base, &end);
      tc::mbar_wait(&full[s], (it / stages) & 1);
      const long long idx = base + 4ll * t;
      const float* src = ring + static_cast<size_t>(s) * stage_floats + 4 * t;
      if (idx < end) {
        float4 g = *reinterpret_cast<const float4*>(src);
        for (int r = 1; r < a.n_grads; ++r) {
          const float4 x = *reinterpret_cast<const float4*>(src + r * kTmaChunk);
          g.x = __fadd_rn(g.x, x.x); g.y = __fadd_rn(g.y, x.y); g.z = __fadd_rn(g.z, x.z); g.w = __fadd_rn(g.w, x.w);
        }
        float4 v = *reinterpret_cast<const float4*>(src + a.n_grads * kTmaChunk);
        float4 m = make_float4(0.f, 0.f, 0.f, 0.f);
        if (KIND != TFG_OPT_SGD) m = *reinterpret_cast<const float4*>(src + (a.n_grads + 1) * kTmaChunk);
        g.x = __fdiv_rn(g.x, a.divisor); g.y = __fdiv_rn(g.y, a.divisor);
        g.z = __fdiv_rn(g.z, a.divisor); g.w = __fdiv_rn(g.w, a.divisor);
        update1<KIND>(g.x, v.x, m.x, a.lr, a.mu);
        update1<KIND>(g.y, v.y, m.y, a.lr, a.mu);
        update1<KIND>(g.z, v.z, m.z, a.lr, a.mu);
        update1<KIND>(g.w, v.w, m.w, a.lr, a.mu);
        if (KIND != TFG_OPT_SGD) *reinterpret_cast<float4*>(a.accum + idx) = m;
        *reinterpret_cast<float4*>(a.var + idx) = v;
        for (int p = 0; p < a.n_pull; ++p) *reinterpret_cast<float4*>(a.pull[p] + idx) = v;
        if (a.n_lp > 0 && a.lp_mode == 1) {
          const __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
          uint2 pk;
          pk.x = *reinterpret_cast<const uint32_t*>(&lo);
          pk.y = *reinterpret_cast<const uint32_t*>(&hi);
          for (int p = 0; p < a.n_lp; ++p) *reinterpret_cast<uint2*>(static_cast<__nv_bfloat16*>(a.lp[p]) + idx) = pk;
        } else if (a.n_lp > 0) {
          const float4 lo = make_float4(tc::tf32_lo(v.x), tc::tf32_lo(v.y), tc::tf32_lo(v.z), tc::tf32_lo(v.w));
          for (int p = 0; p < a.n_lp; ++p) *reinterpret_cast<float4*>(static_cast<float*>(a.lp[p]) + idx) = lo;
        }
      }
      __syncwarp();
      if ((t & 31) == 0) tc::mbar_arrive(&empty[s]);      // this warp has read its part of the stage
    }
  }
  if (SYNC) {
    __shared__ int last;
    XCHG_STAMP(blockIdx.x == 0 && t == 0, 2);                    // CTA 0: its chunks are done
    __threadfence_system();
    __syncthreads();
    XCHG_STAMP(blockIdx.x == 0 && t == 0, 3);                    // CTA 0: its stores have landed
    if (t == 0) last = atomicAdd(a.done, 1u) == gridDim.x - 1;
    __syncthreads();
    if (last) {
      XCHG_STAMP(t == 0, 4);                                     // the last CTA is through
      __threadfence_system();
      signal_peers(a, kFlagsPulled, epoch, t);
      if (!wait_peers(a, kFlagsPulled, epoch, t)) atomicExch(a.err_out, 1);
      __syncthreads();
      XCHG_STAMP(t == 0, 5);                                     // every peer's stores have landed here
      if (t == 0) { *a.done = 0u; *a.epoch = epoch; }
    }
  }
}

// ring depth that fits the budget for this many arrays per stage; < 2 = use the register kernel
static int tma_stages(int n_arr) { return std::min(8, kTmaSmemBudget / (n_arr * kTmaChunk * 4)); }

template <int KIND>
static cudaError_t launch_apply_tma_k(int grid, size_t smem, int stages, cudaStream_t s, const ApplyArgs& a) {
  static bool attr[2] = {false, false};
  const int sy = a.sync ? 1 : 0;
  if (!attr[sy]) {
    cudaError_t e = a.sync ? cudaFuncSetAttribute(push_apply_pull_tma_kernel<KIND, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kTmaSmemBudget + 256)
                           : cudaFuncSetAttribute(push_apply_pull_tma_kernel<KIND, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kTmaSmemBudget + 256);
    if (e != cudaSuccess) return e;
    attr[sy] = true;
  }
  if (a.sync) push_apply_pull_tma_kernel<KIND, true><<<grid, kTmaThreads, smem, s>>>(a, stages);
  else push_apply_pull_tma_kernel<KIND, false><<<grid, kTmaThreads, smem, s>>>(a, stages);
  return cudaGetLastError();
}

static cudaError_t launch_apply_tma(int kind, int grid, size_t smem, int stages, cudaStream_t s, const ApplyArgs& a) {
  if (kind == TFG_OPT_SGD) return launch_apply_tma_k<TFG_OPT_SGD>(grid, smem, stages, s, a);
  if (kind == TFG_OPT_MOMENTUM) return launch_apply_tma_k<TFG_OPT_MOMENTUM>(grid, smem, stages, s, a);
  return launch_apply_tma_k<TFG_OPT_NESTEROV>(grid, smem, stages, s, a);
}

// scalar path for caller buffers whose length or alignment rules out float4
template <int KIND>
__global__ void __launch_bounds__(kApplyThreads)
apply_scalar_kernel(const __grid_constant__ ApplyArgs a, long long count) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < count;
       i += (long long)gridDim.x * blockDim.x) {
    float g = a.grads[0][i];
    for (int r = 1; r < a.n_grads; ++r) g = __fadd_rn(g, a.grads[r][i]);
    g = __fdiv_rn(g, a.divisor);
    float v = a.var[i], m = (KIND != TFG_OPT_SGD) ? a.accum[i] : 0.f;
    update1<KIND>(g, v, m, a.lr, a.mu);
    if (KIND != TFG_OPT_SGD) a.accum[i] = m;
    a.var[i] = v;
    for (int p = 0; p < a.n_pull; ++p) a.pull[p][i] = v;
  }
}

static int grid_for(long long n_chunks) {
  long long g = n_chunks < (long long)kNumSMs * 8 ? n_chunks : (long long)kNumSMs * 8;
  return (int)(g < 1 ? 1 : g);
}

template <int KIND, int VEC>
static void launch_apply_kv(int grid, cudaStream_t s, const ApplyArgs& a) {
  if (a.sync) push_apply_pull_kernel<KIND, VEC, true><<<grid, kApplyThreads, 0, s>>>(a);
  else push_apply_pull_kernel<KIND, VEC, false><<<grid, kApplyThreads, 0, s>>>(a);
}

static cudaError_t launch_apply(int kind, int vec, int grid, cudaStream_t s, const ApplyArgs& a) {
  switch (kind * 2 + (vec == 4 ? 1 : 0)) {
    case TFG_OPT_SGD * 2: launch_apply_kv<TFG_OPT_SGD, 1>(grid, s, a); break;
    case TFG_OPT_SGD * 2 + 1: launch_apply_kv<TFG_OPT_SGD, 4>(grid, s, a); break;
    case TFG_OPT_MOMENTUM * 2: launch_apply_kv<TFG_OPT_MOMENTUM, 1>(grid, s, a); break;
    case TFG_OPT_MOMENTUM * 2 + 1: launch_apply_kv<TFG_OPT_MOMENTUM, 4>(grid, s, a); break;
    case TFG_OPT_NESTEROV * 2: launch_apply_kv<TFG_OPT_NESTEROV, 1>(grid, s, a); break;
    default: launch_apply_kv<TFG_OPT_NESTEROV, 4>(grid, s, a); break;
  }
  return cudaGetLastError();
}

template <typename K>
static cudaError_t launch_kind(int kind, K&& f) {
  switch (kind) {
    case TFG_OPT_SGD: f(std::integral_constant<int, TFG_OPT_SGD>()); break;
    case TFG_OPT_MOMENTUM: f(std::integral_constant<int, TFG_OPT_MOMENTUM>()); break;
    default: f(std::integral_constant<int, TFG_OPT_NESTEROV>()); break;
  }
  return cudaGetLastError();
}

int launch_job_apply(tfg_job* j, bool summed_local, int n_summed, bool whole, bool sync) {
  ApplyArgs a = {};
  if (summed_local) {
    a.grads[0] = j->grads[0];
    a.n_grads = 1;
    a.divisor = (float)n_summed;
    a.var = j->params;
    a.n_pull = 0;
    a.n_loss = 0;
    if (j->params_lp) a.lp[a.n_lp++] = j->params_lp;   // owned ranges only
  } else {
    for (int r = 0; r < j->n_replicas; ++r) a.grads[r] = j->peer_grads[r];
    a.n_grads = j->n_replicas;
    a.divisor = (float)j->n_replicas;
    a.var = j->params;
    a.n_pull = 0;
    for (int g = 0; g < j->n_gpus; ++g)
      if (g != j->gpu_rank) a.pull[a.n_pull++] = j->peer_params[g];
    if (j->params_lp && !lp_derived_per_step(j))
      for (int g = 0; g < j->n_gpus; ++g) a.lp[a.n_lp++] = j->peer_params_lp[g];
    a.n_loss = (j->model || j->resnet) ? j->n_replicas : 0;
    for (int r = 0; r < a.n_loss; ++r) a.loss_src[r] = j->peer_loss[r % j->n_gpus] + r;
  }
  a.lp_mode = j->lp_mode;
  a.err = j->n_gpus > 1 ? j->dev_err : nullptr;
  a.loss_dst = j->loss + TFG_MAX_REPLICAS;
  a.accum = j->accum;
  a.segs = j->d_segs;
  a.n_segs = (int)j->segs.size();
  a.n_chunks = j->n_chunks;
  int vec = j->chunk_elems == kChunkBig ? 4 : 1;
  if (whole) {   // every rank applies the whole flat space (the all-reduce baseline)
    a.segs = nullptr;
    a.seg0 = Segment{0, j->flat_count, 0};
    a.n_segs = 1;
    vec = 4;
    a.n_chunks = (j->flat_count + kChunkBig - 1) / kChunkBig;
  }
  a.kind = j->opt_kind;
  a.lr = j->lr;
  a.mu = j->mu;
  a.sync = sync && j->n_gpus > 1 && !summed_local;
  if (a.sync) {
    a.me = j->gpu_rank;
    a.n_ranks = j->n_gpus;
    for (int g = 0; g < j->n_gpus; ++g) a.flags[g] = j->peer_flags[g];
    a.epoch = j->dev_epoch + 1;
    a.done = reinterpret_cast<unsigned int*>(j->dev_epoch + 2);
    a.err_out = j->dev_err;
  }
  if (a.n_segs == 0 && a.n_loss == 0 && !a.sync) return TFG_OK;  // this rank owns nothing and has nobody to wait for
  if (a.n_segs == 0) a.n_chunks = 0;
  // ranks of a multi-GPU job stream their operands through shared memory (bulk copies); the segment
  // table is cut into 1024-element chunks for it
  const int n_arr = a.n_grads + 1 + (j->opt_kind != TFG_OPT_SGD ? 1 : 0);
  const int stages = tma_stages(n_arr);
  cudaError_t e;
  if (j->n_gpus > 1 && !summed_local && !whole && j->chunk_elems == kTmaChunk && stages >= 2) {
    const size_t smem = (size_t)stages * n_arr * kTmaChunk * 4 + 2 * stages * sizeof(uint64_t);
    const int per_sm = std::max(1, (int)((227 * 1024) / (smem + 1024)));
    const int grid = (int)std::max<long long>(1, std::min<long long>(a.n_chunks, (long long)kNumSMs * per_sm));
    e = launch_apply_tma(j->opt_kind, grid, smem, stages, j->stream, a);
  } else {
    e = launch_apply(j->opt_kind, vec, grid_for(a.n_chunks), j->stream, a);
  }
  prof_mark(j, "push_apply_pull_kernel");
  if (e != cudaSuccess) return fail_cuda(j, e, "push_apply_pull_kernel");
  return TFG_OK;
}

int launch_dense_apply(int kind, float* var, float* accum, const float* const* grads, int n_grads,
                       float divisor, int64_t count, float lr, float mu, float* const* pull,
                       int n_pull, cudaStream_t s) {
  ApplyArgs a = {};
  bool aligned = ((uintptr_t)var % 16 == 0) && (!accum || (uintptr_t)accum % 16 == 0);
  for (int r = 0; r < n_grads; ++r) {
    if (!grads[r]) return fail(nullptr, TFG_EINVAL, "NULL gradient pointer");
    a.grads[r] = grads[r];
    aligned = aligned && ((uintptr_t)grads[r] % 16 == 0);
  }
  for (int p = 0; p < n_pull; ++p) {
    if (!pull[p]) return fail(nullptr, TFG_EINVAL, "NULL pull pointer");
    a.pull[p] = pull[p];
    aligned = aligned && ((uintptr_t)pull[p] % 16 == 0);
  }
  a.n_grads = n_grads;
  a.n_pull = n_pull;
  a.var = var;
  a.accum = accum;
  a.kind = kind;
  a.divisor = divisor;
  a.lr = lr;
  a.mu = mu;
  a.n_loss = 0;
  cudaError_t e = cudaSuccess;
  int64_t vec_count = aligned ? (count / 4) * 4 : 0;
  if (vec_count > 0) {
    a.segs = nullptr;
    a.seg0 = Segment{0, vec_count, 0};
    a.n_segs = 1;
    const int vec = vec_count >= (long long)kChunkBig * 2 * kNumSMs ? 4 : 1;
    const int chunk = vec == 4 ? kChunkBig : kChunkSmall;
    a.n_chunks = (vec_count + chunk - 1) / chunk;
    int grid = grid_for(a.n_chunks);
    e = launch_apply(kind, vec, grid, s, a);
    if (e != cudaSuccess) return fail_cuda(nullptr, e, "push_apply_pull_kernel");
  }
  if (vec_count < count) {
    ApplyArgs t = a;
    for (int r = 0; r < n_grads; ++r) t.grads[r] = a.grads[r] + vec_count;
    for (int p = 0; p < n_pull; ++p) t.pull[p] = a.pull[p] + vec_count;
    t.var = var + vec_count;
    t.accum = accum ? accum + vec_count : nullptr;
    long long rem = count - vec_count;
    int grid = (int)((rem + kApplyThreads - 1) / kApplyThreads);
    if (grid > kNumSMs * 8) grid = kNumSMs * 8;
    e = launch_kind(kind, [&](auto k) {
      apply_scalar_kernel<decltype(k)::value><<<grid, kApplyThreads, 0, s>>>(t, rem);
    });
    if (e != cudaSuccess) return fail_cuda(nullptr, e, "apply_scalar_kernel");
  }
  return TFG_OK;
}

__global__ void __launch_bounds__(256)
cast_lp_kernel(const float* __restrict__ src, void* __restrict__ dst, long long n, int lp_mode) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    if (lp_mode == 1) static_cast<__nv_bfloat16*>(dst)[i] = __float2bfloat16_rn(src[i]);
    else static_cast<float*>(dst)[i] = tc::tf32_lo(src[i]);
  }
}

// the low-order plane, four values per thread (16-byte aligned buffers, n a multiple of 4)
__global__ void __launch_bounds__(256)
split_lo4_kernel(const float4* __restrict__ src, float4* __restrict__ dst, long long n4) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    const float4 v = src[i];
    dst[i] = make_float4(tc::tf32_lo(v.x), tc::tf32_lo(v.y), tc::tf32_lo(v.z), tc::tf32_lo(v.w));
  }
}

int launch_cast_lp(const float* src, void* dst, long long n, int lp_mode, cudaStream_t s) {
  if (n <= 0) return TFG_OK;
  if (lp_mode == 2 && (n & 3) == 0 && ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15) == 0) {
    long long blocks = (n / 4 + 255) / 256;
    if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
    split_lo4_kernel<<<(int)blocks, 256, 0, s>>>(reinterpret_cast<const float4*>(src), static_cast<float4*>(dst), n / 4);
    return cudaGetLastError() == cudaSuccess ? TFG_OK : TFG_ECUDA;
  }
  long long blocks = (n + 255) / 256;
  if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
  cast_lp_kernel<<<(int)blocks, 256, 0, s>>>(src, dst, n, lp_mode);
  return cudaGetLastError() == cudaSuccess ? TFG_OK : TFG_ECUDA;
}

// ---- cross-rank barrier ----------------------------------------------------------------
struct BarrierArgs {
  uint32_t* flags[TFG_MAX_GPUS];  // every rank's flag array (own one local, others peer-mapped)
  int* err;
  int me, n;
  uint32_t* epoch_ctr;   // device-resident: the kernel advances it, so a captured launch replays correctly
};

__global__ void barrier_kernel(const __grid_constant__ BarrierArgs a) {
  int t = threadIdx.x;
  __shared__ uint32_t epoch_s;
  if (t == 0) epoch_s = ++(*a.epoch_ctr);   // every rank runs the same number of barriers
  __syncthreads();
  const uint32_t epoch = epoch_s;
  if (t >= a.n || t == a.me) return;
  __threadfence_system();  // my earlier stores (gradients, pulled parameters) before the flag
  asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(a.flags[t] + a.me), "r"(epoch) : "memory");
  unsigned long long t0, t1;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  for (;;) {
    uint32_t seen;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(seen) : "l"(a.flags[a.me] + t) : "memory");
    if ((int32_t)(seen - epoch) >= 0) break;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
    if (t1 - t0 > 20ull * 1000 * 1000 * 1000) {  // 20 s: a peer rank died; fail, do not hang
      atomicExch(a.err, 1);
      break;
    }
    __nanosleep(200);
  }
}

int launch_barrier(tfg_job* j) {
  BarrierArgs a = {};
  for (int g = 0; g < j->n_gpus; ++g) a.flags[g] = j->peer_flags[g];
  a.err = j->dev_err;
  a.me = j->gpu_rank;
  a.n = j->n_gpus;
  a.epoch_ctr = j->dev_epoch;
  barrier_kernel<<<1, 32, 0, j->stream>>>(a);
  prof_mark(j, "barrier_kernel");
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail_cuda(j, e, "barrier_kernel");
  return TFG_OK;
}

}  // namespace tfg

#ifdef TFG_TRACE
extern "C" int tfg_debug_xchg_trace(unsigned long long* out /* [8] */) {
  cudaError_t e = cudaMemcpyFromSymbol(out, tfg::g_xchg_trace, sizeof(unsigned long long) * 8);
  return e == cudaSuccess ? TFG_OK : tfg::fail_cuda(nullptr, e, "cudaMemcpyFromSymbol(trace)");
}
#endif
