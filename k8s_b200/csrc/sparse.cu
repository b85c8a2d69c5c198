// Sparse embedding PS path: partition index math, row gather and
// scatter-sub with duplicate ids applied in order of appearance.
//
// HBM-bound random-row traffic.  Algorithmic bytes per id (fp32 rows of width D, int64
// ids): gather 8 + 8*D; scatter-sub 8 + 12*D.
//
// scatter-sub = stable LSD radix sort of (local row, position) pairs in scratch memory,
// then one warp per distinct row walks its run in position order, keeping the row in
// registers: table[row] = fl(table[row] - fl(scale * rows[i])) for each i in turn, which
// is what a sequential CPU loop over i computes ([TF1.3] scatter_sub on the CPU).
#include <atomic>
#include <cstring>

#include <cooperative_groups.h>

#include "common.cuh"

namespace tfg {
namespace cg = cooperative_groups;

// Where the ids and the update rows of a (possibly multi-source) batch live.  A batch is the
// concatenation of n_src equally long pieces, source after source -- on several GPUs the pieces are the
// replicas' own buffers, read in place through peer mappings (position p of the batch is element
// p % per of source p / per).  per == 0: one source, position = index.
struct IdSrc {
  const long long* base[TFG_MAX_REPLICAS];
  long long per;
  __device__ __forceinline__ long long at(long long i) const { return per == 0 ? base[0][i] : base[i / per][i % per]; }
};
struct RowSrc {
  const float* base[TFG_MAX_REPLICAS];
  uint32_t per;
  __device__ __forceinline__ const float* at(uint32_t v, int dim) const {
    return per == 0 ? base[0] + (long long)v * dim : base[v / per] + (long long)(v % per) * dim;
  }
};
struct TableSrc {               // the shards of one table: shard k's rows start at base[k] (local or peer memory)
  const float* base[TFG_MAX_REPLICAS];
};

constexpr int kSortThreads = 256;
constexpr int kSortWarps = kSortThreads / 32;
constexpr int kKeysPerWarp = 512;
constexpr int kSortTile = kSortWarps * kKeysPerWarp;  // 4096 keys per block (2048: 10 % slower at a million ids, measured)
constexpr int kMaxDigitBits = 11;                     // 2048 bins: 8 warps x 2048 counters = 64 KB of shared memory
constexpr int kMaxBins = 1 << kMaxDigitBits;
constexpr int kMaxPasses = 3;                         // 32-bit keys

struct Part {
  long long vocab;
  long long ipp;     // vocab / n_shards
  long long extras;  // vocab % n_shards
  int n_shards, shard, strategy;
};

// [TF1.3] embedding_ops: "mod": p = id % S, row = id / S.  "div": p = max(id / (ipp+1),
// (id - extras) / ipp), row = p < extras ? id % (ipp+1) : (id - extras) % ipp.
__host__ __device__ inline void locate(const Part& pt, long long id, int* shard, long long* row) {
  if (pt.strategy == TFG_PART_MOD) {
    *shard = (int)(id % pt.n_shards);
    *row = id / pt.n_shards;
    return;
  }
  if (pt.ipp == 0) {  // fewer rows than shards: shard k holds row k
    *shard = (int)id;
    *row = 0;
    return;
  }
  long long a = id / (pt.ipp + 1), b = (id - pt.extras) / pt.ipp;
  long long p = a > b ? a : b;
  *shard = (int)p;
  *row = p < pt.extras ? id % (pt.ipp + 1) : (id - pt.extras) % pt.ipp;
}

static Part make_part(int64_t vocab, int n_shards, int shard, int strategy) {
  Part p;
  p.vocab = vocab;
  p.ipp = vocab / n_shards;
  p.extras = vocab % n_shards;
  p.n_shards = n_shards;
  p.shard = shard;
  p.strategy = strategy;
  return p;
}

static long long shard_rows(int64_t vocab, int n_shards, int shard, int strategy) {
  if (strategy == TFG_PART_MOD) return shard < vocab ? (vocab - shard + n_shards - 1) / n_shards : 0;
  return vocab / n_shards + (shard < vocab % n_shards ? 1 : 0);
}

// ---- gather ------------------------------------------------------------------------------
// LANES threads per id: a whole warp for rows of more than 64 floats, 16 or 8 lanes for narrower rows
template <int LANES>
__device__ __forceinline__ void copy_row(const float* __restrict__ src, float* __restrict__ dst, int dim, int lane) {
  if ((dim & 3) == 0) {
    for (int c = lane * 4; c < dim; c += LANES * 4)
      *reinterpret_cast<float4*>(dst + c) = *reinterpret_cast<const float4*>(src + c);
  } else {
    for (int c = lane; c < dim; c += LANES) dst[c] = src[c];
  }
}

static int gather_lanes(int dim) {
  const int per = (dim & 3) == 0 ? 4 : 1;
  return dim <= 8 * per ? 8 : dim <= 16 * per ? 16 : 32;
}

template <int LANES>
__global__ void __launch_bounds__(256)
gather_kernel(const float* __restrict__ table, Part pt, const long long* __restrict__ ids,
              long long n, int dim, float* __restrict__ out, int* bad) {
  const int lane = threadIdx.x % LANES;
  const long long groups = (long long)gridDim.x * (blockDim.x / LANES);
  for (long long i = blockIdx.x * (long long)(blockDim.x / LANES) + threadIdx.x / LANES; i < n; i += groups) {
    long long id = ids[i];
    if (id < 0 || id >= pt.vocab) {
      if (lane == 0 && bad) atomicAdd(bad, 1);
      continue;
    }
    int sh;
    long long row;
    locate(pt, id, &sh, &row);
    if (sh != pt.shard) continue;
    copy_row<LANES>(table + row * (long long)dim, out + i * (long long)dim, dim, lane);
  }
}

// the same with every shard of the table addressable from this GPU (own shards local, the others
// peer-mapped): a worker fetches the rows of all its ids in one launch, straight from their owners
template <int LANES>
__global__ void __launch_bounds__(256)
gather_multi_kernel(const __grid_constant__ TableSrc tables, Part pt, const long long* __restrict__ ids,
                    long long n, int dim, float* __restrict__ out, int* bad) {
  const int lane = threadIdx.x % LANES;
  const long long groups = (long long)gridDim.x * (blockDim.x / LANES);
  for (long long i = blockIdx.x * (long long)(blockDim.x / LANES) + threadIdx.x / LANES; i < n; i += groups) {
    long long id = ids[i];
    if (id < 0 || id >= pt.vocab) {
      if (lane == 0 && bad) atomicAdd(bad, 1);
      continue;
    }
    int sh;
    long long row;
    locate(pt, id, &sh, &row);
    copy_row<LANES>(tables.base[sh] + row * (long long)dim, out + i * (long long)dim, dim, lane);
  }
}

// ---- stable radix sort of (row key, position) ------------------------------------------
__global__ void __launch_bounds__(256)
make_keys_kernel(Part pt, const __grid_constant__ IdSrc ids, long long n, uint32_t sentinel,
                 uint32_t* __restrict__ keys, uint32_t* __restrict__ vals) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x) {
    long long id = ids.at(i);
    uint32_t k = sentinel;
    if (id >= 0 && id < pt.vocab) {
      int sh;
      long long row;
      locate(pt, id, &sh, &row);
      if (sh == pt.shard) k = (uint32_t)row;
    }
    keys[i] = k;
    vals[i] = (uint32_t)i;
  }
}

// ---- owned ids only ---------------------------------------------------------------------------
// With several shards most ids of a batch belong to other shards (all but 1/N of them when every GPU
// of the box owns one).  Three small kernels keep the owned ones, in batch order, before anything is
// sorted: (1) row keys of every id (sentinel = not mine) and the count per tile, (2) scan of the tile
// counts, (3) stable compaction of (key, position).  The number kept stays on the device (*n_live);
// the kernels after this read it there, under launch shapes sized for the whole batch.
__global__ void __launch_bounds__(kSortThreads)
owned_keys_kernel(Part pt, const __grid_constant__ IdSrc ids, long long n, uint32_t sentinel,
                  uint32_t* __restrict__ keys_all, uint32_t* __restrict__ tile_count) {
  __shared__ uint32_t warp_cnt[kSortWarps];
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long base = blockIdx.x * (long long)kSortTile + w * kKeysPerWarp;
  uint32_t mine = 0;
  for (int c = 0; c < kKeysPerWarp / 32; ++c) {
    const long long i = base + c * 32 + lane;
    if (i >= n) break;
    const long long id = ids.at(i);
    uint32_t k = sentinel;
    if (id >= 0 && id < pt.vocab) {
      int sh;
      long long row;
      locate(pt, id, &sh, &row);
      if (sh == pt.shard) { k = (uint32_t)row; ++mine; }
    }
    keys_all[i] = k;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mine += __shfl_xor_sync(0xffffffffu, mine, o);
  if (lane == 0) warp_cnt[w] = mine;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t t = 0;
#pragma unroll
    for (int x = 0; x < kSortWarps; ++x) t += warp_cnt[x];
    tile_count[blockIdx.x] = t;
  }
}

// exclusive scan of the tile counts in place (one block); the total goes to *n_live
__global__ void __launch_bounds__(1024) owned_scan_kernel(uint32_t* tile_count, int n_tiles, uint32_t* n_live) {
  __shared__ uint32_t warp_sums[32];
  __shared__ uint32_t carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  for (int base = 0; base < n_tiles; base += 1024) {
    const int i = base + threadIdx.x;
    const uint32_t v = i < n_tiles ? tile_count[i] : 0u;
    uint32_t x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += y;
    }
    if (lane == 31) warp_sums[w] = x;
    __syncthreads();
    if (w == 0) {
      uint32_t sm = warp_sums[lane], t = sm;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        uint32_t y = __shfl_up_sync(0xffffffffu, t, o);
        if (lane >= o) t += y;
      }
      warp_sums[lane] = t - sm;
    }
    __syncthreads();
    const uint32_t excl = carry + warp_sums[w] + x - v;
    if (i < n_tiles) tile_count[i] = excl;
    __syncthreads();
    if (threadIdx.x == 1023) carry = excl + v;
    __syncthreads();
  }
  if (threadIdx.x == 0) *n_live = carry;
}

__global__ void __launch_bounds__(kSortThreads)
owned_compact_kernel(const uint32_t* __restrict__ keys_all, long long n, uint32_t sentinel,
                     const uint32_t* __restrict__ tile_off, uint32_t* __restrict__ keys, uint32_t* __restrict__ vals) {
  __shared__ uint32_t warp_cnt[kSortWarps];
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long base = blockIdx.x * (long long)kSortTile + w * kKeysPerWarp;
  uint32_t mine = 0;
  for (int c = 0; c < kKeysPerWarp / 32; ++c) {
    const long long i = base + c * 32 + lane;
    mine += __popc(__ballot_sync(0xffffffffu, i < n && keys_all[i] != sentinel));
  }
  if (lane == 0) warp_cnt[w] = mine;
  __syncthreads();
  uint32_t pos = tile_off[blockIdx.x];
  for (int x = 0; x < w; ++x) pos += warp_cnt[x];
  for (int c = 0; c < kKeysPerWarp / 32; ++c) {
    const long long i = base + c * 32 + lane;
    const uint32_t k = i < n ? keys_all[i] : sentinel;
    const uint32_t keep = __ballot_sync(0xffffffffu, k != sentinel);
    if (k != sentinel) {
      const uint32_t at = pos + __popc(keep & ((1u << lane) - 1u));
      keys[at] = k;
      vals[at] = (uint32_t)i;
    }
    pos += __popc(keep);
  }
}

// digit counts of this warp's sub-tile into cnt[warp * bins + digit]
// *n_live, when given, is the number of keys (fewer than the launch was shaped for)
__device__ __forceinline__ long long live_count(long long n, const uint32_t* n_live) { return n_live ? (long long)*n_live : n; }

__device__ __forceinline__ void warp_digit_counts(const uint32_t* keys, long long n, int shift, int bins,
                                                  uint32_t* cnt) {
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint32_t* mine = cnt + w * bins;
  for (int d = lane; d < bins; d += 32) mine[d] = 0;
  __syncwarp();
  const long long base = blockIdx.x * (long long)kSortTile + w * kKeysPerWarp;
  const uint32_t mask = (uint32_t)bins - 1u;
  for (int c = 0; c < kKeysPerWarp / 32; ++c) {
    long long i = base + c * 32 + lane;
    if (i < n) atomicAdd(&mine[(keys[i] >> shift) & mask], 1u);
  }
  __syncwarp();
}

// hist[digit][block] (digit-major) and the digit totals of the whole array
__global__ void __launch_bounds__(kSortThreads)
sort_hist_kernel(const uint32_t* __restrict__ keys, long long n, const uint32_t* __restrict__ n_live, int shift, int bins,
                 uint32_t* __restrict__ hist, int n_blocks, uint32_t* __restrict__ totals) {
  extern __shared__ uint32_t cnt[];
  n = live_count(n, n_live);
  if (blockIdx.x * (long long)kSortTile >= n) return;   // tiles past the live keys: their hist columns are never read
  warp_digit_counts(keys, n, shift, bins, cnt);
  __syncthreads();
  for (int d = threadIdx.x; d < bins; d += kSortThreads) {
    uint32_t s = 0;
#pragma unroll
    for (int w = 0; w < kSortWarps; ++w) s += cnt[w * bins + d];
    hist[(size_t)d * n_blocks + blockIdx.x] = s;
    if (s) atomicAdd(&totals[d], s);
  }
}

// Exclusive scan of hist in place, one warp per digit: the digit's base is the sum of the totals of
// the digits below it, then the warp scans its own row of n_blocks counts.  (The first version scanned
// all bins * n_blocks entries with one block: 128 dependent iterations per pass for a million ids,
// more time than the rest of the sort together.)
__global__ void __launch_bounds__(kSortThreads)
sort_scan_kernel(uint32_t* __restrict__ hist, long long n, const uint32_t* __restrict__ n_live, int n_blocks, int bins,
                 const uint32_t* __restrict__ totals) {
  const int live_blocks = (int)((live_count(n, n_live) + kSortTile - 1) / kSortTile);
  const int lane = threadIdx.x & 31;
  const int d = blockIdx.x * kSortWarps + (threadIdx.x >> 5);
  if (d >= bins) return;
  uint32_t run = 0;
  for (int e = lane; e < d; e += 32) run += totals[e];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) run += __shfl_xor_sync(0xffffffffu, run, o);
  uint32_t* row = hist + (size_t)d * n_blocks;
  for (int b0 = 0; b0 < live_blocks; b0 += 32) {
    const int b = b0 + lane;
    const uint32_t v = b < live_blocks ? row[b] : 0u;
    uint32_t x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += y;
    }
    if (b < live_blocks) row[b] = run + x - v;
    run += __shfl_sync(0xffffffffu, x, 31);
  }
}

__global__ void __launch_bounds__(kSortThreads)
sort_scatter_kernel(const uint32_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                    long long n, const uint32_t* __restrict__ n_live, int shift, int bins,
                    const uint32_t* __restrict__ hist, int n_blocks,
                    uint32_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out) {
  extern __shared__ uint32_t cnt[];
  n = live_count(n, n_live);
  if (blockIdx.x * (long long)kSortTile >= n) return;
  warp_digit_counts(keys_in, n, shift, bins, cnt);
  __syncthreads();
  // cnt[w][d] <- global position of the first key of digit d in warp w's sub-tile
  for (int d = threadIdx.x; d < bins; d += kSortThreads) {
    uint32_t run = hist[(size_t)d * n_blocks + blockIdx.x];
#pragma unroll
    for (int w = 0; w < kSortWarps; ++w) {
      uint32_t c = cnt[w * bins + d];
      cnt[w * bins + d] = run;
      run += c;
    }
  }
  __syncthreads();
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint32_t* mine = cnt + w * bins;
  const uint32_t mask = (uint32_t)bins - 1u;
  const long long base = blockIdx.x * (long long)kSortTile + w * kKeysPerWarp;
  for (int c = 0; c < kKeysPerWarp / 32; ++c) {
    long long i = base + c * 32 + lane;
    bool live = i < n;
    uint32_t k = live ? keys_in[i] : 0u, v = live ? vals_in[i] : 0u;
    uint32_t d = live ? ((k >> shift) & mask) : (uint32_t)bins + lane;  // dead lanes match nobody
    uint32_t peers = __match_any_sync(0xffffffffu, d);
    uint32_t before = __popc(peers & ((1u << lane) - 1u));
    uint32_t pos = 0;
    if (live) pos = mine[d] + before;
    __syncwarp();
    if (live && before == 0) mine[d] += __popc(peers);  // lowest lane of the group
    __syncwarp();
    if (live) {
      keys_out[pos] = k;
      vals_out[pos] = v;
    }
  }
}

// ---- apply the sorted runs -----------------------------------------------------------------
template <int VEC>
__global__ void __launch_bounds__(256)
scatter_sub_runs_kernel(float* __restrict__ table, int dim, const uint32_t* __restrict__ keys,
                        const uint32_t* __restrict__ vals, long long n, const uint32_t* __restrict__ n_live,
                        uint32_t sentinel, const __grid_constant__ RowSrc rows, float scale) {
  n = live_count(n, n_live);
  const int lane = threadIdx.x & 31;
  const long long warps = (long long)gridDim.x * (blockDim.x >> 5);
  for (long long p = blockIdx.x * (long long)(blockDim.x >> 5) + (threadIdx.x >> 5); p < n; p += warps) {
    const uint32_t k = keys[p];
    if (k == sentinel) continue;
    if (p > 0 && keys[p - 1] == k) continue;  // not the head of its run
    float* trow = table + (long long)k * dim;
    for (int c0 = 0; c0 < dim; c0 += 32 * VEC) {  // column panel kept in registers
      const int c = c0 + lane * VEC;
      float acc[VEC];
      if (c < dim) {
        if (VEC == 4) {
          float4 t = *reinterpret_cast<const float4*>(trow + c);
          acc[0] = t.x; acc[1 % VEC] = t.y; acc[2 % VEC] = t.z; acc[3 % VEC] = t.w;
        } else {
#pragma unroll
          for (int u = 0; u < VEC; ++u) acc[u] = trow[c + u];
        }
      }
      for (long long q = p; q < n && keys[q] == k; ++q) {
        const float* g = rows.at(vals[q], dim);
        if (c < dim) {
          float gv[VEC];
          if (VEC == 4) {
            float4 t = *reinterpret_cast<const float4*>(g + c);
            gv[0] = t.x; gv[1 % VEC] = t.y; gv[2 % VEC] = t.z; gv[3 % VEC] = t.w;
          } else {
#pragma unroll
            for (int u = 0; u < VEC; ++u) gv[u] = g[c + u];
          }
#pragma unroll
          for (int u = 0; u < VEC; ++u) acc[u] = __fsub_rn(acc[u], __fmul_rn(scale, gv[u]));
        }
      }
      if (c < dim) {
        if (VEC == 4) {
          *reinterpret_cast<float4*>(trow + c) = make_float4(acc[0], acc[1 % VEC], acc[2 % VEC], acc[3 % VEC]);
        } else {
#pragma unroll
          for (int u = 0; u < VEC; ++u) trow[c + u] = acc[u];
        }
      }
    }
  }
}

// ---- long runs (heavy-tailed ids) ------------------------------------------------------------
// A row hit by more than kLongRun ids would serialise one warp for the whole run.  With enough
// scratch such a run is cut into chunks of kRunChunk positions: one warp per chunk sums
// fl(scale * rows[i]) in position order into carry[chunk], then one warp per run subtracts the
// chunk sums from the table row in chunk order.  Deterministic (fixed chunking, fixed order),
// but rounded differently from the one-by-one definition, which short runs still follow
// exactly.
constexpr int kLongRun = 64;
constexpr int kRunChunk = 128;

struct LongRun { uint32_t start, len, key, chunk_base; };
struct ChunkDesc { uint32_t pos, len; };

template <int VEC>
__device__ __forceinline__ void load_vec(const float* p, float (&v)[VEC]) {
  if (VEC == 4) {
    const float4 t = *reinterpret_cast<const float4*>(p);
    v[0] = t.x; v[1 % VEC] = t.y; v[2 % VEC] = t.z; v[3 % VEC] = t.w;
  } else {
#pragma unroll
    for (int u = 0; u < VEC; ++u) v[u] = p[u];
  }
}
template <int VEC>
__device__ __forceinline__ void store_vec(float* p, const float (&v)[VEC]) {
  if (VEC == 4) {
    *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1 % VEC], v[2 % VEC], v[3 % VEC]);
  } else {
#pragma unroll
    for (int u = 0; u < VEC; ++u) p[u] = v[u];
  }
}

// One group of LANES threads per sorted position (a whole warp for rows of more than 64 floats, fewer
// lanes for narrower rows so that no lane idles); only run heads act.  Runs of at most kLongRun ids
// are applied here, one id after another; longer runs are registered for the two kernels below.
template <int VEC, int LANES>
__global__ void __launch_bounds__(256)
scatter_sub_heads_kernel(float* __restrict__ table, int dim, const uint32_t* __restrict__ keys,
                         const uint32_t* __restrict__ vals, long long n, const uint32_t* __restrict__ n_live,
                         uint32_t sentinel, const __grid_constant__ RowSrc rows, float scale, LongRun* __restrict__ runs,
                         ChunkDesc* __restrict__ chunks, uint32_t* __restrict__ counters) {
  static_assert(kLongRun % LANES == 0, "the probe walks whole groups");
  n = live_count(n, n_live);
  const auto tile = cg::tiled_partition<LANES>(cg::this_thread_block());
  const int lane = tile.thread_rank();
  constexpr uint32_t kAll = LANES == 32 ? 0xffffffffu : (1u << LANES) - 1u;
  const long long groups = (long long)gridDim.x * (blockDim.x / LANES);
  for (long long p = blockIdx.x * (long long)(blockDim.x / LANES) + threadIdx.x / LANES; p < n; p += groups) {
    const uint32_t k = keys[p];
    if (k == sentinel) continue;
    if (p > 0 && keys[p - 1] == k) continue;  // not the head of its run
    // run length, probing kLongRun positions ahead
    long long len = 1;
    bool is_long = true;
    for (int blk = 0; blk < kLongRun / LANES; ++blk) {
      const long long q = p + 1 + blk * LANES + lane;
      const uint32_t same = tile.ballot(q < n && keys[q] == k);
      if (same != kAll) { len += __ffs(~same) - 1; is_long = false; break; }
      len += LANES;
    }
    float* trow = table + (long long)k * dim;
    if (!is_long) {
      for (int c0 = 0; c0 < dim; c0 += LANES * VEC) {  // column panel kept in registers
        const int c = c0 + lane * VEC;
        float acc[VEC], gv[VEC];
        if (c < dim) load_vec<VEC>(trow + c, acc);
        for (long long q = p; q < p + len; ++q) {
          const float* g = rows.at(vals[q], dim);
          if (c < dim) {
            load_vec<VEC>(g + c, gv);
#pragma unroll
            for (int u = 0; u < VEC; ++u) acc[u] = __fsub_rn(acc[u], __fmul_rn(scale, gv[u]));
          }
        }
        if (c < dim) store_vec<VEC>(trow + c, acc);
      }
      continue;
    }
    // long run: its end is the first position whose key differs (keys are sorted)
    uint32_t base = 0;
    if (lane == 0) {
      long long lo = p + kLongRun, hi = n;          // keys[lo] == k is known; answer in (lo, n]
      while (lo + 1 < hi) {
        const long long mid = (lo + hi) >> 1;
        if (keys[mid] == k) lo = mid; else hi = mid;
      }
      len = hi - p;
      const uint32_t nch = (uint32_t)((len + kRunChunk - 1) / kRunChunk);
      const uint32_t r = atomicAdd(&counters[0], 1u);
      base = atomicAdd(&counters[1], nch);
      runs[r] = LongRun{(uint32_t)p, (uint32_t)len, k, base};
    }
    len = tile.shfl(len, 0);
    base = tile.shfl(base, 0);
    const uint32_t nch = (uint32_t)((len + kRunChunk - 1) / kRunChunk);
    for (uint32_t jc = lane; jc < nch; jc += LANES) {
      const long long off = (long long)jc * kRunChunk;
      chunks[base + jc] = ChunkDesc{(uint32_t)(p + off), (uint32_t)min((long long)kRunChunk, len - off)};
    }
  }
}

template <int VEC>
__global__ void __launch_bounds__(256)
long_chunk_sum_kernel(int dim, const uint32_t* __restrict__ vals, const __grid_constant__ RowSrc rows, float scale,
                      const ChunkDesc* __restrict__ chunks, const uint32_t* __restrict__ counters,
                      float* __restrict__ carry) {
  const int lane = threadIdx.x & 31;
  const uint32_t n_chunks = counters[1];
  const uint32_t warps = gridDim.x * (blockDim.x >> 5);
  for (uint32_t slot = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); slot < n_chunks; slot += warps) {
    const ChunkDesc d = chunks[slot];
    for (int c0 = 0; c0 < dim; c0 += 32 * VEC) {
      const int c = c0 + lane * VEC;
      float acc[VEC];
#pragma unroll
      for (int u = 0; u < VEC; ++u) acc[u] = 0.f;
      for (uint32_t i0 = 0; i0 < d.len; i0 += 8) {       // 8 independent row loads in flight
        float gv[8][VEC];
#pragma unroll
        for (int t = 0; t < 8; ++t)
          if (i0 + t < d.len && c < dim) load_vec<VEC>(rows.at(vals[d.pos + i0 + t], dim) + c, gv[t]);
#pragma unroll
        for (int t = 0; t < 8; ++t)
          if (i0 + t < d.len && c < dim) {
#pragma unroll
            for (int u = 0; u < VEC; ++u) acc[u] = __fadd_rn(acc[u], __fmul_rn(scale, gv[t][u]));
          }
      }
      if (c < dim) store_vec<VEC>(carry + (long long)slot * dim + c, acc);
    }
  }
}

// One block per long run.  Its warps sum contiguous segments of the run's chunk sums (in chunk order,
// 8 loads in flight), then warp 0 subtracts the segment sums from the table row in segment order: the
// row hit by half of a Zipf batch has thousands of chunks, which one warp alone walked for 0.3 ms.
constexpr int kApplyWarps = 16;

template <int VEC>
__global__ void __launch_bounds__(kApplyWarps * 32)
long_run_apply_kernel(float* __restrict__ table, int dim, const LongRun* __restrict__ runs,
                      const uint32_t* __restrict__ counters, const float* __restrict__ carry) {
  __shared__ float seg[kApplyWarps][32 * VEC];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const uint32_t n_runs = counters[0];
  for (uint32_t r = blockIdx.x; r < n_runs; r += gridDim.x) {
    const LongRun run = runs[r];
    const uint32_t nch = (run.len + kRunChunk - 1) / kRunChunk;
    const uint32_t per = (nch + kApplyWarps - 1) / kApplyWarps;
    const uint32_t j_begin = min(nch, w * per), j_end = min(nch, j_begin + per);
    float* trow = table + (long long)run.key * dim;
    for (int c0 = 0; c0 < dim; c0 += 32 * VEC) {
      const int c = c0 + lane * VEC;
      float sum[VEC];
#pragma unroll
      for (int u = 0; u < VEC; ++u) sum[u] = 0.f;
      for (uint32_t j0 = j_begin; j0 < j_end; j0 += 8) {
        float cv[8][VEC];
#pragma unroll
        for (int t = 0; t < 8; ++t)
          if (j0 + t < j_end && c < dim) load_vec<VEC>(carry + (long long)(run.chunk_base + j0 + t) * dim + c, cv[t]);
#pragma unroll
        for (int t = 0; t < 8; ++t)
          if (j0 + t < j_end && c < dim) {
#pragma unroll
            for (int u = 0; u < VEC; ++u) sum[u] = __fadd_rn(sum[u], cv[t][u]);
          }
      }
#pragma unroll
      for (int u = 0; u < VEC; ++u) seg[w][lane * VEC + u] = sum[u];
      __syncthreads();
      if (w == 0 && c < dim) {
        float acc[VEC];
        load_vec<VEC>(trow + c, acc);
        for (int x = 0; x < kApplyWarps; ++x) {
#pragma unroll
          for (int u = 0; u < VEC; ++u) acc[u] = __fsub_rn(acc[u], seg[x][lane * VEC + u]);
        }
        store_vec<VEC>(trow + c, acc);
      }
      __syncthreads();
    }
  }
}

static int bits_for(uint32_t v) {
  int b = 0;
  while (b < 32 && (v >> b) != 0) ++b;
  return b < 1 ? 1 : b;
}

}  // namespace tfg

using namespace tfg;

extern "C" {

int64_t tfg_sparse_shard_rows(int64_t vocab, int n_shards, int shard, int strategy) {
  if (vocab < 0 || n_shards < 1 || shard < 0 || shard >= n_shards ||
      (strategy != TFG_PART_MOD && strategy != TFG_PART_DIV))
    return -1;
  return shard_rows(vocab, n_shards, shard, strategy);
}

static int check_sparse(int64_t vocab, int dim, int n_shards, int shard, int strategy, int64_t n) {
  if (vocab < 1 || dim < 1 || n < 0) return fail(nullptr, TFG_EINVAL, "bad table shape or id count");
  if (n_shards < 1 || shard < 0 || shard >= n_shards) return fail(nullptr, TFG_EINVAL, "bad shard");
  if (strategy != TFG_PART_MOD && strategy != TFG_PART_DIV)
    return fail(nullptr, TFG_EINVAL, "unknown partition strategy");
  return TFG_OK;
}

int tfg_sparse_gather(const float* table, int64_t vocab, int dim, int n_shards, int shard,
                      int strategy, const int64_t* ids, int64_t n, float* out, int* bad, void* stream) {
  int rc = check_sparse(vocab, dim, n_shards, shard, strategy, n);
  if (rc) return rc;
  if (n == 0) return TFG_OK;
  if (!table || !ids || !out) return fail(nullptr, TFG_EINVAL, "NULL argument");
  Part pt = make_part(vocab, n_shards, shard, strategy);
  const int lanes = gather_lanes(dim), per_block = 256 / lanes;
  long long blocks = (n + per_block - 1) / per_block;
  if (blocks > kNumSMs * 16) blocks = kNumSMs * 16;
  const long long* idp = (const long long*)ids;
  cudaStream_t st = (cudaStream_t)stream;
  if (lanes == 8) gather_kernel<8><<<(int)blocks, 256, 0, st>>>(table, pt, idp, n, dim, out, bad);
  else if (lanes == 16) gather_kernel<16><<<(int)blocks, 256, 0, st>>>(table, pt, idp, n, dim, out, bad);
  else gather_kernel<32><<<(int)blocks, 256, 0, st>>>(table, pt, idp, n, dim, out, bad);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail_cuda(nullptr, e, "gather_kernel");
  return TFG_OK;
}

size_t tfg_sparse_scratch_bytes(int64_t n) {
  if (n < 0) return 0;
  // two (key, position) array pairs, hist[kMaxBins][blocks], the digit totals of every pass
  size_t nb = (size_t)((n + kSortTile - 1) / kSortTile);
  size_t arr = ((size_t)n * 4 + 255) / 256 * 256;
  return 4 * arr + nb * kMaxBins * 4 + kMaxPasses * kMaxBins * 4 + 256;
}

// scratch for the chunk-parallel treatment of long runs, on top of tfg_sparse_scratch_bytes(n)
static size_t long_run_extra_bytes(int64_t n, int dim, size_t* max_runs, size_t* max_chunks) {
  const size_t runs = (size_t)(n / (kLongRun + 1)) + 1;
  const size_t chunks = (size_t)(n / kRunChunk) + runs + 1;
  if (max_runs) *max_runs = runs;
  if (max_chunks) *max_chunks = chunks;
  auto up = [](size_t b) { return (b + 255) / 256 * 256; };
  return 256 + up(runs * sizeof(LongRun)) + up(chunks * sizeof(ChunkDesc)) + up(chunks * (size_t)dim * sizeof(float));
}

size_t tfg_sparse_scratch_bytes_fast(int64_t n, int dim) {
  if (n < 0 || dim < 1) return 0;
  return tfg_sparse_scratch_bytes(n) + long_run_extra_bytes(n, dim, nullptr, nullptr);
}

// per CUDA device ordinal: the sort kernels may use 64 KB of dynamic shared memory
static std::atomic<bool> g_sort_attr_set[64];

static int scatter_sub_impl(float* table, int64_t vocab, int dim, int n_shards, int shard, int strategy,
                            const IdSrc& ids, int64_t n, const RowSrc& rows, float scale,
                            void* scratch, size_t scratch_bytes, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  int rc = check_sparse(vocab, dim, n_shards, shard, strategy, n);
  if (rc) return rc;
  if (n == 0) return TFG_OK;
  if (!table || !scratch) return fail(nullptr, TFG_EINVAL, "NULL argument");
  if (n >= (1ll << 31)) return fail(nullptr, TFG_ELIMIT, "more than 2^31-1 ids in one call");
  if (scratch_bytes < tfg_sparse_scratch_bytes(n)) return fail(nullptr, TFG_EINVAL, "scratch too small");
  if ((uintptr_t)scratch % 256 != 0) return fail(nullptr, TFG_EINVAL, "scratch must be 256-B aligned");
  long long srows = shard_rows(vocab, n_shards, shard, strategy);
  if (srows >= 0xffffffffll) return fail(nullptr, TFG_ELIMIT, "shard has more than 2^32-2 rows");
  Part pt = make_part(vocab, n_shards, shard, strategy);
  const uint32_t sentinel = (uint32_t)srows;  // one past the last valid row: sorts last
  const int n_blocks = (int)((n + kSortTile - 1) / kSortTile);
  size_t arr = ((size_t)n * 4 + 255) / 256 * 256;
  char* base = (char*)scratch;
  uint32_t* keys[2] = {(uint32_t*)base, (uint32_t*)(base + arr)};
  uint32_t* vals[2] = {(uint32_t*)(base + 2 * arr), (uint32_t*)(base + 3 * arr)};
  uint32_t* hist = (uint32_t*)(base + 4 * arr);
  uint32_t* totals = hist + (size_t)n_blocks * kMaxBins;
  uint32_t* n_live = nullptr;
  if (n_shards > 1) {
    // most ids belong to other shards: keep the owned ones (in order) and sort only those
    n_live = totals + (size_t)kMaxPasses * kMaxBins;          // the trailing 256 B of the scratch
    owned_keys_kernel<<<n_blocks, kSortThreads, 0, s>>>(pt, ids, n, sentinel, keys[1], hist);
    owned_scan_kernel<<<1, 1024, 0, s>>>(hist, n_blocks, n_live);
    owned_compact_kernel<<<n_blocks, kSortThreads, 0, s>>>(keys[1], n, sentinel, hist, keys[0], vals[0]);
  } else {
    int kb = (int)((n + 255) / 256);
    if (kb > kNumSMs * 16) kb = kNumSMs * 16;
    make_keys_kernel<<<kb, 256, 0, s>>>(pt, ids, n, sentinel, keys[0], vals[0]);
  }
  // the fewest passes of at most kMaxDigitBits bits that cover the key: two for up to 2^22 rows per shard
  const int key_bits = bits_for(sentinel);
  const int passes = (key_bits + kMaxDigitBits - 1) / kMaxDigitBits;
  const int bits = (key_bits + passes - 1) / passes, bins = 1 << bits;
  const size_t smem = (size_t)kSortWarps * bins * sizeof(uint32_t);
  if (smem > 48 * 1024) {
    int dev = -1;
    cudaError_t ea = cudaGetDevice(&dev);
    if (ea != cudaSuccess) return fail_cuda(nullptr, ea, "cudaGetDevice");
    if (dev < 0 || dev >= 64) return fail(nullptr, TFG_ELIMIT, "CUDA device ordinal beyond 63");
    if (!g_sort_attr_set[dev].load(std::memory_order_acquire)) {
      const int most = kSortWarps * kMaxBins * (int)sizeof(uint32_t);
      ea = cudaFuncSetAttribute(sort_hist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, most);
      if (ea == cudaSuccess) ea = cudaFuncSetAttribute(sort_scatter_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, most);
      if (ea != cudaSuccess) return fail_cuda(nullptr, ea, "cudaFuncSetAttribute(sort kernels)");
      g_sort_attr_set[dev].store(true, std::memory_order_release);
    }
  }
  cudaError_t ez = cudaMemsetAsync(totals, 0, (size_t)passes * kMaxBins * sizeof(uint32_t), s);
  if (ez != cudaSuccess) return fail_cuda(nullptr, ez, "cudaMemsetAsync(digit totals)");
  int cur = 0;
  for (int p = 0; p < passes; ++p) {
    uint32_t* tot = totals + (size_t)p * kMaxBins;
    sort_hist_kernel<<<n_blocks, kSortThreads, smem, s>>>(keys[cur], n, n_live, bits * p, bins, hist, n_blocks, tot);
    sort_scan_kernel<<<(bins + kSortWarps - 1) / kSortWarps, kSortThreads, 0, s>>>(hist, n, n_live, n_blocks, bins, tot);
    sort_scatter_kernel<<<n_blocks, kSortThreads, smem, s>>>(keys[cur], vals[cur], n, n_live, bits * p, bins, hist,
                                                             n_blocks, keys[cur ^ 1], vals[cur ^ 1]);
    cur ^= 1;
  }
  long long blocks = (n + 7) / 8;
  if (blocks > kNumSMs * 16) blocks = kNumSMs * 16;
  const size_t base_bytes = tfg_sparse_scratch_bytes(n);
  size_t max_runs = 0, max_chunks = 0;
  const size_t extra = long_run_extra_bytes(n, dim, &max_runs, &max_chunks);
  if (scratch_bytes >= base_bytes + extra) {
    // enough scratch: long runs go through the chunk-parallel kernels
    auto up = [](size_t b) { return (b + 255) / 256 * 256; };
    char* x = base + base_bytes;
    uint32_t* counters = (uint32_t*)x;                       x += 256;
    LongRun* runs = (LongRun*)x;                             x += up(max_runs * sizeof(LongRun));
    ChunkDesc* chunks = (ChunkDesc*)x;                       x += up(max_chunks * sizeof(ChunkDesc));
    float* carry = (float*)x;
    cudaError_t em = cudaMemsetAsync(counters, 0, 8, s);     // the two counters must start at zero
    if (em != cudaSuccess) return fail_cuda(nullptr, em, "cudaMemsetAsync(counters)");
    const int gl = kNumSMs * 8;
    // lanes per position: the fewest of 8 / 16 / 32 whose vectors cover a row
    const int vec = (dim & 3) == 0 ? 4 : 1;
    const int lanes = dim <= 8 * vec ? 8 : dim <= 16 * vec ? 16 : 32;
    long long hb = (n + 256 / lanes - 1) / (256 / lanes);
    if (hb > kNumSMs * 16) hb = kNumSMs * 16;
#define TFG_HEADS(V, LN)                                                                                          \
  scatter_sub_heads_kernel<V, LN><<<(int)hb, 256, 0, s>>>(table, dim, keys[cur], vals[cur], n, n_live, sentinel, rows, scale, \
                                                          runs, chunks, counters)
    if (vec == 4) {
      if (lanes == 8) TFG_HEADS(4, 8); else if (lanes == 16) TFG_HEADS(4, 16); else TFG_HEADS(4, 32);
      long_chunk_sum_kernel<4><<<gl, 256, 0, s>>>(dim, vals[cur], rows, scale, chunks, counters, carry);
      long_run_apply_kernel<4><<<gl, kApplyWarps * 32, 0, s>>>(table, dim, runs, counters, carry);
    } else {
      if (lanes == 8) TFG_HEADS(1, 8); else if (lanes == 16) TFG_HEADS(1, 16); else TFG_HEADS(1, 32);
      long_chunk_sum_kernel<1><<<gl, 256, 0, s>>>(dim, vals[cur], rows, scale, chunks, counters, carry);
      long_run_apply_kernel<1><<<gl, kApplyWarps * 32, 0, s>>>(table, dim, runs, counters, carry);
    }
#undef TFG_HEADS
  } else if ((dim & 3) == 0)
    scatter_sub_runs_kernel<4><<<(int)blocks, 256, 0, s>>>(table, dim, keys[cur], vals[cur], n, n_live, sentinel,
                                                           rows, scale);
  else
    scatter_sub_runs_kernel<1><<<(int)blocks, 256, 0, s>>>(table, dim, keys[cur], vals[cur], n, n_live, sentinel,
                                                           rows, scale);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail_cuda(nullptr, e, "sparse scatter_sub kernels");
  return TFG_OK;
}

int tfg_sparse_scatter_sub(float* table, int64_t vocab, int dim, int n_shards, int shard, int strategy,
                           const int64_t* ids, int64_t n, const float* rows, float scale,
                           void* scratch, size_t scratch_bytes, void* stream) {
  if (n > 0 && (!ids || !rows)) return fail(nullptr, TFG_EINVAL, "NULL argument");
  IdSrc is = {};
  RowSrc rs = {};
  is.base[0] = (const long long*)ids;
  rs.base[0] = rows;
  return scatter_sub_impl(table, vocab, dim, n_shards, shard, strategy, is, n, rs, scale, scratch, scratch_bytes, stream);
}

int tfg_sparse_scatter_sub_multi(float* table, int64_t vocab, int dim, int n_shards, int shard, int strategy,
                                 const int64_t* const* ids_devs, const float* const* rows_devs, int n_sources,
                                 int64_t n_per_source, float scale, void* scratch, size_t scratch_bytes, void* stream) {
  if (n_sources < 1 || n_sources > TFG_MAX_REPLICAS) return fail(nullptr, TFG_ELIMIT, "n_sources out of range");
  if (n_per_source < 0 || !ids_devs || !rows_devs) return fail(nullptr, TFG_EINVAL, "bad argument");
  if (n_per_source * n_sources >= (1ll << 31)) return fail(nullptr, TFG_ELIMIT, "more than 2^31-1 ids in one call");
  IdSrc is = {};
  RowSrc rs = {};
  for (int r = 0; r < n_sources; ++r) {
    if (n_per_source > 0 && (!ids_devs[r] || !rows_devs[r])) return fail(nullptr, TFG_EINVAL, "NULL source pointer");
    is.base[r] = (const long long*)ids_devs[r];
    rs.base[r] = rows_devs[r];
  }
  is.per = n_sources > 1 ? n_per_source : 0;
  rs.per = n_sources > 1 ? (uint32_t)n_per_source : 0;
  return scatter_sub_impl(table, vocab, dim, n_shards, shard, strategy, is, n_per_source * n_sources, rs, scale, scratch,
                          scratch_bytes, stream);
}

int tfg_sparse_gather_multi(const float* const* table_shards, int64_t vocab, int dim, int n_shards, int strategy,
                            const int64_t* ids, int64_t n, float* out, int* bad, void* stream) {
  int rc = check_sparse(vocab, dim, n_shards, 0, strategy, n);
  if (rc) return rc;
  if (n_shards > TFG_MAX_REPLICAS) return fail(nullptr, TFG_ELIMIT, "more than TFG_MAX_REPLICAS shards");
  if (n == 0) return TFG_OK;
  if (!table_shards || !ids || !out) return fail(nullptr, TFG_EINVAL, "NULL argument");
  TableSrc ts = {};
  for (int k = 0; k < n_shards; ++k) {
    if (!table_shards[k] && shard_rows(vocab, n_shards, k, strategy) > 0) return fail(nullptr, TFG_EINVAL, "NULL shard pointer");
    ts.base[k] = table_shards[k];
  }
  Part pt = make_part(vocab, n_shards, 0, strategy);
  const int lanes = gather_lanes(dim), per_block = 256 / lanes;
  long long blocks = (n + per_block - 1) / per_block;
  if (blocks > kNumSMs * 16) blocks = kNumSMs * 16;
  const long long* idp = (const long long*)ids;
  cudaStream_t st = (cudaStream_t)stream;
  if (lanes == 8) gather_multi_kernel<8><<<(int)blocks, 256, 0, st>>>(ts, pt, idp, n, dim, out, bad);
  else if (lanes == 16) gather_multi_kernel<16><<<(int)blocks, 256, 0, st>>>(ts, pt, idp, n, dim, out, bad);
  else gather_multi_kernel<32><<<(int)blocks, 256, 0, st>>>(ts, pt, idp, n, dim, out, bad);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail_cuda(nullptr, e, "gather_multi_kernel");
  return TFG_OK;
}

// CUDA-IPC handles of caller-owned device buffers (table shards, id and row staging): how the
// ranks of one box make them addressable from each other, as tfg_job_ipc_export does for a job's own.
// cudaMalloc / cudaFree behind the C ABI: a CUDA-IPC handle names a whole allocation, so a buffer that
// is to be shared must be an allocation of its own (a tensor carved out of a framework's caching
// allocator is not: its handle opens at the base of the cached segment, not at the tensor)
int tfg_dev_alloc(size_t bytes, void** dev_ptr_out) {
  if (!dev_ptr_out) return fail(nullptr, TFG_EINVAL, "NULL argument");
  cudaError_t e = cudaMalloc(dev_ptr_out, bytes ? bytes : 256);
  if (e != cudaSuccess) return fail_cuda(nullptr, e, "cudaMalloc");
  e = cudaMemset(*dev_ptr_out, 0, bytes ? bytes : 256);
  if (e != cudaSuccess) return fail_cuda(nullptr, e, "cudaMemset");
  return TFG_OK;
}
int tfg_dev_free(void* dev_ptr) {
  if (!dev_ptr) return TFG_OK;
  cudaError_t e = cudaFree(dev_ptr);
  if (e != cudaSuccess) return fail_cuda(nullptr, e, "cudaFree");
  return TFG_OK;
}

int tfg_ipc_export(const void* dev_ptr, void* handle_out) {
  if (!dev_ptr || !handle_out) return fail(nullptr, TFG_EINVAL, "NULL argument");
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, const_cast<void*>(dev_ptr));
  if (e != cudaSuccess) { fail_cuda(nullptr, e, "cudaIpcGetMemHandle"); return TFG_EPEER; }
  memcpy(handle_out, &h, sizeof h);
  return TFG_OK;
}
int tfg_ipc_open(const void* handle, void** dev_ptr_out) {
  if (!handle || !dev_ptr_out) return fail(nullptr, TFG_EINVAL, "NULL argument");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof h);
  cudaError_t e = cudaIpcOpenMemHandle(dev_ptr_out, h, cudaIpcMemLazyEnablePeerAccess);
  if (e != cudaSuccess) { fail_cuda(nullptr, e, "cudaIpcOpenMemHandle"); return TFG_EPEER; }
  return TFG_OK;
}
int tfg_ipc_close(void* dev_ptr) {
  if (!dev_ptr) return TFG_OK;
  cudaError_t e = cudaIpcCloseMemHandle(dev_ptr);
  if (e != cudaSuccess) { fail_cuda(nullptr, e, "cudaIpcCloseMemHandle"); return TFG_EPEER; }
  return TFG_OK;
}

// ---- scatter-sub without any allocation ---------------------------------------------------
// For callers that cannot cudaMalloc (see include/tfgraft.h): the sort scratch is this
// file-scope device array, part of the library image on every GPU that loads it.  In host
// code `g_static_scratch` is only a handle for the runtime, not a device address, so the
// address is fetched once per device with cudaGetSymbolAddress and that pointer is what the
// kernels receive.
static __device__ __align__(256) unsigned char g_static_scratch[TFG_SPARSE_STATIC_SCRATCH_BYTES];

// per CUDA device ordinal, filled on first use; atomic because calls may come from several host threads
// (two first calls on a device race only to store the same address)
static std::atomic<void*> g_static_scratch_addr[64];

int tfg_sparse_scatter_sub_static(float* table, int64_t vocab, int dim, int n_shards, int shard,
                                  int strategy, const int64_t* ids, int64_t n, const float* rows,
                                  float scale, void* stream) {
  if (n > TFG_SPARSE_STATIC_MAX_IDS)
    return fail(nullptr, TFG_ELIMIT,
                "tfg_sparse_scatter_sub_static: n = " + std::to_string(n) + " ids exceeds the built-in "
                "scratch (TFG_SPARSE_STATIC_MAX_IDS = " + std::to_string(TFG_SPARSE_STATIC_MAX_IDS) +
                "); split the batch or use tfg_sparse_scatter_sub with your own scratch");
  static_assert(TFG_SPARSE_STATIC_SCRATCH_BYTES % 256 == 0, "scratch size keeps 256-B alignment");
  if (tfg_sparse_scratch_bytes(TFG_SPARSE_STATIC_MAX_IDS) > TFG_SPARSE_STATIC_SCRATCH_BYTES)
    return fail(nullptr, TFG_ESTATE, "built-in scratch is smaller than the sort needs");
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return fail_cuda(nullptr, e, "cudaGetDevice");
  if (dev < 0 || dev >= 64) return fail(nullptr, TFG_ELIMIT, "CUDA device ordinal beyond 63");
  void* scratch = g_static_scratch_addr[dev].load(std::memory_order_acquire);
  if (!scratch) {  // once per device, with that device current
    e = cudaGetSymbolAddress(&scratch, g_static_scratch);
    if (e != cudaSuccess) return fail_cuda(nullptr, e, "cudaGetSymbolAddress(g_static_scratch)");
    g_static_scratch_addr[dev].store(scratch, std::memory_order_release);
  }
  // Hand over exactly the sort scratch this n needs: every duplicate is then applied in order,
  // whatever n is (with more scratch tfg_sparse_scatter_sub would switch long runs to its
  // chunk-parallel path, which rounds differently).
  return tfg_sparse_scatter_sub(table, vocab, dim, n_shards, shard, strategy, ids, n, rows, scale, scratch,
                                tfg_sparse_scratch_bytes(n), stream);
}

}  // extern "C"
