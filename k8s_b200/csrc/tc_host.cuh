// Host-side interface of the tcgen05 GEMM (gemm_tc.cu) for the model translation units.
#pragma once
#include "common.cuh"
#include "gemm_tc.cuh"

namespace tfg {

struct TcGemm {
  CUtensorMap a, b;
  CUtensorMap a_lo, b_lo;      // x3: maps of the operands' low-order planes (same geometry as a / b)
  bool x3 = false;             // 3xTF32 split product (gemm_x3.cuh; kind tf32 only): fp32-equivalent result
  int chunk_kb = 0;            // x3: K blocks accumulated inside the tensor core per chunk (0 = default 2)
  CUtensorMap c;               // bf16 output [M, ldc], box {64, 128} (make_map(.., M, N, ldc, 128)); used when tma_out != 0
  int tma_out = 0;             // 1: TMA store, 2: TMA reduce-add into the output; forces the persistent kernel
  float* C = nullptr;          // [M, ldc]
  float* partial = nullptr;    // [splits][M][N] when splits > 1
  int M = 0, N = 0, K = 0, ldc = 0, block_n = 128, splits = 1;
  bool a_mn = false, b_mn = false;
  bool reduce_later = false;   // splits > 1: leave the partial sums [splits][M][N] for the consumer kernel to add (tc_splits() of them)
  tc::Epilogue epi = {nullptr, nullptr, 0, nullptr, 0, 0};
  const char* name = "gemm";
};


// Operand with `rows` (M or N) and `cols` (K).  K-major: stored [rows, K], pitch ld, box = 128 B of
// K x box_rows.  MN-major: stored [K, rows], pitch ld, box = 128 B of rows x one stage of K.
int make_map(CUtensorMap* map, int kind, const void* ptr, int64_t rows, int64_t cols, int64_t ld, int box_rows,
             bool mn_major = false);
// j may be NULL (stateless entry point).  Launches the GEMM (and the split-K reduce when needed).
int launch_tc(tfg_job* j, const TcGemm& g, int kind, cudaStream_t s);
// number of K splits launch_tc really uses for g (no split is left empty)
int tc_splits(const TcGemm& g, int kind);

}  // namespace tfg
