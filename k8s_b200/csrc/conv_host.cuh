// Host-side interface of the implicit-GEMM convolutions (conv_tc.cu).
#pragma once
#include "common.cuh"
#include "conv_tc.cuh"

namespace tfg {

struct ConvGemm {              // forward (b_mn) or stride-1 data gradient (!b_mn)
  CUtensorMap a, b;
  CUtensorMap c;               // bf16 output [Nimg, OH, OW, N], box {64, OW, TH, TN} (make_map_4d, stride 1); used when tma_out != 0
  int tma_out = 0;             // 1: TMA store, 2: TMA reduce-add into the output
  tc::ConvShape cs;
  float* C = nullptr;          // fp32 output, or NULL when epi.out_bf16 is set
  int N = 0, ldc = 0, block_n = 128, b_rows_per_tap = 0;
  bool b_mn = true;
  tc::Epilogue epi = {nullptr, nullptr, 0, nullptr, 1, 0};
  const char* name = "conv";
};

struct HaloConv {              // 3x3 stride-1 forward (b_mn) or data gradient (!b_mn), input tile resident in smem
  CUtensorMap a, b, c;         // a: box {64, OW + 2, TH + 2, 1} of the gathered tensor; c: box {64, OW + 2, TH, 1} of the output
  tc::HaloShape hs;
  int N = 0, block_n = 128, b_rows_per_tap = 0, tma_out = 1;
  bool b_mn = true;
  tc::Epilogue epi = {nullptr, nullptr, 0, nullptr, 1, 1, nullptr};
  const char* name = "conv halo";
};

struct ConvWgrad {
  CUtensorMap x, dz;
  tc::ConvShape cs;
  float* C = nullptr;          // [M, N] fp32
  float* partial = nullptr;    // [splits][M][N]
  int M = 0, N = 0, block_n = 128, splits = 1;
  const char* name = "conv wgrad";
};

int make_map_4d(CUtensorMap* map, const void* ptr, int N, int H, int W, int C, int bw, int bh, int bn, int stride);
void pick_conv_tile(int Nimg, int OH, int OW, bool rows_mult16, int* TN, int* TH);
void fill_conv_shape(tc::ConvShape* cs, int Nimg, int OH, int OW, int TN, int TH, int R, int S, int stride, int pad,
                     int C, int flip);
int fill_dgrad_s2_class(tc::ConvShape* cs, int R, int ph, int pw);
int launch_conv_gemm(tfg_job* j, const ConvGemm& g, cudaStream_t s);
int launch_conv_wgrad(tfg_job* j, const ConvWgrad& g, cudaStream_t s);
// 4-D map with an explicit box {64, bw, bh, 1} (element strides 1); bw may exceed W (clipped / zero-filled)
int make_map_4d_box(CUtensorMap* map, const void* ptr, int N, int H, int W, int C, int bw, int bh);
bool halo_fits(int OH, int OW, int C, int R, int stride);   // can conv3x3_halo_kernel take this layer?
int halo_mt(int Nimg, int OH, int OW, int tiles_n);           // M tiles per CTA (1 or 2) for this layer
void fill_halo_shape(tc::HaloShape* hs, int Nimg, int OH, int OW, int C, int flip, int tiles_n);
int halo_grid_rows(const tc::HaloShape& hs, int tiles_n);    // CTA rows of the launch (column-sum records / parts)
int launch_conv_halo(tfg_job* j, const HaloConv& g, cudaStream_t s);

}  // namespace tfg
