// Shared between model_mnist.cu (CUDA-core kernels, orchestration) and gemm_tc.cu
// (tcgen05 tensor path).
#pragma once
#include <cuda_bf16.h>

#include <vector>

#include "common.cuh"

namespace tfg {

constexpr int kSplitW2 = 32;   // split-K factor of the fp32 conv2 weight-gradient GEMM

template <typename T> __device__ __forceinline__ T cvt(float v);
template <> __device__ __forceinline__ float cvt<float>(float v) { return v; }
template <> __device__ __forceinline__ __nv_bfloat16 cvt<__nv_bfloat16>(float v) {
  return __float2bfloat16_rn(v);
}
__device__ __forceinline__ void store4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
__device__ __forceinline__ void store4(__nv_bfloat16* p, float4 v) {
  __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
  uint2 u;
  u.x = *reinterpret_cast<uint32_t*>(&lo);
  u.y = *reinterpret_cast<uint32_t*>(&hi);
  *reinterpret_cast<uint2*>(p) = u;
}

struct MnistSlot {       // activations of one local replica
  float* x_stage = nullptr;      // [B,784] staging for host inputs
  int32_t* lab_stage = nullptr;
  float* a1 = nullptr;           // [B,14,14,32] pooled relu(conv1)
  uint8_t* idx1 = nullptr;
  float* col = nullptr;          // [B*196, 800] im2col of a1 (fp32 math mode)
  float* y2 = nullptr;           // [B*196, 64] conv2 + bias
  float* a2 = nullptr;           // [B,7,7,64] = [B,3136]
  uint8_t* idx2 = nullptr;
  float* h1 = nullptr;           // [B,1024]
  float* logits = nullptr;       // [B,10]
  float* dlogits = nullptr;
  float* loss_vec = nullptr;
  unsigned int* head_done = nullptr;   // blocks of head_fwd_kernel that have finished (the last one takes the mean loss)
  float* dh1 = nullptr;
  float* da2 = nullptr;
  float* da2m = nullptr;
  float* dy2 = nullptr;          // [B*196, 64]
  float* dcol = nullptr;         // [B*196, 800]
  float* g1m = nullptr;          // [B,14,14,32]
  float* scratch = nullptr;
  float* scratch_c1 = nullptr;           // per-image partials of the conv1 gradients
  size_t scratch_floats = 0;
  void* tc = nullptr;            // tensor-path buffers (gemm_tc.cu)
};

struct MnistModel {
  int B = 0;
  int math = TFG_MATH_FP32;
  int var[8] = {};
  std::vector<MnistSlot> slots;
  void* tc = nullptr;            // tensor-path shared state (weight copies, tensor maps)
};

// model_mnist.cu: y2 -> a2, idx2 (shared by both math paths)
int mnist_pool2(tfg_job* j, int slot, void* a2_bf16, float* a2_lo);
// out[c] = sum over the R rows of in[R, C], rows added in a fixed order (reduce_rows_kernel)
int mnist_reduce_rows(tfg_job* j, const float* in, int R, int C, float* out, cudaStream_t st);

// gemm_tc.cu
int mnist_tc_build(tfg_job* j);
void mnist_tc_destroy(tfg_job* j);
int mnist_tc_forward(tfg_job* j, int slot);    // a1 -> y2/a2/idx2 -> h1
int mnist_tc_backward(tfg_job* j, int slot);   // dh1 -> dW3, da2, dy2, dW2, dcol
void* mnist_tc_dh1_bf16(tfg_job* j, int slot); // where fc2_bwd drops its bf16 copy of dh1 (NULL unless bf16 math)
float* mnist_tc_dh1_lo(tfg_job* j, int slot);
// 3xTF32: fc1's K-split partial sums [splits][B][1024], which head_fwd_kernel adds (NULL otherwise: h1 is complete)
const float* mnist_tc_fc1_partial(tfg_job* j, int slot, int* splits);
float* mnist_tc_a1_lo(tfg_job* j, int slot);   // where conv1 drops the low-order plane of a1 (NULL unless 3xTF32 math)  // ... and the low-order plane of dh1 (NULL unless 3xTF32 math)

}  // namespace tfg
