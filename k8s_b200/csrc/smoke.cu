// The tf_sample smoke workload: per cluster task, c = a * b with
// a = b = arange(100).reshape(10, 10), int32
// (examples/tf_sample/tf_sample/tf_smoke.py:57-59).
#include "common.cuh"

namespace tfg {

__global__ void smoke_mul_kernel(int32_t* __restrict__ out, int n_tasks) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_tasks * 100) return;
  int32_t a = i % 100;  // tf.constant(range(width * height), shape=[height, width])
  int32_t b = i % 100;
  out[i] = a * b;       // tf.multiply(a, b)
}

int launch_smoke(int32_t* out_dev, int n_tasks, cudaStream_t s) {
  int n = n_tasks * 100;
  smoke_mul_kernel<<<(n + 127) / 128, 128, 0, s>>>(out_dev, n_tasks);
  return cudaGetLastError() == cudaSuccess ? TFG_OK : TFG_ECUDA;
}

}  // namespace tfg
