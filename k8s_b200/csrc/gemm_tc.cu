// Tensor path of the sample model -- placeholder until the tcgen05 GEMM lands.
#include "common.cuh"
#include "model_mnist.cuh"

namespace tfg {
int mnist_tc_build(tfg_job* j) { return fail(j, TFG_ESTATE, "tensor math modes are not built yet"); }
void mnist_tc_destroy(tfg_job*) {}
int mnist_tc_forward(tfg_job* j, int) { return fail(j, TFG_ESTATE, "tensor math modes are not built yet"); }
int mnist_tc_backward(tfg_job* j, int) { return fail(j, TFG_ESTATE, "tensor math modes are not built yet"); }
}  // namespace tfg
