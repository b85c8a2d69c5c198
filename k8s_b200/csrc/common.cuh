// Internal declarations shared by the libtfgraft translation units.
#pragma once
#include <cuda_runtime.h>
#include <nvtx3/nvToolsExt.h>
#include <stdint.h>
#include <stdio.h>
#include <string>
#include <utility>
#include <vector>

#include "../../include/tfgraft.h"

namespace tfg {

constexpr int kPadFloats = 32;          // every variable starts on a 128-B boundary
constexpr int kNumSMs = 148;            // B200
// rows of a rank's flag block (each TFG_MAX_GPUS words, written by the peers): the standalone barrier,
// and the two synchronisation points folded into the push/apply/pull kernel
constexpr int kFlagsBarrier = 0, kFlagsReady = 1, kFlagsPulled = 2, kFlagRows = 3;

void set_thread_error(const std::string& s);
const char* thread_error();

struct VarInfo {
  std::string name, op_type;
  int dtype = TFG_DT_F32;
  std::vector<int64_t> shape;
  bool trainable = true;
  int ps_task = -1;      // /job:ps/task:k, -1 = worker-local (no ps job)
  int64_t count = 0;     // elements
  int64_t offset = -1;   // element offset in the flat fp32 space, -1 = not in it
  int64_t padded = 0;    // count rounded up to kPadFloats
};

struct Segment {         // a range of the flat space whose update this rank applies
  int64_t off;           // element offset (multiple of 4)
  int64_t cnt;           // elements (multiple of 4)
  int64_t chunk0;        // index of the segment's first work chunk
};

struct MnistModel;       // model_mnist.cu
struct ResnetModel;      // resnet.cu

}  // namespace tfg

struct tfg_job {
  // cluster (jobs sorted by name, as Go marshals the ClusterSpec map)
  std::vector<std::pair<std::string, std::vector<std::string>>> cluster;
  int gpu_rank = 0, n_gpus = 1;
  int n_ps = 0, n_replicas = 1;
  std::vector<int> local_replicas;     // global replica ids hosted here, ascending
  // variables
  std::vector<tfg::VarInfo> vars;
  int rr_next = 0;
  bool finalized = false;
  int layout = TFG_LAYOUT_LOGICAL;
  int64_t flat_count = 0;
  std::vector<tfg::Segment> segs;
  int64_t n_chunks = 0;
  int chunk_elems = 4096;              // elements per work chunk of the apply kernel (1024 for small shards)
  // device state
  int device = 0;
  bool bound = false, own_stream = false;
  cudaStream_t stream = nullptr;
  float* params = nullptr;
  float* accum = nullptr;
  void* params_lp = nullptr;           // operand-precision shadow of `params`, kept current by the apply kernel
  int lp_mode = 0;                     // 0 none, 1 = bf16 copy [flat], 2 = fp32 low-order plane of the 3xTF32 split [flat]
  std::vector<float*> grads;           // one flat buffer per local replica slot
  float* loss = nullptr;               // [TFG_MAX_REPLICAS + 2]; [r] = replica r, then the mean, then the error word (dev_err)
  uint32_t* flags = nullptr;           // [TFG_MAX_GPUS] barrier epochs written by peers
  int* dev_err = nullptr;              // device-side error word (barrier timeout)
  tfg::Segment* d_segs = nullptr;
  float* h_loss = nullptr;             // pinned
  int* h_err = nullptr;                // pinned
  // peers (index = GPU rank / global replica id); own entries point at own buffers
  float* peer_params[TFG_MAX_GPUS] = {};
  void* peer_params_lp[TFG_MAX_GPUS] = {};
  float* peer_loss[TFG_MAX_GPUS] = {};
  uint32_t* peer_flags[TFG_MAX_GPUS] = {};
  float* peer_grads[TFG_MAX_REPLICAS] = {};
  std::vector<void*> ipc_opened;
  uint32_t* dev_epoch = nullptr;       // [0] barrier epoch (advanced by the barrier kernel), [1] epoch and [2] finished-CTA
                                       // counter of the push/apply/pull kernel
  // optimizer
  int opt_kind = TFG_OPT_SGD;
  float lr = 0.01f, mu = 0.9f;
  int64_t global_step = 0;
  // model
  tfg::MnistModel* model = nullptr;
  tfg::ResnetModel* resnet = nullptr;
  int64_t launches = 0;
  // Side branch of the step: kernels that are leaves of the dependency graph (the ResNet weight
  // gradients) run on side_stream between side_fork() and side_join(); inside the captured step they
  // become a parallel branch of the graph.  With per-kernel timing on, the side stream IS the main one.
  cudaStream_t side_stream = nullptr;
  cudaEvent_t ev_side[2] = {nullptr, nullptr};
  bool side_open = false;
  // optional per-kernel timing: one event after every launch (see prof_mark)
  bool prof_on = false;
  std::vector<std::pair<const char*, cudaEvent_t>> prof_ev;
  size_t prof_used = 0;
  // CUDA graphs of the whole step, keyed by the input pointers they were captured with
  struct StepGraph {
    std::vector<const void*> key;
    cudaGraphExec_t exec = nullptr;
    int64_t launches = 0;
    int calls = 0;
  };
  std::vector<StepGraph> step_graphs;
  bool use_graphs = true;
  std::string err;
};

namespace tfg {

// NVTX range over a phase of the step (forward, backward, push+apply+pull): visible in Nsight Systems /
// Compute timelines, a no-op without a profiler attached (header-only NVTX 3)
struct NvtxRange {
  explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
  ~NvtxRange() { nvtxRangePop(); }
  NvtxRange(const NvtxRange&) = delete;
  NvtxRange& operator=(const NvtxRange&) = delete;
};

int fail(tfg_job* j, int rc, const std::string& msg);
int fail_cuda(tfg_job* j, cudaError_t e, const char* what);

#define TFG_CUDA(j, call)                                         \
  do {                                                            \
    cudaError_t e__ = (call);                                     \
    if (e__ != cudaSuccess) return tfg::fail_cuda((j), e__, #call); \
  } while (0)

// Count a launch and, when profiling is on, drop an event behind it.  The job's stream is
// serial, so the span between consecutive events is the kernel's duration plus any gap in
// front of it.
void prof_mark(tfg_job* j, const char* what);
// side branch (see tfg_job::side_stream): side_fork makes the side stream wait for everything
// enqueued on j->stream so far and returns the stream to launch on; side_join makes j->stream wait
// for the side branch (no-op when none is open)
cudaStream_t side_fork(tfg_job* j);
void side_join(tfg_job* j);

// apply.cu
int launch_job_apply(tfg_job* j, bool summed_local, int n_summed, bool whole = false, bool sync = false);
int launch_barrier(tfg_job* j);
int launch_dense_apply(int kind, float* var, float* accum, const float* const* grads, int n_grads,
                       float divisor, int64_t count, float lr, float mu, float* const* pull,
                       int n_pull, cudaStream_t s);
// smoke.cu
int launch_smoke(int32_t* out_dev, int n_tasks, cudaStream_t s);
// model_mnist.cu
int mnist_build(tfg_job* j, int batch, int math);
void mnist_destroy(tfg_job* j);
int mnist_forward_backward(tfg_job* j, int slot, const float* images, const int32_t* labels);
int mnist_logits(tfg_job* j, int slot, float* host, size_t bytes);
// resnet.cu
int resnet_build(tfg_job* j, int batch, int image_hw, int classes);
void resnet_destroy(tfg_job* j);
int resnet_forward_backward(tfg_job* j, int slot, const float* images, const int32_t* labels);
int resnet_logits(tfg_job* j, int slot, float* host, size_t bytes);
int resnet_batch(const tfg_job* j);
size_t resnet_image_floats(const tfg_job* j);
float* resnet_stage_images(tfg_job* j, int slot);
int32_t* resnet_stage_labels(tfg_job* j, int slot);
int resnet_tensor_get(tfg_job* j, int slot, const char* name, float* host, size_t cap_floats, size_t* n_out);
int mnist_batch(const tfg_job* j);
int mnist_math(const tfg_job* j);
int launch_cast_lp(const float* src, void* dst, long long n, int lp_mode, cudaStream_t s);   // apply.cu
// 3xTF32 on several GPUs: the exchange moves the fp32 parameters only (8 (N-1)/N instead of 12 (N-1)/N bytes per
// parameter and NVLink direction); every rank derives the low-order plane from its own copy at the start of the next
// forward, beside conv1, which does not read it (model_mnist.cu).
inline bool lp_derived_per_step(const tfg_job* j) { return j->lp_mode == 2 && j->n_gpus > 1; }
float* mnist_stage_images(tfg_job* j, int slot);
int32_t* mnist_stage_labels(tfg_job* j, int slot);

}  // namespace tfg
