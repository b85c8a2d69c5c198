// Job lifecycle, cluster bookkeeping, variable placement, peer mapping and step
// orchestration of libtfgraft.  Reference citations are in include/tfgraft.h.
#include <algorithm>
#include <cstring>
#include <new>

#include "common.cuh"
#include "json_min.hpp"

namespace tfg {

static thread_local std::string g_thread_err;
void set_thread_error(const std::string& s) { g_thread_err = s; }
const char* thread_error() { return g_thread_err.c_str(); }

int fail(tfg_job* j, int rc, const std::string& msg) {
  if (j) j->err = msg;
  g_thread_err = msg;
  return rc;
}

int fail_cuda(tfg_job* j, cudaError_t e, const char* what) {
  std::string m = std::string("CUDA error ") + cudaGetErrorName(e) + " (" + cudaGetErrorString(e) +
                  ") in " + what;
  int rc = (e == cudaErrorMemoryAllocation) ? TFG_ENOMEM : TFG_ECUDA;
  return fail(j, rc, m);
}

void prof_mark(tfg_job* j, const char* what) {
  j->launches++;
  if (!j->prof_on) return;
  if (j->prof_used == j->prof_ev.size()) {
    cudaEvent_t e;
    if (cudaEventCreate(&e) != cudaSuccess) { j->prof_on = false; return; }
    j->prof_ev.emplace_back(what, e);
  }
  j->prof_ev[j->prof_used].first = what;
  cudaEventRecord(j->prof_ev[j->prof_used].second, j->stream);
  j->prof_used++;
}

cudaStream_t side_fork(tfg_job* j) {
  if (j->prof_on || !j->side_stream) return j->stream;   // per-kernel timing attributes time on one stream
  if (cudaEventRecord(j->ev_side[0], j->stream) != cudaSuccess ||
      cudaStreamWaitEvent(j->side_stream, j->ev_side[0], 0) != cudaSuccess) {
    cudaGetLastError();
    return j->stream;
  }
  j->side_open = true;
  return j->side_stream;
}

void side_join(tfg_job* j) {
  if (!j->side_open) return;
  j->side_open = false;
  cudaEventRecord(j->ev_side[1], j->side_stream);
  cudaStreamWaitEvent(j->stream, j->ev_side[1], 0);
}

// [TF1.3] replica_device_setter default ps_ops (python/training/device_setter.py).
static bool is_ps_op(const std::string& t) {
  return t == "Variable" || t == "VariableV2" || t == "VarHandleOp";
}

static int64_t round_up(int64_t v, int64_t m) { return (v + m - 1) / m * m; }

static int ensure_bound(tfg_job* j) {
  if (j->bound) {
    cudaError_t e = cudaSetDevice(j->device);
    if (e != cudaSuccess) return fail_cuda(j, e, "cudaSetDevice");
    return TFG_OK;
  }
  return tfg_job_bind(j, j->gpu_rank, nullptr) == TFG_OK ? TFG_OK : TFG_ECUDA;
}

// owned ranges of this rank inside [lo, hi) of the flat space, cut into work chunks
static std::vector<Segment> segments_in(tfg_job* j, int64_t lo_lim, int64_t hi_lim, int64_t kChunk, int64_t* n_chunks) {
  std::vector<Segment> segs;
  auto add = [&](int64_t off, int64_t cnt) {
    const int64_t lo = std::max(off, lo_lim), hi = std::min(off + cnt, hi_lim);
    if (hi <= lo) return;
    if (!segs.empty() && segs.back().off + segs.back().cnt == lo && (segs.back().cnt % kChunk) == 0) {
      segs.back().cnt += hi - lo;  // merge only when chunk boundaries stay aligned
    } else {
      segs.push_back({lo, hi - lo, 0});
    }
  };
  if (j->layout == TFG_LAYOUT_BALANCED) {
    int64_t part = round_up((j->flat_count + j->n_gpus - 1) / j->n_gpus, kPadFloats);
    int64_t lo = std::min<int64_t>(part * j->gpu_rank, j->flat_count);
    int64_t hi = std::min<int64_t>(lo + part, j->flat_count);
    add(lo, hi - lo);
  } else {
    for (auto& v : j->vars) {
      if (v.offset < 0) continue;
      int owner = v.ps_task < 0 ? 0 : v.ps_task % j->n_gpus;
      if (owner == j->gpu_rank) add(v.offset, v.padded);
    }
  }
  int64_t chunk = 0;
  for (auto& sg : segs) {
    sg.chunk0 = chunk;
    chunk += (sg.cnt + kChunk - 1) / kChunk;
  }
  *n_chunks = chunk;
  return segs;
}

static void build_segments(tfg_job* j) {
  // elements per work chunk of the apply kernel: small shards take small chunks so that the kernel
  // still has a few CTAs per SM (the MNIST model on 8 ranks: 409 k parameters per rank)
  int64_t owned = 0;
  if (j->layout == TFG_LAYOUT_BALANCED) {
    owned = (j->flat_count + j->n_gpus - 1) / j->n_gpus;
  } else {
    for (auto& v : j->vars)
      if (v.offset >= 0 && (v.ps_task < 0 ? 0 : v.ps_task % j->n_gpus) == j->gpu_rank) owned += v.padded;
  }
  // multi-GPU ranks always use 1024-element chunks: their exchange kernel streams 4-KB pieces through
  // shared memory (apply.cu, push_apply_pull_tma_kernel)
  const int64_t kChunk = (owned >= 4096ll * 2 * kNumSMs && j->n_gpus == 1) ? 4096 : 1024;
  j->chunk_elems = (int)kChunk;
  j->segs = segments_in(j, 0, j->flat_count, kChunk, &j->n_chunks);
}

}  // namespace tfg

using namespace tfg;

extern "C" {

int tfg_version(void) { return 100; }

int tfg_device_count(int* n_out) {
  if (!n_out) return fail(nullptr, TFG_EINVAL, "n_out is NULL");
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess) {
    *n_out = 0;
    cudaGetLastError();
    return fail_cuda(nullptr, e, "cudaGetDeviceCount");
  }
  *n_out = n;
  return TFG_OK;
}

const char* tfg_last_error(const tfg_job* job) { return job ? job->err.c_str() : thread_error(); }

int tfg_job_create(const char* tf_config_json, int gpu_rank, int n_gpus, tfg_job** out) {
  if (!out) return fail(nullptr, TFG_EINVAL, "out is NULL");
  *out = nullptr;
  if (n_gpus < 1 || n_gpus > TFG_MAX_GPUS)
    return fail(nullptr, TFG_EINVAL, "n_gpus must be in [1, " + std::to_string(TFG_MAX_GPUS) + "]");
  if (gpu_rank < 0 || gpu_rank >= n_gpus) return fail(nullptr, TFG_EINVAL, "gpu_rank out of range");
  json::Value root;
  std::string perr;
  const char* text = (tf_config_json && *tf_config_json) ? tf_config_json : "{}";
  if (!json::Parser(text).parse(&root, &perr)) return fail(nullptr, TFG_ESPEC, "TF_CONFIG: " + perr);
  if (root.kind != json::Value::Obj) return fail(nullptr, TFG_ESPEC, "TF_CONFIG: not a JSON object");
  tfg_job* j = new (std::nothrow) tfg_job();
  if (!j) return fail(nullptr, TFG_ENOMEM, "out of host memory");
  j->gpu_rank = gpu_rank;
  j->n_gpus = n_gpus;
  const json::Value* cl = root.get("cluster");
  if (cl && cl->kind == json::Value::Obj) {
    for (auto& kv : cl->obj) {
      if (kv.second.kind != json::Value::Arr) {
        delete j;
        return fail(nullptr, TFG_ESPEC, "TF_CONFIG: cluster." + kv.first + " is not a list");
      }
      std::vector<std::string> tasks;
      for (auto& t : kv.second.arr) {
        if (t.kind != json::Value::Str) {
          delete j;
          return fail(nullptr, TFG_ESPEC, "TF_CONFIG: cluster." + kv.first + " holds a non-string task");
        }
        tasks.push_back(t.str);
      }
      j->cluster.emplace_back(kv.first, std::move(tasks));
    }
    std::stable_sort(j->cluster.begin(), j->cluster.end(),
                     [](const auto& a, const auto& b) { return a.first < b.first; });
  } else if (cl && cl->kind != json::Value::Null) {
    delete j;
    return fail(nullptr, TFG_ESPEC, "TF_CONFIG: cluster is not an object");
  }
  int n_master = tfg_cluster_num_tasks(j, "master"), n_worker = tfg_cluster_num_tasks(j, "worker");
  j->n_ps = std::max(0, tfg_cluster_num_tasks(j, "ps"));
  j->n_replicas = std::max(0, n_master) + std::max(0, n_worker);
  if (j->n_replicas == 0) j->n_replicas = 1;  // empty TF_CONFIG: the local, single-replica case
  if (j->n_replicas > TFG_MAX_REPLICAS) {
    delete j;
    return fail(nullptr, TFG_ELIMIT, "more than TFG_MAX_REPLICAS training replicas");
  }
  for (int r = 0; r < j->n_replicas; ++r)
    if (r % n_gpus == gpu_rank) j->local_replicas.push_back(r);
  *out = j;
  return TFG_OK;
}

int tfg_job_destroy(tfg_job* j) {
  if (!j) return TFG_OK;
  if (j->bound) {
    cudaSetDevice(j->device);
    cudaStreamSynchronize(j->stream);
    mnist_destroy(j);
    resnet_destroy(j);
    for (auto& pe : j->prof_ev) cudaEventDestroy(pe.second);
    for (void* p : j->ipc_opened) cudaIpcCloseMemHandle(p);
    cudaFree(j->params);
    cudaFree(j->accum);
    cudaFree(j->params_lp);
    for (float* g : j->grads) cudaFree(g);
    cudaFree(j->loss);
    cudaFree(j->flags);
    cudaFree(j->dev_epoch);
    for (auto& g : j->step_graphs)
      if (g.exec) cudaGraphExecDestroy(g.exec);
    cudaFree(j->d_segs);
    if (j->h_loss) cudaFreeHost(j->h_loss);
    if (j->side_stream) cudaStreamDestroy(j->side_stream);
    for (cudaEvent_t e : j->ev_side)
      if (e) cudaEventDestroy(e);
    if (j->own_stream && j->stream) cudaStreamDestroy(j->stream);
    cudaGetLastError();
  }
  delete j;
  return TFG_OK;
}

int tfg_job_bind(tfg_job* j, int cuda_device, void* stream) {
  if (!j) return fail(nullptr, TFG_EINVAL, "job is NULL");
  if (j->bound) return fail(j, TFG_ESTATE, "job is already bound to a device");
  TFG_CUDA(j, cudaSetDevice(cuda_device));
  j->device = cuda_device;
  if (stream) {
    j->stream = (cudaStream_t)stream;
  } else {
    TFG_CUDA(j, cudaStreamCreateWithFlags(&j->stream, cudaStreamNonBlocking));
    j->own_stream = true;
  }
  TFG_CUDA(j, cudaStreamCreateWithFlags(&j->side_stream, cudaStreamNonBlocking));
  for (cudaEvent_t& e : j->ev_side) TFG_CUDA(j, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  // [replica losses | mean loss | error word]: the last two words are read back by ONE 8-byte copy per step
  TFG_CUDA(j, cudaMalloc(&j->loss, sizeof(float) * (TFG_MAX_REPLICAS + 2)));
  TFG_CUDA(j, cudaMemset(j->loss, 0, sizeof(float) * (TFG_MAX_REPLICAS + 2)));
  j->dev_err = reinterpret_cast<int*>(j->loss + TFG_MAX_REPLICAS + 1);
  TFG_CUDA(j, cudaMalloc(&j->flags, sizeof(uint32_t) * TFG_MAX_GPUS * kFlagRows));
  TFG_CUDA(j, cudaMemset(j->flags, 0, sizeof(uint32_t) * TFG_MAX_GPUS * kFlagRows));
  TFG_CUDA(j, cudaMalloc(&j->dev_epoch, sizeof(uint32_t) * 4));
  TFG_CUDA(j, cudaMemset(j->dev_epoch, 0, sizeof(uint32_t) * 4));
  if (const char* g = getenv("TFGRAFT_GRAPH")) j->use_graphs = atoi(g) != 0;
  TFG_CUDA(j, cudaMallocHost(&j->h_loss, sizeof(float) * (TFG_MAX_REPLICAS + 2)));   // same layout, pinned
  j->h_err = reinterpret_cast<int*>(j->h_loss + TFG_MAX_REPLICAS + 1);
  *j->h_err = 0;
  j->peer_loss[j->gpu_rank] = j->loss;
  j->peer_flags[j->gpu_rank] = j->flags;
  j->bound = true;
  return TFG_OK;
}

int tfg_cluster_num_tasks(const tfg_job* j, const char* job_name) {
  if (!j || !job_name) return -1;
  for (auto& kv : j->cluster)
    if (kv.first == job_name) return (int)kv.second.size();
  return -1;
}
int tfg_num_replicas(const tfg_job* j) { return j ? j->n_replicas : -1; }
int tfg_num_ps(const tfg_job* j) { return j ? j->n_ps : -1; }
int tfg_replica_gpu(const tfg_job* j, int r) {
  return (!j || r < 0 || r >= j->n_replicas) ? -1 : r % j->n_gpus;
}
int tfg_ps_gpu(const tfg_job* j, int k) { return (!j || k < 0 || k >= j->n_ps) ? -1 : k % j->n_gpus; }
int tfg_num_local_replicas(const tfg_job* j) { return j ? (int)j->local_replicas.size() : -1; }
int tfg_local_replica(const tfg_job* j, int slot) {
  return (!j || slot < 0 || slot >= (int)j->local_replicas.size()) ? -1 : j->local_replicas[slot];
}

int tfg_parse_cluster_spec(const char* cluster_spec, char* json_out, size_t cap) {
  if (!cluster_spec || !json_out || cap == 0) return fail(nullptr, TFG_EINVAL, "NULL argument");
  std::string s(cluster_spec);
  if (s.empty()) return fail(nullptr, TFG_ESPEC, "Empty cluster_spec string");
  std::vector<std::pair<std::string, std::vector<std::string>>> jobs;
  size_t pos = 0;
  for (;;) {
    size_t comma = s.find(',', pos);
    std::string js = s.substr(pos, comma == std::string::npos ? std::string::npos : comma - pos);
    if (std::count(js.begin(), js.end(), '|') != 1)
      return fail(nullptr, TFG_ESPEC, "Not exactly one instance of '|' in cluster_spec");
    size_t bar = js.find('|');
    std::string name = js.substr(0, bar);
    if (name.empty()) return fail(nullptr, TFG_ESPEC, "Empty job_name in cluster_spec");
    std::vector<std::string> tasks;
    std::string rest = js.substr(bar + 1);
    size_t tp = 0;
    for (int i = 0;; ++i) {
      size_t semi = rest.find(';', tp);
      std::string t = rest.substr(tp, semi == std::string::npos ? std::string::npos : semi - tp);
      if (t.empty())
        return fail(nullptr, TFG_ESPEC, "Empty task string at position " + std::to_string(i));
      tasks.push_back(t);
      if (semi == std::string::npos) break;
      tp = semi + 1;
    }
    bool merged = false;
    for (auto& kv : jobs)
      if (kv.first == name) { kv.second = tasks; merged = true; }
    if (!merged) jobs.emplace_back(name, tasks);
    if (comma == std::string::npos) break;
    pos = comma + 1;
  }
  std::string out = "{";
  for (size_t a = 0; a < jobs.size(); ++a) {
    if (a) out += ",";
    json::escape(jobs[a].first, &out);
    out += ":[";
    for (size_t b = 0; b < jobs[a].second.size(); ++b) {
      if (b) out += ",";
      json::escape(jobs[a].second[b], &out);
    }
    out += "]";
  }
  out += "}";
  if (out.size() + 1 > cap) return fail(nullptr, TFG_ELIMIT, "json_out too small");
  memcpy(json_out, out.c_str(), out.size() + 1);
  return TFG_OK;
}

int tfg_var_register(tfg_job* j, const char* name, const char* op_type, int dtype,
                     const int64_t* shape, int rank, int trainable, int* var_out, int* ps_task_out) {
  if (!j) return fail(nullptr, TFG_EINVAL, "job is NULL");
  if (j->finalized) return fail(j, TFG_ESTATE, "variables are finalized");
  if (!name || !op_type || rank < 0 || (rank > 0 && !shape)) return fail(j, TFG_EINVAL, "bad argument");
  if (dtype != TFG_DT_F32 && dtype != TFG_DT_I64) return fail(j, TFG_EINVAL, "unknown dtype");
  VarInfo v;
  v.name = name;
  v.op_type = op_type;
  v.dtype = dtype;
  v.trainable = trainable != 0;
  v.count = 1;
  for (int i = 0; i < rank; ++i) {
    if (shape[i] < 0) return fail(j, TFG_EINVAL, "negative dimension");
    v.shape.push_back(shape[i]);
    v.count *= shape[i];
  }
  // replica_device_setter: round-robin over ps tasks, one tick per PS op, in creation order
  if (j->n_ps > 0 && is_ps_op(v.op_type)) {
    v.ps_task = j->rr_next;
    j->rr_next = (j->rr_next + 1) % j->n_ps;
  } else {
    v.ps_task = -1;
  }
  j->vars.push_back(v);
  if (var_out) *var_out = (int)j->vars.size() - 1;
  if (ps_task_out) *ps_task_out = v.ps_task;
  return TFG_OK;
}

int tfg_num_vars(const tfg_job* j) { return j ? (int)j->vars.size() : -1; }

int tfg_var_info(const tfg_job* j, int var, char* name_out, size_t name_cap, int64_t* shape_out, int* rank_out,
                 int* trainable_out) {
  if (!j || var < 0 || var >= (int)j->vars.size()) return fail(nullptr, TFG_EINVAL, "bad variable index");
  const VarInfo& v = j->vars[var];
  if (name_out && name_cap) {
    const size_t n = std::min(name_cap - 1, v.name.size());
    memcpy(name_out, v.name.data(), n);
    name_out[n] = 0;
  }
  if (rank_out) *rank_out = (int)v.shape.size();
  if (shape_out)
    for (size_t i = 0; i < v.shape.size() && i < 8; ++i) shape_out[i] = v.shape[i];
  if (trainable_out) *trainable_out = (v.trainable && v.dtype == TFG_DT_F32) ? 1 : 0;
  return TFG_OK;
}

int tfg_vars_finalize(tfg_job* j, int layout) {
  if (!j) return fail(nullptr, TFG_EINVAL, "job is NULL");
  if (j->finalized) return fail(j, TFG_ESTATE, "variables are already finalized");
  if (layout != TFG_LAYOUT_LOGICAL && layout != TFG_LAYOUT_BALANCED)
    return fail(j, TFG_EINVAL, "unknown layout");
  int rc = ensure_bound(j);
  if (rc) return rc;
  int64_t off = 0;
  for (auto& v : j->vars) {
    if (v.dtype == TFG_DT_F32 && v.trainable) {
      v.offset = off;
      v.padded = round_up(std::max<int64_t>(v.count, 1), kPadFloats);
      off += v.padded;
    }
  }
  j->flat_count = off;
  j->layout = layout;
  size_t bytes = sizeof(float) * (size_t)std::max<int64_t>(off, kPadFloats);
  TFG_CUDA(j, cudaMalloc(&j->params, bytes));
  TFG_CUDA(j, cudaMemsetAsync(j->params, 0, bytes, j->stream));
  TFG_CUDA(j, cudaMalloc(&j->accum, bytes));
  TFG_CUDA(j, cudaMemsetAsync(j->accum, 0, bytes, j->stream));
  for (size_t s = 0; s < j->local_replicas.size(); ++s) {
    float* g = nullptr;
    TFG_CUDA(j, cudaMalloc(&g, bytes));
    TFG_CUDA(j, cudaMemsetAsync(g, 0, bytes, j->stream));
    j->grads.push_back(g);
    j->peer_grads[j->local_replicas[s]] = g;
  }
  j->peer_params[j->gpu_rank] = j->params;
  // operand-precision shadow of the parameters, kept current by the apply kernel: bf16 copy (bf16
  // math, ResNet-50) or the fp32 low-order plane of the 3xTF32 split
  if (j->model && mnist_math(j) == TFG_MATH_TF32X3) j->lp_mode = 2;
  else if ((j->model && mnist_math(j) == TFG_MATH_BF16) || j->resnet) j->lp_mode = 1;
  if (j->lp_mode) {
    const size_t lp_bytes = j->lp_mode == 2 ? bytes : bytes / 2;
    TFG_CUDA(j, cudaMalloc(&j->params_lp, lp_bytes));
    TFG_CUDA(j, cudaMemsetAsync(j->params_lp, 0, lp_bytes, j->stream));
    j->peer_params_lp[j->gpu_rank] = j->params_lp;
  }
  build_segments(j);
  if (!j->segs.empty()) {
    TFG_CUDA(j, cudaMalloc(&j->d_segs, sizeof(Segment) * j->segs.size()));
    TFG_CUDA(j, cudaMemcpyAsync(j->d_segs, j->segs.data(), sizeof(Segment) * j->segs.size(),
                                cudaMemcpyHostToDevice, j->stream));
  }
  TFG_CUDA(j, cudaStreamSynchronize(j->stream));
  j->finalized = true;
  return TFG_OK;
}

int tfg_placement(const tfg_job* j, int var, int* ps_task, int* gpu, int64_t* offset, int64_t* count) {
  if (!j || var < 0 || var >= (int)j->vars.size()) return fail(nullptr, TFG_EINVAL, "bad variable index");
  const VarInfo& v = j->vars[var];
  if (ps_task) *ps_task = v.ps_task;
  if (gpu) *gpu = v.ps_task < 0 ? 0 : v.ps_task % j->n_gpus;
  if (offset) *offset = v.offset;
  if (count) *count = v.count;
  return TFG_OK;
}

int64_t tfg_flat_count(const tfg_job* j) { return j ? j->flat_count : -1; }

static int var_copy(tfg_job* j, float* base, int var, void* host_out, const void* host_in, size_t bytes) {
  if (!j) return fail(nullptr, TFG_EINVAL, "job is NULL");
  if (!j->finalized) return fail(j, TFG_ESTATE, "variables are not finalized");
  if (var < 0 || var >= (int)j->vars.size()) return fail(j, TFG_EINVAL, "bad variable index");
  const VarInfo& v = j->vars[var];
  if (v.offset < 0) return fail(j, TFG_EINVAL, "variable " + v.name + " is not in the flat fp32 space");
  if (bytes != sizeof(float) * (size_t)v.count)
    return fail(j, TFG_EINVAL, "byte count does not match variable " + v.name);
  if (!base) return fail(j, TFG_EINVAL, "buffer does not exist");
  TFG_CUDA(j, cudaSetDevice(j->device));
  if (host_in)
    TFG_CUDA(j, cudaMemcpyAsync(base + v.offset, host_in, bytes, cudaMemcpyHostToDevice, j->stream));
  else
    TFG_CUDA(j, cudaMemcpyAsync(host_out, base + v.offset, bytes, cudaMemcpyDeviceToHost, j->stream));
  TFG_CUDA(j, cudaStreamSynchronize(j->stream));
  return TFG_OK;
}

int tfg_var_set(tfg_job* j, int var, const void* host, size_t bytes) {
  int rc = var_copy(j, j ? j->params : nullptr, var, nullptr, host, bytes);
  if (rc == TFG_OK && j->params_lp) {   // keep the operand-precision shadow current
    const VarInfo& v = j->vars[var];
    const size_t es = j->lp_mode == 2 ? 4 : 2;
    if (launch_cast_lp(j->params + v.offset, static_cast<char*>(j->params_lp) + es * v.offset, v.count, j->lp_mode, j->stream))
      return fail(j, TFG_ECUDA, "parameter shadow cast failed");
    TFG_CUDA(j, cudaStreamSynchronize(j->stream));
  }
  return rc;
}
int tfg_var_get(tfg_job* j, int var, void* host, size_t bytes) {
  return var_copy(j, j ? j->params : nullptr, var, host, nullptr, bytes);
}
int tfg_accum_get(tfg_job* j, int var, void* host, size_t bytes) {
  return var_copy(j, j ? j->accum : nullptr, var, host, nullptr, bytes);
}
int tfg_grad_set(tfg_job* j, int slot, int var, const void* host, size_t bytes) {
  if (!j || slot < 0 || slot >= (int)j->grads.size()) return fail(j, TFG_EINVAL, "bad replica slot");
  return var_copy(j, j->grads[slot], var, nullptr, host, bytes);
}
int tfg_grad_get(tfg_job* j, int slot, int var, void* host, size_t bytes) {
  if (!j || slot < 0 || slot >= (int)j->grads.size()) return fail(j, TFG_EINVAL, "bad replica slot");
  return var_copy(j, j->grads[slot], var, host, nullptr, bytes);
}

int tfg_job_buffer(tfg_job* j, int which, int slot, void** dev_out, size_t* bytes_out) {
  if (!j || !dev_out) return fail(j, TFG_EINVAL, "NULL argument");
  size_t flat = sizeof(float) * (size_t)j->flat_count;
  void* p = nullptr;
  size_t b = 0;
  switch (which) {
    case TFG_BUF_PARAMS: p = j->params; b = flat; break;
    case TFG_BUF_ACCUM: p = j->accum; b = flat; break;
    case TFG_BUF_GRADS:
      if (slot < 0 || slot >= (int)j->grads.size()) return fail(j, TFG_EINVAL, "bad replica slot");
      p = j->grads[slot]; b = flat; break;
    case TFG_BUF_LOSS: p = j->loss; b = sizeof(float) * (TFG_MAX_REPLICAS + 1); break;
    case TFG_BUF_FLAGS: p = j->flags; b = sizeof(uint32_t) * TFG_MAX_GPUS * kFlagRows; break;
    default: return fail(j, TFG_EINVAL, "unknown buffer");
  }
  if (!p) return fail(j, TFG_ESTATE, "buffer is not allocated yet");
  *dev_out = p;
  if (bytes_out) *bytes_out = b;
  return TFG_OK;
}

int tfg_opt_config(tfg_job* j, int kind, float lr, float momentum) {
  if (!j) return fail(nullptr, TFG_EINVAL, "job is NULL");
  if (kind != TFG_OPT_SGD && kind != TFG_OPT_MOMENTUM && kind != TFG_OPT_NESTEROV)
    return fail(j, TFG_EINVAL, "unknown optimizer kind");
  if (kind != j->opt_kind || lr != j->lr || momentum != j->mu) {
    // the captured step graphs carry the optimizer's constants as kernel arguments: drop them, the
    // next step with each input set runs eagerly once and is captured again
    if (j->bound && !j->step_graphs.empty()) {
      TFG_CUDA(j, cudaSetDevice(j->device));
      TFG_CUDA(j, cudaStreamSynchronize(j->stream));
      for (auto& g : j->step_graphs)
        if (g.exec) cudaGraphExecDestroy(g.exec);
      j->step_graphs.clear();
    }
  }
  j->opt_kind = kind;
  j->lr = lr;
  j->mu = momentum;
  return TFG_OK;
}

int64_t tfg_global_step(const tfg_job* j) { return j ? j->global_step : -1; }

// ---- raw access to the flat space (checkpoint of the sharded variables) ---------------------
int tfg_global_step_set(tfg_job* j, int64_t step) {
  if (!j || step < 0) return fail(j, TFG_EINVAL, "bad argument");
  j->global_step = step;
  return TFG_OK;
}

int tfg_owned_ranges(const tfg_job* j, int64_t* offsets_out, int64_t* counts_out, int cap, int* n_out) {
  if (!j || !n_out) return fail(nullptr, TFG_EINVAL, "NULL argument");
  if (!j->finalized) return fail(nullptr, TFG_ESTATE, "variables are not finalized");
  *n_out = (int)j->segs.size();
  if (cap < (int)j->segs.size()) return (offsets_out || counts_out) ? fail(nullptr, TFG_ELIMIT, "range arrays too small") : TFG_OK;
  for (size_t i = 0; i < j->segs.size(); ++i) {
    if (offsets_out) offsets_out[i] = j->segs[i].off;
    if (counts_out) counts_out[i] = j->segs[i].cnt;
  }
  return TFG_OK;
}

static int flat_io(tfg_job* j, int which, int64_t offset, int64_t count, void* host_out, const void* host_in) {
  if (!j) return fail(nullptr, TFG_EINVAL, "job is NULL");
  if (!j->finalized) return fail(j, TFG_ESTATE, "variables are not finalized");
  if (which != TFG_BUF_PARAMS && which != TFG_BUF_ACCUM) return fail(j, TFG_EINVAL, "flat access takes TFG_BUF_PARAMS or TFG_BUF_ACCUM");
  if (offset < 0 || count < 0 || offset + count > j->flat_count) return fail(j, TFG_EINVAL, "range outside the flat space");
  if (count == 0) return TFG_OK;
  if (!host_out && !host_in) return fail(j, TFG_EINVAL, "NULL host buffer");
  float* base = which == TFG_BUF_PARAMS ? j->params : j->accum;
  TFG_CUDA(j, cudaSetDevice(j->device));
  const size_t bytes = sizeof(float) * (size_t)count;
  if (host_in) {
    TFG_CUDA(j, cudaMemcpyAsync(base + offset, host_in, bytes, cudaMemcpyHostToDevice, j->stream));
    if (which == TFG_BUF_PARAMS && j->params_lp) {   // keep the operand-precision shadow current
      const size_t es = j->lp_mode == 2 ? 4 : 2;
      if (launch_cast_lp(j->params + offset, static_cast<char*>(j->params_lp) + es * offset, count, j->lp_mode, j->stream))
        return fail(j, TFG_ECUDA, "parameter shadow cast failed");
    }
  } else {
    TFG_CUDA(j, cudaMemcpyAsync(host_out, base + offset, bytes, cudaMemcpyDeviceToHost, j->stream));
  }
  TFG_CUDA(j, cudaStreamSynchronize(j->stream));
  return TFG_OK;
}

int tfg_flat_read(tfg_job* j, int which, int64_t offset, int64_t count, void* host) {
  return flat_io(j, which, offset, count, host, nullptr);
}
int tfg_flat_write(tfg_job* j, int which, int64_t offset, int64_t count, const void* host) {
  return flat_io(j, which, offset, count, nullptr, host);
}

// ---- peer mapping --------------------------------------------------------------------
// blob = [n_local_replicas:int32][global replica ids...] then handles in the order
// params, loss, flags, grads[slot]...
int tfg_job_ipc_export(tfg_job* j, void* blob_out, size_t cap, size_t* bytes_out) {
  if (!j || !blob_out) return fail(j, TFG_EINVAL, "NULL argument");
  if (!j->finalized) return fail(j, TFG_ESTATE, "variables are not finalized");
  size_t nl = j->local_replicas.size();
  const size_t has_lp = j->params_lp ? 1 : 0;
  size_t need = 4 + 4 * nl + TFG_IPC_HANDLE_BYTES * (3 + nl + has_lp);
  if (bytes_out) *bytes_out = need;
  if (cap < need) return fail(j, TFG_ELIMIT, "blob_out too small");
  static_assert(sizeof(cudaIpcMemHandle_t) == TFG_IPC_HANDLE_BYTES, "IPC handle size");
  TFG_CUDA(j, cudaSetDevice(j->device));
  char* p = (char*)blob_out;
  int32_t n32 = (int32_t)nl | (has_lp ? 0x10000 : 0);   // bit 16: a bf16 shadow handle follows the gradients
  memcpy(p, &n32, 4); p += 4;
  for (size_t s = 0; s < nl; ++s) { int32_t r = j->local_replicas[s]; memcpy(p, &r, 4); p += 4; }
  void* bufs[3] = {j->params, j->loss, j->flags};
  for (void* b : bufs) {
    cudaIpcMemHandle_t h;
    TFG_CUDA(j, cudaIpcGetMemHandle(&h, b));
    memcpy(p, &h, sizeof h); p += sizeof h;
  }
  for (size_t s = 0; s < nl; ++s) {
    cudaIpcMemHandle_t h;
    TFG_CUDA(j, cudaIpcGetMemHandle(&h, j->grads[s]));
    memcpy(p, &h, sizeof h); p += sizeof h;
  }
  if (has_lp) {
    cudaIpcMemHandle_t h;
    TFG_CUDA(j, cudaIpcGetMemHandle(&h, j->params_lp));
    memcpy(p, &h, sizeof h); p += sizeof h;
  }
  return TFG_OK;
}

int tfg_job_ipc_open(tfg_job* j, int peer, const void* blob, size_t bytes) {
  if (!j || !blob) return fail(j, TFG_EINVAL, "NULL argument");
  if (peer < 0 || peer >= j->n_gpus || peer == j->gpu_rank) return fail(j, TFG_EINVAL, "bad peer rank");
  if (!j->finalized) return fail(j, TFG_ESTATE, "variables are not finalized");
  const char* p = (const char*)blob;
  if (bytes < 4) return fail(j, TFG_EINVAL, "short blob");
  int32_t nl;
  memcpy(&nl, p, 4); p += 4;
  const bool has_lp = (nl & 0x10000) != 0;
  nl &= 0xffff;
  if (nl < 0 || nl > TFG_MAX_REPLICAS ||
      bytes < 4 + 4 * (size_t)nl + TFG_IPC_HANDLE_BYTES * (3 + (size_t)nl + (has_lp ? 1 : 0)))
    return fail(j, TFG_EINVAL, "malformed blob");
  if (has_lp != (j->params_lp != nullptr)) return fail(j, TFG_EINVAL, "peer runs a different math mode");
  std::vector<int32_t> reps(nl);
  for (int s = 0; s < nl; ++s) { memcpy(&reps[s], p, 4); p += 4; }
  TFG_CUDA(j, cudaSetDevice(j->device));
  auto open = [&](void** out) -> cudaError_t {
    cudaIpcMemHandle_t h;
    memcpy(&h, p, sizeof h); p += sizeof h;
    cudaError_t e = cudaIpcOpenMemHandle(out, h, cudaIpcMemLazyEnablePeerAccess);
    if (e == cudaSuccess) j->ipc_opened.push_back(*out);
    return e;
  };
  void* q = nullptr;
  cudaError_t e;
  if ((e = open(&q)) != cudaSuccess) { fail_cuda(j, e, "cudaIpcOpenMemHandle(params)"); return TFG_EPEER; }
  j->peer_params[peer] = (float*)q;
  if ((e = open(&q)) != cudaSuccess) { fail_cuda(j, e, "cudaIpcOpenMemHandle(loss)"); return TFG_EPEER; }
  j->peer_loss[peer] = (float*)q;
  if ((e = open(&q)) != cudaSuccess) { fail_cuda(j, e, "cudaIpcOpenMemHandle(flags)"); return TFG_EPEER; }
  j->peer_flags[peer] = (uint32_t*)q;
  for (int s = 0; s < nl; ++s) {
    if (reps[s] < 0 || reps[s] >= j->n_replicas) return fail(j, TFG_EINVAL, "peer replica id out of range");
    if ((e = open(&q)) != cudaSuccess) { fail_cuda(j, e, "cudaIpcOpenMemHandle(grads)"); return TFG_EPEER; }
    j->peer_grads[reps[s]] = (float*)q;
  }
  if (has_lp) {
    if ((e = open(&q)) != cudaSuccess) { fail_cuda(j, e, "cudaIpcOpenMemHandle(params_lp)"); return TFG_EPEER; }
    j->peer_params_lp[peer] = q;
  }
  return TFG_OK;
}

int tfg_job_attach_local_peer(tfg_job* j, int peer_rank, tfg_job* peer) {
  if (!j || !peer) return fail(j, TFG_EINVAL, "NULL argument");
  if (peer_rank < 0 || peer_rank >= j->n_gpus || peer_rank == j->gpu_rank)
    return fail(j, TFG_EINVAL, "bad peer rank");
  if (!j->finalized || !peer->finalized) return fail(j, TFG_ESTATE, "variables are not finalized");
  if (peer->flat_count != j->flat_count) return fail(j, TFG_EINVAL, "peer has a different flat space");
  j->peer_params[peer_rank] = peer->params;
  j->peer_params_lp[peer_rank] = peer->params_lp;
  if ((peer->params_lp != nullptr) != (j->params_lp != nullptr)) return fail(j, TFG_EINVAL, "peer runs a different math mode");
  j->peer_loss[peer_rank] = peer->loss;
  j->peer_flags[peer_rank] = peer->flags;
  for (size_t s = 0; s < peer->local_replicas.size(); ++s)
    j->peer_grads[peer->local_replicas[s]] = peer->grads[s];
  return TFG_OK;
}

static int check_peers(tfg_job* j) {
  for (int g = 0; g < j->n_gpus; ++g)
    if (!j->peer_params[g] || !j->peer_flags[g] || !j->peer_loss[g] || (j->params_lp && !j->peer_params_lp[g]))
      return fail(j, TFG_ESTATE, "GPU rank " + std::to_string(g) + " is not mapped (tfg_job_ipc_open)");
  for (int r = 0; r < j->n_replicas; ++r)
    if (!j->peer_grads[r])
      return fail(j, TFG_ESTATE, "replica " + std::to_string(r) + " gradients are not mapped");
  return TFG_OK;
}

int tfg_job_barrier(tfg_job* j) {
  if (!j || !j->finalized) return fail(j, TFG_ESTATE, "job is not finalized");
  if (j->n_gpus == 1) return TFG_OK;
  int rc = check_peers(j);
  if (rc) return rc;
  TFG_CUDA(j, cudaSetDevice(j->device));
  return launch_barrier(j);
}

int tfg_push_apply_pull(tfg_job* j, int with_barriers) {
  if (!j || !j->finalized) return fail(j, TFG_ESTATE, "job is not finalized");
  int rc = check_peers(j);
  if (rc) return rc;
  TFG_CUDA(j, cudaSetDevice(j->device));
  // with_barriers: the two cross-rank synchronisation points are part of the kernel itself
  if ((rc = launch_job_apply(j, false, j->n_replicas, false, with_barriers != 0))) return rc;
  j->global_step += 1;
  return TFG_OK;
}

int tfg_apply_owned(tfg_job* j, int n_summed) {
  if (!j || !j->finalized) return fail(j, TFG_ESTATE, "job is not finalized");
  if (n_summed < 1) return fail(j, TFG_EINVAL, "n_summed must be >= 1");
  if (j->grads.empty()) return fail(j, TFG_ESTATE, "this rank hosts no replica");
  TFG_CUDA(j, cudaSetDevice(j->device));
  int rc = launch_job_apply(j, true, n_summed);
  if (rc) return rc;
  j->global_step += 1;
  return TFG_OK;
}

int tfg_apply_all(tfg_job* j, int n_summed) {
  if (!j || !j->finalized) return fail(j, TFG_ESTATE, "job is not finalized");
  if (n_summed < 1) return fail(j, TFG_EINVAL, "n_summed must be >= 1");
  if (j->grads.empty()) return fail(j, TFG_ESTATE, "this rank hosts no replica");
  TFG_CUDA(j, cudaSetDevice(j->device));
  int rc = launch_job_apply(j, true, n_summed, true);
  if (rc) return rc;
  j->global_step += 1;
  return TFG_OK;
}

int tfg_dense_apply(int kind, float* var, float* accum, const float* const* grads, int n_grads,
                    float divisor, int64_t count, float lr, float mu, float* const* pull, int n_pull,
                    void* stream) {
  if (!var || !grads || count < 0) return fail(nullptr, TFG_EINVAL, "NULL argument");
  if (kind != TFG_OPT_SGD && kind != TFG_OPT_MOMENTUM && kind != TFG_OPT_NESTEROV)
    return fail(nullptr, TFG_EINVAL, "unknown optimizer kind");
  if (kind != TFG_OPT_SGD && !accum) return fail(nullptr, TFG_EINVAL, "momentum needs an accumulator");
  if (n_grads < 1 || n_grads > TFG_MAX_REPLICAS) return fail(nullptr, TFG_ELIMIT, "n_grads out of range");
  if (n_pull < 0 || n_pull > TFG_MAX_GPUS) return fail(nullptr, TFG_ELIMIT, "n_pull out of range");
  if (n_pull > 0 && !pull) return fail(nullptr, TFG_EINVAL, "pull is NULL");
  if (count == 0) return TFG_OK;
  return launch_dense_apply(kind, var, accum, grads, n_grads, divisor, count, lr, mu, pull, n_pull,
                            (cudaStream_t)stream);
}

// ---- model + step --------------------------------------------------------------------
static bool has_model(const tfg_job* j) { return j->model || j->resnet; }
static int model_fb(tfg_job* j, int slot, const float* x, const int32_t* y) {
  return j->resnet ? resnet_forward_backward(j, slot, x, y) : mnist_forward_backward(j, slot, x, y);
}
static int model_batch(const tfg_job* j) { return j->resnet ? resnet_batch(j) : mnist_batch(j); }
static size_t model_image_floats(const tfg_job* j) { return j->resnet ? resnet_image_floats(j) : 784; }
static float* model_stage_images(tfg_job* j, int s) { return j->resnet ? resnet_stage_images(j, s) : mnist_stage_images(j, s); }
static int32_t* model_stage_labels(tfg_job* j, int s) { return j->resnet ? resnet_stage_labels(j, s) : mnist_stage_labels(j, s); }

int tfg_model_resnet50(tfg_job* j, int batch, int image_hw, int classes) {
  if (!j) return fail(nullptr, TFG_EINVAL, "job is NULL");
  if (has_model(j)) return fail(j, TFG_ESTATE, "a model is already built");
  if (j->finalized) return fail(j, TFG_ESTATE, "variables are already finalized");
  if (batch < 1 || batch > 4096) return fail(j, TFG_EINVAL, "batch_per_replica out of range");
  if (image_hw < 32 || image_hw % 32 != 0 || image_hw > 512)
    return fail(j, TFG_EINVAL, "image size must be a multiple of 32 in [32, 512]");
  if (classes < 2 || classes % 8 != 0) return fail(j, TFG_EINVAL, "classes must be a multiple of 8");
  int rc = ensure_bound(j);
  if (rc) return rc;
  return resnet_build(j, batch, image_hw, classes);
}

int tfg_tensor_get(tfg_job* j, int slot, const char* name, float* host, size_t cap_floats, size_t* n_out) {
  if (!j || !j->resnet) return fail(j, TFG_ESTATE, "no ResNet model is built");
  if (slot < 0 || slot >= (int)j->grads.size() || !name || !host) return fail(j, TFG_EINVAL, "bad argument");
  TFG_CUDA(j, cudaSetDevice(j->device));
  return resnet_tensor_get(j, slot, name, host, cap_floats, n_out);
}

int tfg_model_mnist_cnn(tfg_job* j, int batch, int math) {
  if (!j) return fail(nullptr, TFG_EINVAL, "job is NULL");
  if (has_model(j)) return fail(j, TFG_ESTATE, "a model is already built");
  if (j->finalized) return fail(j, TFG_ESTATE, "variables are already finalized");
  if (batch < 1 || batch > 65536) return fail(j, TFG_EINVAL, "batch_per_replica out of range");
  if (math != TFG_MATH_FP32 && math != TFG_MATH_BF16 && math != TFG_MATH_TF32 && math != TFG_MATH_TF32X3)
    return fail(j, TFG_EINVAL, "unknown math mode");
  int rc = ensure_bound(j);
  if (rc) return rc;
  return mnist_build(j, batch, math);
}

int tfg_forward_backward(tfg_job* j, int slot, const float* images, const int32_t* labels) {
  if (!j || !has_model(j)) return fail(j, TFG_ESTATE, "no model is built");
  if (!j->finalized) return fail(j, TFG_ESTATE, "variables are not finalized");
  if (slot < 0 || slot >= (int)j->grads.size()) return fail(j, TFG_EINVAL, "bad replica slot");
  if (!images || !labels) return fail(j, TFG_EINVAL, "NULL input");
  TFG_CUDA(j, cudaSetDevice(j->device));
  return model_fb(j, slot, images, labels);
}

int tfg_logits_get(tfg_job* j, int slot, float* host, size_t bytes) {
  if (!j || !has_model(j)) return fail(j, TFG_ESTATE, "no model is built");
  if (slot < 0 || slot >= (int)j->grads.size()) return fail(j, TFG_EINVAL, "bad replica slot");
  TFG_CUDA(j, cudaSetDevice(j->device));
  return j->resnet ? resnet_logits(j, slot, host, bytes) : mnist_logits(j, slot, host, bytes);
}

// Device work of one step: forward/backward of every hosted replica, then the fused push + apply +
// pull kernel, which also carries both cross-rank synchronisation points.  Pure stream work with
// fixed pointers, hence capturable.
static int step_device_work(tfg_job* j, const float* const* images, const int32_t* const* labels) {
  int rc;
  for (size_t s = 0; s < j->grads.size(); ++s)
    if ((rc = model_fb(j, (int)s, images[s], labels[s]))) return rc;
  NvtxRange nvtx("push + apply + pull");
  return launch_job_apply(j, false, j->n_replicas, false, true);   // the cross-rank synchronisation is inside the kernel
}

// Runs the step eagerly the first time a set of input pointers is seen (this also performs
// every one-time set-up), captures it into a CUDA graph the second time, replays afterwards:
// the ~30 launches of a step cost one graph launch.  TFGRAFT_GRAPH=0 or profiling = eager.
static int step_run(tfg_job* j, const float* const* images, const int32_t* const* labels) {
  int rc = check_peers(j);
  if (rc) return rc;
  const size_t ns = j->grads.size();
  tfg_job::StepGraph* g = nullptr;
  if (j->use_graphs && !j->prof_on) {
    std::vector<const void*> key;
    for (size_t s = 0; s < ns; ++s) { key.push_back(images[s]); key.push_back(labels[s]); }
    for (auto& e : j->step_graphs)
      if (e.key == key) g = &e;
    if (!g) {
      if (j->step_graphs.size() >= 32) {   // bounded cache: the oldest input set makes room
        TFG_CUDA(j, cudaStreamSynchronize(j->stream));
        if (j->step_graphs.front().exec) cudaGraphExecDestroy(j->step_graphs.front().exec);
        j->step_graphs.erase(j->step_graphs.begin());
      }
      j->step_graphs.emplace_back();
      g = &j->step_graphs.back();
      g->key = key;
    }
  }
  if (!g || ++g->calls == 1) {
    rc = step_device_work(j, images, labels);
  } else if (!g->exec) {
    const int64_t l0 = j->launches;
    cudaGraph_t graph = nullptr;
    cudaError_t e = cudaStreamBeginCapture(j->stream, cudaStreamCaptureModeRelaxed);
    if (e != cudaSuccess) return fail_cuda(j, e, "cudaStreamBeginCapture");
    rc = step_device_work(j, images, labels);
    e = cudaStreamEndCapture(j->stream, &graph);
    if (rc == TFG_OK && e == cudaSuccess) e = cudaGraphInstantiate(&g->exec, graph, 0);
    if (graph) cudaGraphDestroy(graph);
    g->launches = j->launches - l0;
    j->launches = l0;
    if (rc) return rc;
    if (e != cudaSuccess) { g->exec = nullptr; j->use_graphs = false; return fail_cuda(j, e, "step graph capture"); }
    TFG_CUDA(j, cudaGraphLaunch(g->exec, j->stream));
    j->launches += g->launches;
  } else {
    TFG_CUDA(j, cudaGraphLaunch(g->exec, j->stream));
    j->launches += g->launches;
  }
  if (rc) return rc;
  j->global_step += 1;
  return TFG_OK;
}

static int step_finish(tfg_job* j, float* loss_out) {
  if (loss_out) {
    TFG_CUDA(j, cudaMemcpyAsync(j->h_loss + TFG_MAX_REPLICAS, j->loss + TFG_MAX_REPLICAS, 2 * sizeof(float),
                                cudaMemcpyDeviceToHost, j->stream));   // mean loss + error word
    TFG_CUDA(j, cudaStreamSynchronize(j->stream));
    if (*j->h_err) return fail(j, TFG_EPEER, "cross-rank barrier timed out (a peer rank is gone?)");
    *loss_out = j->h_loss[TFG_MAX_REPLICAS];
  }
  return TFG_OK;
}

int tfg_step_dense_dev(tfg_job* j, const float* const* images, const int32_t* const* labels,
                       float* loss_out) {
  if (!j || !has_model(j)) return fail(j, TFG_ESTATE, "no model is built");
  if (!j->finalized) return fail(j, TFG_ESTATE, "variables are not finalized");
  if (!images || !labels) return fail(j, TFG_EINVAL, "NULL input");
  TFG_CUDA(j, cudaSetDevice(j->device));
  for (size_t s = 0; s < j->grads.size(); ++s)
    if (!images[s] || !labels[s]) return fail(j, TFG_EINVAL, "NULL input for a replica slot");
  int rc = step_run(j, images, labels);
  if (rc) return rc;
  return step_finish(j, loss_out);
}

int tfg_step_dense(tfg_job* j, const float* const* images_host, const int32_t* const* labels_host,
                   float* loss_out) {
  if (!j || !has_model(j)) return fail(j, TFG_ESTATE, "no model is built");
  if (!j->finalized) return fail(j, TFG_ESTATE, "variables are not finalized");
  if (!images_host || !labels_host || !loss_out) return fail(j, TFG_EINVAL, "NULL argument");
  TFG_CUDA(j, cudaSetDevice(j->device));
  const int B = model_batch(j);
  const size_t img = model_image_floats(j);
  std::vector<const float*> di(j->grads.size());
  std::vector<const int32_t*> dl(j->grads.size());
  for (size_t s = 0; s < j->grads.size(); ++s) {
    if (!images_host[s] || !labels_host[s]) return fail(j, TFG_EINVAL, "NULL input for a replica slot");
    di[s] = model_stage_images(j, (int)s);
    dl[s] = model_stage_labels(j, (int)s);
    TFG_CUDA(j, cudaMemcpyAsync(const_cast<float*>(di[s]), images_host[s], sizeof(float) * img * (size_t)B,
                                cudaMemcpyHostToDevice, j->stream));
    TFG_CUDA(j, cudaMemcpyAsync(const_cast<int32_t*>(dl[s]), labels_host[s], sizeof(int32_t) * (size_t)B,
                                cudaMemcpyHostToDevice, j->stream));
  }
  int rc = step_run(j, di.data(), dl.data());
  if (rc) return rc;
  return step_finish(j, loss_out);
}

int tfg_loss_get(tfg_job* j, int replica, float* loss_out) {
  if (!j || !j->bound || !loss_out) return fail(j, TFG_EINVAL, "NULL argument or unbound job");
  if (replica < -1 || replica >= j->n_replicas) return fail(j, TFG_EINVAL, "replica out of range");
  TFG_CUDA(j, cudaSetDevice(j->device));
  const float* src = j->loss + (replica < 0 ? TFG_MAX_REPLICAS : replica);
  TFG_CUDA(j, cudaMemcpyAsync(j->h_loss, src, sizeof(float), cudaMemcpyDeviceToHost, j->stream));
  TFG_CUDA(j, cudaStreamSynchronize(j->stream));
  *loss_out = j->h_loss[0];
  return TFG_OK;
}

int tfg_job_sync(tfg_job* j) {
  if (!j || !j->bound) return fail(j, TFG_ESTATE, "job is not bound to a device");
  TFG_CUDA(j, cudaSetDevice(j->device));
  TFG_CUDA(j, cudaStreamSynchronize(j->stream));
  TFG_CUDA(j, cudaMemcpy(j->h_err, j->dev_err, sizeof(int), cudaMemcpyDeviceToHost));
  if (*j->h_err) return fail(j, TFG_EPEER, "cross-rank barrier timed out (a peer rank is gone?)");
  return TFG_OK;
}

int64_t tfg_launch_count(const tfg_job* j) { return j ? j->launches : -1; }

int tfg_profile_enable(tfg_job* j, int on) {
  if (!j || !j->bound) return fail(j, TFG_ESTATE, "job is not bound to a device");
  TFG_CUDA(j, cudaSetDevice(j->device));
  TFG_CUDA(j, cudaStreamSynchronize(j->stream));
  j->prof_used = 0;
  j->prof_on = on != 0;
  if (j->prof_on) prof_mark(j, "(start)");
  return TFG_OK;
}

int tfg_profile_report(tfg_job* j, char* json_out, size_t cap) {
  if (!j || !j->bound || !json_out || cap < 3) return fail(j, TFG_EINVAL, "bad argument");
  TFG_CUDA(j, cudaSetDevice(j->device));
  TFG_CUDA(j, cudaStreamSynchronize(j->stream));
  std::vector<std::pair<std::string, std::pair<long long, double>>> agg;
  for (size_t i = 1; i < j->prof_used; ++i) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, j->prof_ev[i - 1].second, j->prof_ev[i].second) != cudaSuccess) continue;
    std::string name = j->prof_ev[i].first;
    if (name == "(start)") continue;   // a marker, not a kernel
    bool found = false;
    for (auto& a : agg)
      if (a.first == name) { a.second.first++; a.second.second += ms; found = true; break; }
    if (!found) agg.push_back({name, {1, (double)ms}});
  }
  std::string out = "{";
  for (size_t i = 0; i < agg.size(); ++i) {
    char buf[96];
    snprintf(buf, sizeof buf, "{\"count\":%lld,\"ms\":%.6f}", agg[i].second.first, agg[i].second.second);
    if (i) out += ",";
    json::escape(agg[i].first, &out);
    out += ":";
    out += buf;
  }
  out += "}";
  j->prof_used = 0;
  if (j->prof_on) prof_mark(j, "(start)");
  if (out.size() + 1 > cap) return fail(j, TFG_ELIMIT, "json_out too small");
  memcpy(json_out, out.c_str(), out.size() + 1);
  return TFG_OK;
}

// ---- smoke (A9) ----------------------------------------------------------------------
int tfg_smoke(tfg_job* j, int all_tasks, int32_t* out_host, size_t cap, int* n_tasks_out) {
  if (!j || !out_host) return fail(j, TFG_EINVAL, "NULL argument");
  int n_tasks = 0;
  for (auto& kv : j->cluster) n_tasks += (int)kv.second.size();
  if (n_tasks_out) *n_tasks_out = n_tasks;
  if (cap < (size_t)n_tasks * 100) return fail(j, TFG_ELIMIT, "out_host too small");
  if (n_tasks == 0) return TFG_OK;  // TF_CONFIG unset: the loop body never runs (tf_smoke.py:53-54)
  int rc = ensure_bound(j);
  if (rc) return rc;
  int32_t* d = nullptr;
  TFG_CUDA(j, cudaMalloc(&d, sizeof(int32_t) * 100 * (size_t)n_tasks));
  TFG_CUDA(j, cudaMemsetAsync(d, 0, sizeof(int32_t) * 100 * (size_t)n_tasks, j->stream));
  // task t of a job is hosted by GPU t % n_gpus (master/worker: replica order; ps: owner)
  int t = 0;
  for (auto& kv : j->cluster) {
    int replica_base = 0;
    if (kv.first == "worker") replica_base = std::max(0, tfg_cluster_num_tasks(j, "master"));
    for (size_t i = 0; i < kv.second.size(); ++i, ++t) {
      int gpu = (int)((replica_base + i) % j->n_gpus);
      if (all_tasks || gpu == j->gpu_rank) {
        rc = launch_smoke(d + 100 * (size_t)t, 1, j->stream);
        prof_mark(j, "smoke_mul_kernel");
        if (rc) { cudaFree(d); return fail(j, rc, "smoke kernel launch failed"); }
      }
    }
  }
  cudaError_t e = cudaMemcpyAsync(out_host, d, sizeof(int32_t) * 100 * (size_t)n_tasks,
                                  cudaMemcpyDeviceToHost, j->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(j->stream);
  cudaFree(d);
  if (e != cudaSuccess) return fail_cuda(j, e, "smoke copy-back");
  return TFG_OK;
}

}  // extern "C"
