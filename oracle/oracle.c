/* ORACLE -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Plain-C restatement, on the CPU, of the arithmetic of the synchronous parameter-server
 * step that mitake/k8s orchestrates.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load it; the product (k8s_b200/, libtfgraft)
 * never does.
 *
 * PARITY UNPINNED for everything except orc_smoke: the reference tree holds no training
 * arithmetic (SURVEY.md par. 0.1).  The algorithms live in TensorFlow 1.3.0, a dependency
 * pinned only as a Docker image tag (pkg/spec/tf_job.go:87,
 * examples/tf_sample/Dockerfile.template:1) whose source is not under /root/reference and
 * which cannot be installed here.  What follows restates TF 1.3's published kernels from
 * their documented semantics; no golden vector of the reference exists to check them
 * against.  The reference's own call sites that anchor each function are cited.
 *
 * orc_smoke IS pinned: it restates examples/tf_sample/tf_sample/tf_smoke.py:57-59, whose
 * result is fixed by construction (c[i][j] = (10 i + j)^2).
 *
 * Build: see oracle/Makefile (gcc -O2 -ffp-contract=off: every product is rounded before
 * it is added, as in TF 1.3's non-FMA Eigen build).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORC_SGD 0
#define ORC_MOMENTUM 1
#define ORC_NESTEROV 2
#define ORC_MOD 0
#define ORC_DIV 1

/* tf_smoke.py:57-59: a = b = tf.constant(range(100), shape=[10,10]); c = tf.multiply(a, b),
 * once per task of the cluster (tf_smoke.py:53-56). */
void orc_smoke(int32_t* out, int n_tasks) {
  for (int t = 0; t < n_tasks; ++t)
    for (int i = 0; i < 100; ++i) {
      int32_t a = i, b = i;
      out[t * 100 + i] = a * b;
    }
}

/* [TF1.3] tf.train.replica_device_setter, call site tf_smoke.py:116-118: round-robin over
 * ps tasks, one tick per PS op ("Variable", "VariableV2", "VarHandleOp") in creation order.
 * is_ps_op[i] != 0 marks PS ops; task_out[i] = -1 for ops that stay on the worker. */
void orc_round_robin(const int* is_ps_op, int n_ops, int n_ps, int* task_out) {
  int next = 0;
  for (int i = 0; i < n_ops; ++i) {
    if (n_ps > 0 && is_ps_op[i]) {
      task_out[i] = next;
      next = (next + 1) % n_ps;
    } else {
      task_out[i] = -1;
    }
  }
}

/* [TF1.3] sync-PS dense update of one shard (SURVEY.md par. 8a A12+A13):
 *   ConditionalAccumulator: g = (g_0 + g_1 + ... + g_{n-1}) / n   (take_grad = sum / count;
 *     arrival order is unspecified in TF; the build fixes replica order)
 *   ApplyGradientDescent:   var -= g * lr
 *   ApplyMomentum:          accum = accum * mu + g ; var -= accum * lr
 *     use_nesterov:         var -= g * lr + accum * mu * lr
 * Anchor in the reference: the PS processes that host these kernels are started by
 * grpc_tensorflow_server/grpc_tensorflow_server.py:112-115. */
void orc_dense_apply(int kind, float* var, float* accum, const float* const* grads, int n_grads,
                     float divisor, int64_t count, float lr, float mu) {
  for (int64_t i = 0; i < count; ++i) {
    float g = grads[0][i];
    for (int r = 1; r < n_grads; ++r) g = g + grads[r][i];
    g = g / divisor;
    if (kind == ORC_SGD) {
      float d = g * lr;
      var[i] = var[i] - d;
    } else {
      float a = accum[i] * mu;
      a = a + g;
      accum[i] = a;
      if (kind == ORC_MOMENTUM) {
        float d = a * lr;
        var[i] = var[i] - d;
      } else {
        float d1 = g * lr;
        float d2 = a * mu;
        d2 = d2 * lr;
        float d = d1 + d2;
        var[i] = var[i] - d;
      }
    }
  }
}

/* [TF1.3] embedding_lookup partition strategies (SURVEY.md par. 8a A15).
 * "mod": shard = id % S, row = id / S.
 * "div": ipp = V / S, extras = V % S; shard = max(id / (ipp+1), (id - extras) / ipp);
 *        row = shard < extras ? id % (ipp+1) : (id - extras) % ipp. */
void orc_locate(int64_t vocab, int n_shards, int strategy, int64_t id, int* shard, int64_t* row) {
  if (strategy == ORC_MOD) {
    *shard = (int)(id % n_shards);
    *row = id / n_shards;
    return;
  }
  int64_t ipp = vocab / n_shards, extras = vocab % n_shards;
  if (ipp == 0) { *shard = (int)id; *row = 0; return; }
  int64_t a = id / (ipp + 1), b = (id - extras) / ipp;
  int64_t p = a > b ? a : b;
  *shard = (int)p;
  *row = p < extras ? id % (ipp + 1) : (id - extras) % ipp;
}

int64_t orc_shard_rows(int64_t vocab, int n_shards, int shard, int strategy) {
  int64_t n = 0;
  if (strategy == ORC_MOD) {
    for (int64_t id = shard; id < vocab; id += n_shards) ++n;
    return n;
  }
  return vocab / n_shards + (shard < vocab % n_shards ? 1 : 0);
}

/* out[i,:] = table_shard[row(ids[i]),:] for the ids that live on `shard`; other rows of
 * `out` are left alone.  Returns the number of ids outside [0, vocab). */
int64_t orc_sparse_gather(const float* table, int64_t vocab, int dim, int n_shards, int shard,
                          int strategy, const int64_t* ids, int64_t n, float* out) {
  int64_t bad = 0;
  for (int64_t i = 0; i < n; ++i) {
    int64_t id = ids[i];
    if (id < 0 || id >= vocab) { ++bad; continue; }
    int sh; int64_t row;
    orc_locate(vocab, n_shards, strategy, id, &sh, &row);
    if (sh != shard) continue;
    memcpy(out + i * (int64_t)dim, table + row * (int64_t)dim, sizeof(float) * (size_t)dim);
  }
  return bad;
}

/* [TF1.3] GradientDescentOptimizer sparse apply = scatter_sub(var, ids, lr * rows); the CPU
 * scatter kernel visits i = 0..n-1 in order, so duplicate ids subtract one after another. */
void orc_sparse_scatter_sub(float* table, int64_t vocab, int dim, int n_shards, int shard,
                            int strategy, const int64_t* ids, int64_t n, const float* rows,
                            float scale) {
  for (int64_t i = 0; i < n; ++i) {
    int64_t id = ids[i];
    if (id < 0 || id >= vocab) continue;
    int sh; int64_t row;
    orc_locate(vocab, n_shards, strategy, id, &sh, &row);
    if (sh != shard) continue;
    float* t = table + row * (int64_t)dim;
    const float* g = rows + i * (int64_t)dim;
    for (int c = 0; c < dim; ++c) {
      float d = scale * g[c];
      t[c] = t[c] - d;
    }
  }
}
