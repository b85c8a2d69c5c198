/* libtfgraft -- C ABI of the B200-native synchronous parameter-server step.
 *
 * The reference (mitake/k8s, the tensorflow/k8s TfJob operator) has no FFI of its own
 * (SURVEY.md §8b): its hot path sits behind three de-facto contracts -- the TfJob
 * document (B1), the TF_CONFIG / --cluster_spec launch contract (B2) and the container
 * exit code (B3).  Each entry point below names the reference interface it stands in
 * for.  Arithmetic semantics that the reference delegates to TensorFlow 1.3.0 (pinned
 * only as an image tag, pkg/spec/tf_job.go:87) are marked [TF1.3].
 *
 * Conventions
 *   - every function returns 0 on success, else an exit-code-compatible int
 *     (pkg/trainer/training.go:203-238, README.md:198-202):
 *       1..127   permanent  (bad TF_CONFIG / invalid argument / wrong call order)
 *       128..255 retryable  (CUDA / driver / peer-mapping failures)
 *     so a replica process can `exit(rc)` and the trainer classifies it unchanged.
 *   - plain pointers and sizes only.  "dev" pointers are CUDA device pointers of the
 *     current device; "host" pointers are ordinary (ideally pinned) host memory.
 *   - `stream` is a cudaStream_t passed as void* (NULL = the legacy default stream).
 *   - one tfg_job is driven by one host thread at a time (the reference drives a job
 *     from one reconcile goroutine, pkg/trainer/training.go:412-456).
 *   - nothing here falls back to the CPU: without a usable CUDA device every compute
 *     entry point fails with a retryable code.
 */
#ifndef TFGRAFT_H_
#define TFGRAFT_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- return codes (B3: pkg/trainer/training.go:220-237) ------------------------- */
#define TFG_OK            0
#define TFG_EINVAL        2   /* permanent: invalid argument                          */
#define TFG_ESPEC         3   /* permanent: malformed TF_CONFIG / cluster spec        */
#define TFG_ESTATE        4   /* permanent: call order (e.g. step before build)       */
#define TFG_ELIMIT        5   /* permanent: size beyond a documented maximum          */
#define TFG_ECUDA       130   /* retryable: CUDA runtime/driver error                 */
#define TFG_ENOMEM      131   /* retryable: device or host allocation failed          */
#define TFG_EPEER       132   /* retryable: peer mapping / cross-rank barrier timeout */

/* ---- enums ------------------------------------------------------------------------ */
/* [TF1.3] training_ops: ApplyGradientDescent / ApplyMomentum(use_nesterov)            */
#define TFG_OPT_SGD        0
#define TFG_OPT_MOMENTUM   1
#define TFG_OPT_NESTEROV   2
/* physical layout of the flat parameter space across GPU ranks        */
#define TFG_LAYOUT_LOGICAL  0  /* a variable lives whole on the GPU of its PS task   */
#define TFG_LAYOUT_BALANCED 1  /* flat space cut into n_gpus equal 128-B aligned parts */
/* arithmetic of the model's contractions                                              */
#define TFG_MATH_FP32      0   /* CUDA-core FFMA, fp32 everywhere                      */
#define TFG_MATH_BF16      1   /* tcgen05 kind::f16, bf16 operands, fp32 accumulate    */
#define TFG_MATH_TF32      2   /* tcgen05 kind::tf32, fp32 storage                     */
#define TFG_MATH_TF32X3    3   /* tcgen05 kind::tf32, three MMAs per product (hi*hi +  */
                               /* hi*lo + lo*hi, fp32 accumulate): fp32-equivalent,    */
                               /* held to the fp32 parity bounds                       */
/* [TF1.3] embedding_lookup partition_strategy                                         */
#define TFG_PART_MOD       0
#define TFG_PART_DIV       1
/* dtypes of tfg_var_register                                                          */
#define TFG_DT_F32         0
#define TFG_DT_I64         1
/* buffers of tfg_job_buffer / tfg_job_ipc_export                                      */
#define TFG_BUF_PARAMS     0
#define TFG_BUF_GRADS      1   /* + local replica slot in `slot`                       */
#define TFG_BUF_ACCUM      2
#define TFG_BUF_LOSS       3
#define TFG_BUF_FLAGS      4

#define TFG_MAX_REPLICAS  64   /* training replicas (MASTER + WORKER) per job          */
#define TFG_MAX_GPUS      16
#define TFG_IPC_HANDLE_BYTES 64

typedef struct tfg_job tfg_job;

/* ---- library ---------------------------------------------------------------------- */
int tfg_version(void);
/* number of CUDA devices; replaces "how many nvidia.com/gpu does the node offer"
 * (the signal ConfigureAccelerators keys on, pkg/spec/tf_job.go:189-199). */
int tfg_device_count(int* n_out);
/* last error text: of `job`, or of the calling thread when job == NULL. */
const char* tfg_last_error(const tfg_job* job);

/* ---- job lifecycle (B2) ------------------------------------------------------------
 * tf_config_json is the TF_CONFIG value the trainer builds per replica
 * (pkg/trainer/replicas.go:189-203): {"cluster":{job:[host:port...]},"task":{...},
 * "environment":"cloud"}.  Only "cluster" is interpreted; host strings are rank labels.
 * Training replicas are master tasks then worker tasks in index order; replica r runs
 * on GPU r % n_gpus, PS task k is owned by GPU k % n_gpus.  gpu_rank is this process's
 * GPU rank in [0, n_gpus).  Touches no CUDA state (usable on a CPU-only host for
 * placement queries).  Replaces: tf.train.Server(server_def) construction,
 * examples/tf_sample/tf_sample/tf_smoke.py:101-112 and
 * grpc_tensorflow_server/grpc_tensorflow_server.py:93-112. */
int tfg_job_create(const char* tf_config_json, int gpu_rank, int n_gpus, tfg_job** out);
/* idempotent; replaces TFReplicaSet.Delete, pkg/trainer/replicas.go:299-356 */
int tfg_job_destroy(tfg_job* job);
/* bind to CUDA device `cuda_device` and run all work on `stream` (NULL: own stream). */
int tfg_job_bind(tfg_job* job, int cuda_device, void* stream);

/* cluster queries: ClusterSpec(), pkg/trainer/training.go:114-128 */
int tfg_cluster_num_tasks(const tfg_job* job, const char* job_name); /* -1 if absent   */
int tfg_num_replicas(const tfg_job* job);
int tfg_num_ps(const tfg_job* job);
int tfg_replica_gpu(const tfg_job* job, int replica);
int tfg_ps_gpu(const tfg_job* job, int ps_task);
int tfg_num_local_replicas(const tfg_job* job);
int tfg_local_replica(const tfg_job* job, int slot);   /* global replica id of a slot  */

/* Parse the default-PS command-line form "job|h:p;h:p,job|..." into TF_CONFIG's
 * cluster JSON ({"job":["h:p",...],...}, jobs in order of appearance).  Same grammar and
 * failures as parse_cluster_spec, grpc_tensorflow_server/grpc_tensorflow_server.py:46-88
 * (the inverse of transformClusterSpecForDefaultPS, pkg/trainer/replicas.go:102-122).
 * Returns TFG_ESPEC with the reference's message in tfg_last_error(NULL). */
int tfg_parse_cluster_spec(const char* cluster_spec, char* json_out, size_t json_cap);

/* ---- variables and placement (A11) -------------------------------------------------
 * [TF1.3] tf.train.replica_device_setter (call site tf_smoke.py:116-118): an op whose
 * type is a PS op ("Variable", "VariableV2", "VarHandleOp", ...) goes to
 * /job:ps/task:k with k a round-robin counter that advances once per such op in
 * creation order; with no ps job everything stays on the worker device (k = -1).
 * Registration order == creation order.  Returns the variable index in *var_out. */
int tfg_var_register(tfg_job* job, const char* name, const char* op_type, int dtype,
                     const int64_t* shape, int rank, int trainable,
                     int* var_out, int* ps_task_out);
int tfg_num_vars(const tfg_job* job);
/* name (NUL-terminated, truncated to name_cap), shape (up to 8 dims), rank and trainable flag
 * of a registered variable -- what a model-building host needs to seed weights by name. */
int tfg_var_info(const tfg_job* job, int var, char* name_out, size_t name_cap, int64_t* shape_out,
                 int* rank_out, int* trainable_out);
/* allocate the flat parameter / gradient / slot buffers in HBM for the registered
 * trainable fp32 variables.  After this no variable can be registered. */
int tfg_vars_finalize(tfg_job* job, int layout);
/* placement of one variable: its PS task, the GPU rank that applies its update, its
 * element offset and count in the flat space. */
int tfg_placement(const tfg_job* job, int var, int* ps_task, int* gpu,
                  int64_t* offset, int64_t* count);
int64_t tfg_flat_count(const tfg_job* job);           /* padded length of flat space   */
/* host <-> device copies of one variable (seed weights in, parameters out) */
int tfg_var_set(tfg_job* job, int var, const void* host, size_t bytes);
int tfg_var_get(tfg_job* job, int var, void* host, size_t bytes);
int tfg_accum_get(tfg_job* job, int var, void* host, size_t bytes);
/* write a local replica's flat gradient of one variable (tests, PS sweep) */
int tfg_grad_set(tfg_job* job, int slot, int var, const void* host, size_t bytes);
int tfg_grad_get(tfg_job* job, int slot, int var, void* host, size_t bytes);
/* raw device pointer of a job buffer (slot only for TFG_BUF_GRADS) */
int tfg_job_buffer(tfg_job* job, int which, int slot, void** dev_out, size_t* bytes_out);
int tfg_opt_config(tfg_job* job, int kind, float lr, float momentum);
int64_t tfg_global_step(const tfg_job* job);

/* ---- checkpoint of the sharded variables (SURVEY.md par. 8f row 3) -------------------
 * The reference's operator keeps no training state (its only resume hook is "setup is skipped
 * if Phase != None", pkg/trainer/training.go:252-257); TF 1.3's Saver would write the PS
 * variables.  Here a GPU rank reads / writes raw ranges of the flat fp32 space: the element
 * ranges whose update it applies (tfg_owned_ranges: offsets / counts of up to `cap` ranges,
 * *n_out = how many there are), parameters or optimizer slots (which = TFG_BUF_PARAMS /
 * TFG_BUF_ACCUM).  A write to the parameters also refreshes the operand-precision shadow.
 * k8s_b200/checkpoint.py builds the file format (one flat file per shard + placement table). */
int tfg_owned_ranges(const tfg_job* job, int64_t* offsets_out, int64_t* counts_out, int cap, int* n_out);
int tfg_flat_read(tfg_job* job, int which, int64_t offset, int64_t count, void* host);
int tfg_flat_write(tfg_job* job, int which, int64_t offset, int64_t count, const void* host);
int tfg_global_step_set(tfg_job* job, int64_t step);

/* ---- cross-rank peer mapping (one process per GPU) ---------------------------------
 * Each rank exports CUDA IPC handles of its buffers; the host exchanges the blobs
 * (torch.distributed all_gather) and every rank opens every peer's.  Replaces the
 * per-task Service DNS names the reference uses for addressing
 * (pkg/trainer/replicas.go:157-186). */
int tfg_job_ipc_export(tfg_job* job, void* blob_out, size_t cap, size_t* bytes_out);
int tfg_job_ipc_open(tfg_job* job, int peer_gpu_rank, const void* blob, size_t bytes);
/* single-process stand-in used by 1-GPU tests: treat `peer` job (same process, same
 * device) as GPU rank peer_gpu_rank of this job. */
int tfg_job_attach_local_peer(tfg_job* job, int peer_gpu_rank, tfg_job* peer);
/* stream-ordered barrier across all GPU ranks of the job (flags in peer memory). */
int tfg_job_barrier(tfg_job* job);

/* ---- the synchronous PS update (A12 + A13 + A14) -----------------------------------
 * One fused kernel per rank: for every element this rank owns, read all replicas'
 * gradients (peer loads over NVLink, replica order), mean = sum / n_replicas
 * ([TF1.3] ConditionalAccumulator take_grad), apply the optimizer ([TF1.3]
 * ApplyGradientDescent: var -= lr*g ; ApplyMomentum: acc = mu*acc + g,
 * var -= lr*acc ; Nesterov: var -= lr*g + lr*mu*acc, no FMA contraction), and store the
 * new value into every rank's parameter buffer (peer stores = the pull).  global_step
 * advances by one.  with_barriers != 0 brackets the kernel with tfg_job_barrier.
 * Replaces: gRPC RecvTensor push + accumulator + apply + variable Recv of the TF
 * runtime that grpc_tensorflow_server.py:112-115 hosts. */
int tfg_push_apply_pull(tfg_job* job, int with_barriers);
/* split variant for the NCCL baseline: apply on this rank's owned ranges from an
 * already-reduced (summed) gradient in local replica slot 0, writing only locally. */
int tfg_apply_owned(tfg_job* job, int n_summed);
/* the all-reduce baseline: the gradient in local replica slot 0 already holds the sum over all
 * n_summed replicas (e.g. after an NCCL all-reduce); every rank applies the whole flat space
 * to its own copy, no pull.  bench.py --exchange nccl times this against the fused kernel. */
int tfg_apply_all(tfg_job* job, int n_summed);

/* stateless form of the same kernel over caller buffers (PS shard sweep, tests):
 * var/accum [count] fp32 updated in place from n_grads gradient buffers (host array of
 * device pointers, summed in array order, divided by `divisor`), result also stored to
 * n_pull extra destinations.  accum may be NULL for TFG_OPT_SGD. */
int tfg_dense_apply(int kind, float* var_dev, float* accum_dev,
                    const float* const* grads_dev, int n_grads, float divisor,
                    int64_t count, float lr, float momentum,
                    float* const* pull_dev, int n_pull, void* stream);

/* ---- sparse embedding path (A15) ---------------------------------------------------
 * A table of `vocab` rows x `dim` fp32 is partitioned over n_shards PS tasks.
 * [TF1.3] partition_strategy "mod": shard = id % n_shards, local row = id / n_shards;
 * "div": contiguous ranges, the first vocab % n_shards shards hold one extra row.
 * `table_dev` is shard `shard`'s rows.  ids are int64 global ids, n of them. */
int64_t tfg_sparse_shard_rows(int64_t vocab, int n_shards, int shard, int strategy);
/* out[i,:] = table[local(ids[i]),:] for ids of this shard; rows of other shards are
 * left untouched (the owner of that shard fills them). Out-of-range ids -> TFG_EINVAL
 * is reported through *bad_dev (count of ids outside [0,vocab)) when non-NULL. */
int tfg_sparse_gather(const float* table_dev, int64_t vocab, int dim, int n_shards,
                      int shard, int strategy, const int64_t* ids_dev, int64_t n,
                      float* out_dev, int* bad_dev, void* stream);
/* [TF1.3] scatter_sub(var, ids, lr*rows): table[local(ids[i]),:] -= scale*rows[i,:],
 * duplicates applied one after another in order of i (bit-exact with the sequential
 * CPU loop).  Needs scratch_dev of at least tfg_sparse_scratch_bytes(n) bytes.
 * With tfg_sparse_scratch_bytes_fast(n, dim) bytes or more, rows hit by more than 64 ids
 * (heavy-tailed id distributions) are handled chunk-parallel instead of by one warp: their
 * updates are summed in blocks of 128 ids, in order, the block sums in 16 contiguous segments, in order,
 * and the segment sums subtracted in order -- deterministic, as accurate as the one-by-one loop, but not
 * bit-identical to it.  Rows hit
 * by at most 64 ids are bit-exact either way. */
size_t tfg_sparse_scratch_bytes(int64_t n);
size_t tfg_sparse_scratch_bytes_fast(int64_t n, int dim);
int tfg_sparse_scatter_sub(float* table_dev, int64_t vocab, int dim, int n_shards,
                           int shard, int strategy, const int64_t* ids_dev, int64_t n,
                           const float* rows_dev, float scale,
                           void* scratch_dev, size_t scratch_bytes, void* stream);

/* The sparse path across the GPU ranks of a box (SURVEY.md par. 8e).  Shard k of a table is owned by
 * the GPU of PS task k (tfg_ps_gpu).  With every rank's buffers mapped into every other rank
 * (tfg_ipc_export / tfg_ipc_open) neither direction needs a collective:
 *   - pull: a worker fetches the rows of ALL its ids in one launch, each straight from its owner's
 *     shard (table_shards_dev: host array of n_shards device pointers, own shards local, others peer);
 *   - push + apply: the owner of a shard applies [TF1.3] scatter_sub over the concatenation of all
 *     n_sources replicas' batches (source after source, n_per_source ids each), reading the ids and the
 *     update rows in place from the replicas' buffers (host arrays of device pointers).  Duplicates are
 *     applied in (source, position) order, so the result is bit-identical to the sequential loop over
 *     the concatenated batch.  scratch: tfg_sparse_scratch_bytes(n_sources * n_per_source).
 * Replaces: embedding_lookup's per-shard gathers over gRPC and the IndexedSlices push to each PS
 * (hosted by grpc_tensorflow_server.py:112-115). */
int tfg_sparse_gather_multi(const float* const* table_shards_dev, int64_t vocab, int dim, int n_shards,
                            int strategy, const int64_t* ids_dev, int64_t n, float* out_dev, int* bad_dev,
                            void* stream);
int tfg_sparse_scatter_sub_multi(float* table_dev, int64_t vocab, int dim, int n_shards, int shard,
                                 int strategy, const int64_t* const* ids_devs, const float* const* rows_devs,
                                 int n_sources, int64_t n_per_source, float scale, void* scratch_dev,
                                 size_t scratch_bytes, void* stream);
/* CUDA-IPC handle (TFG_IPC_HANDLE_BYTES) of a caller-owned buffer, and its mapping in another process
 * of the same box; replaces the per-task Service DNS names (pkg/trainer/replicas.go:157-186) for
 * buffers the job does not own itself.  dev_ptr must be the START of a cudaMalloc allocation (a handle
 * names the whole allocation): tfg_dev_alloc / tfg_dev_free give such buffers (zero-filled) to hosts
 * whose own allocator sub-allocates. */
int tfg_dev_alloc(size_t bytes, void** dev_ptr_out);
int tfg_dev_free(void* dev_ptr);
int tfg_ipc_export(const void* dev_ptr, void* handle_out);
int tfg_ipc_open(const void* handle, void** dev_ptr_out);
int tfg_ipc_close(void* dev_ptr);

/* Scatter-sub for callers that cannot allocate device memory.  No counterpart in the
 * reference: this entry point exists only for this library's users.
 *
 * tfg_sparse_scatter_sub needs a scratch buffer, which somebody has to cudaMalloc.  Some
 * callers cannot: a PS process that shares its GPU with a framework whose allocator has
 * already claimed all device memory (the reference's default PS starts TensorFlow with
 * per_process_gpu_memory_fraction = 1.0, so a later cudaMalloc in that process fails), or
 * a step loop that must not take the device-wide synchronisation cudaMalloc/cudaFree
 * imply.  For them the scratch lives inside the library image, in a fixed device array of
 * TFG_SPARSE_STATIC_SCRATCH_BYTES (about 1.15 MiB) that every GPU using the library keeps
 * until the process exits.  The caller still passes table, ids and rows; results are
 * identical to tfg_sparse_scatter_sub given tfg_sparse_scratch_bytes(n) of scratch (every
 * duplicate applied in order).
 *   - n > TFG_SPARSE_STATIC_MAX_IDS is rejected with TFG_ELIMIT (split the batch: applying
 *     consecutive slices of ids in order gives the same table).
 *   - The first call on each device looks up the array's device address, once, with that
 *     device current; make that first call before starting any stream capture.
 *   - The call itself performs no cudaMalloc / cudaFree and calls no allocating library.
 *   - All calls on a device share the one array and it keeps the previous call's contents.
 *     Nothing in it has to start at zero (every word a call reads it has written first),
 *     but two calls that overlap on the GPU -- different streams, or different host
 *     threads -- race on it: issue them on one stream, or serialise them yourself.  (The
 *     address look-up itself is safe from several host threads.) */
#define TFG_SPARSE_STATIC_MAX_IDS 65536
#define TFG_SPARSE_STATIC_SCRATCH_BYTES 1204480
int tfg_sparse_scatter_sub_static(float* table_dev, int64_t vocab, int dim, int n_shards,
                                  int shard, int strategy, const int64_t* ids_dev, int64_t n,
                                  const float* rows_dev, float scale, void* stream);

/* ---- the sample model (A16) and the full step --------------------------------------
 * The reference's examples name an image (tf_sample) whose only arithmetic is the smoke
 * multiply below; the MNIST CNN is build-defined with the canonical mnist_deep shapes
 * (SURVEY.md §0.2, §8a A16).  Registers W_conv1,b_conv1,W_conv2,b_conv2,W_fc1,b_fc1,
 * W_fc2,b_fc2,global_step in that creation order and allocates activations for
 * batch_per_replica examples per local replica. */
int tfg_model_mnist_cnn(tfg_job* job, int batch_per_replica, int math);
/* ResNet-50 v1.5 for BASELINE config 4 (tf_cnn_benchmarks-style synthetic ImageNet): bf16
 * activations and operands on the tcgen05 path, fp32 master weights, batch-norm with batch
 * statistics.  image_hw: input height = width, a multiple of 32 (224 for ImageNet); classes a
 * multiple of 8 (1000).  Registers conv / bn / fc variables in creation order (moving
 * statistics included, as non-trainable PS ops) and allocates one replica's activations per
 * local slot.  Inputs of tfg_forward_backward / tfg_step_dense are then images
 * [B,image_hw,image_hw,3] fp32 NHWC and labels [B] int32.  The model is build-defined: the
 * reference tree holds no model code (SURVEY.md par. 0.2). */
int tfg_model_resnet50(tfg_job* job, int batch_per_replica, int image_hw, int classes);
/* Host copy (fp32) of an activation of the last forward of `slot`, by name ("conv1/z",
 * "stage2/block1/c/y", "pool1", ...), or of its gradient with a "grad:" prefix; parity tests. */
int tfg_tensor_get(tfg_job* job, int slot, const char* name, float* host, size_t cap_floats, size_t* n_out);

/* forward + backward of local replica `slot` on device-resident inputs:
 * images [B,28,28,1] fp32 NHWC, labels [B] int32.  Gradients land in the replica's
 * gradient buffer, the mean loss in the job's loss buffer. */
int tfg_forward_backward(tfg_job* job, int slot, const float* images_dev,
                         const int32_t* labels_dev);
/* logits of the last forward of a slot, [B,10] fp32, to host (parity tests) */
int tfg_logits_get(tfg_job* job, int slot, float* host, size_t bytes);
/* One synchronous step through the reference-facing boundary with HOST buffers:
 * copies each local replica's images/labels host->device, runs forward/backward,
 * push+apply+pull across all ranks, and returns the mean loss over all replicas
 * (device->host) in *loss_out.  images_host[slot], labels_host[slot].
 * Replaces one Session.run(train_op) of a between-graph sync-PS replica
 * (tf_smoke.py:71-76 is the reference's only Session.run call site). */
int tfg_step_dense(tfg_job* job, const float* const* images_host,
                   const int32_t* const* labels_host, float* loss_out);
/* same, inputs already resident in HBM (the kernel-only timing arm); loss stays on the
 * device unless loss_out != NULL. */
int tfg_step_dense_dev(tfg_job* job, const float* const* images_dev,
                       const int32_t* const* labels_dev, float* loss_out);
/* loss of the last forward of global replica `replica` hosted here, or (replica == -1) the
 * mean over all replicas computed by the last tfg_push_apply_pull; synchronises the stream. */
int tfg_loss_get(tfg_job* job, int replica, float* loss_out);
int tfg_job_sync(tfg_job* job);
/* kernels launched by this job since creation (bench.py's gpu_launches) */
int64_t tfg_launch_count(const tfg_job* job);

/* The tensor-path GEMM alone (parity tests, profiling): C[M,N] fp32 (dense, ldc = N) =
 * A(M x K) * B(N x K)^T in `math` precision (TFG_MATH_BF16: bf16 elements, TFG_MATH_TF32:
 * fp32 elements).  An operand is K-major (stored [rows, K], x_mn_major = 0) or MN-major
 * (stored [K, rows], x_mn_major = 1); lda/ldb = row pitch of the stored matrix in elements
 * (a multiple of 16 bytes).  tcgen05 tile 128 x block_n (64 or 128).  splits > 1 = split-K
 * through partial_dev (splits*M*N floats), reduced in a fixed order.  bias_dev[N] / relu are
 * applied when splits == 1.  The contractions of the sample model (A16) are instances of it. */
int tfg_gemm_tc(int math, const void* a_dev, int a_mn_major, const void* b_dev, int b_mn_major,
                float* c_dev, int M, int N, int K, int64_t lda, int64_t ldb, int block_n, int splits,
                float* partial_dev, const float* bias_dev, int relu, void* stream);

/* The same GEMM as a 3xTF32 split product (math mode TFG_MATH_TF32X3): fp32 operands plus their
 * low-order planes (same shape and pitch; lo = rn_tf32(x - trunc_tf32(x)), produced by
 * tfg_tf32_split_lo or by the kernel that wrote x), three kind::tf32 MMAs per K step
 * (a_lo*b + a*b_lo + a*b).  The tensor core adds into its accumulator with truncation, so K is cut
 * into chunks of chunk_kb K blocks (32 fp32 each; 0 = the default 2) that accumulate in TMEM from
 * zero and are summed in registers with round-to-nearest: the result carries fp32 summation-level
 * error at tensor-core speed.  This is how the fp32 sample model of BASELINE configs 1-3
 * (SURVEY.md par. 8d "which roofline") runs its contractions. */
int tfg_gemm_tc_x3(const float* a_dev, const float* a_lo_dev, int a_mn_major, const float* b_dev,
                   const float* b_lo_dev, int b_mn_major, float* c_dev, int M, int N, int K, int64_t lda,
                   int64_t ldb, int block_n, int splits, int chunk_kb, float* partial_dev,
                   const float* bias_dev, int relu, void* stream);
int tfg_tf32_split_lo(const float* src_dev, float* lo_dev, int64_t n, void* stream);

/* per-kernel device time of this job's launches, measured with CUDA events on the job's
 * stream (bench.py's roofline leg).  enable(1) starts a fresh record; report() writes
 * {"kernel":{"count":n,"ms":total},...} and starts over. */
int tfg_profile_enable(tfg_job* job, int on);
int tfg_profile_report(tfg_job* job, char* json_out, size_t json_cap);

/* ---- smoke workload (A9) -----------------------------------------------------------
 * For every task of every job in the cluster: c = a * b with
 * a = b = arange(100).reshape(10,10), int32 (tf_smoke.py:53-60).  out_host receives
 * n_tasks x 100 int32 in cluster order (jobs sorted by name, tasks by index); tasks
 * mapped to other GPU ranks are computed by those ranks and left zero here unless
 * all_tasks != 0.  Replaces: run(), examples/tf_sample/tf_sample/tf_smoke.py:34-76. */
int tfg_smoke(tfg_job* job, int all_tasks, int32_t* out_host, size_t out_cap_elems,
              int* n_tasks_out);

#ifdef __cplusplus
}
#endif
#endif /* TFGRAFT_H_ */
