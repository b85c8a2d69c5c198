// Source only: Go is not installed in the image this repository is built in, so this file has
// never been compiled or run.  It shows the binding a maintainer of tensorflow/k8s would add to
// pkg/trainer so that TFReplicaSet.Create (pkg/trainer/replicas.go:124-271 of the reference) starts
// GPU ranks of the local box through libtfgraft's C ABI (include/tfgraft.h) instead of creating a
// Service and a batch/v1 Job per replica.  Every entry point it calls is exercised through the same
// C ABI from Python ctypes (k8s_b200/lib.py) by tests/.
package trainer

/*
#cgo CFLAGS: -I${SRCDIR}/../../../include
#cgo LDFLAGS: -L${SRCDIR}/../../../k8s_b200/csrc -ltfgraft
#include <stdlib.h>
#include "tfgraft.h"
*/
import "C"

import (
	"encoding/json"
	"errors"
	"fmt"
	"runtime"
	"strings"
	"unsafe"
)

// GpuRank is one process-per-GPU rank of a TfJob: the handle libtfgraft gives back for the
// TF_CONFIG the operator would have put into the replica's container environment.
type GpuRank struct {
	job  *C.tfg_job
	Rank int
}

// lastError turns the library's message for the last failure into a Go error.  rc keeps the exit
// code contract of pkg/trainer/training.go:203-238 (0 success, 1..127 permanent, >= 128 retryable),
// so isRetryableTerminationState needs no change.
func lastError(job *C.tfg_job, rc C.int) error {
	return fmt.Errorf("libtfgraft rc=%d: %s", int(rc), C.GoString(C.tfg_last_error(job)))
}

// CreateOnGpuRanks is the GPU-rank twin of TFReplicaSet.Create: the same TF_CONFIG document
// (replicas.go:189-203: cluster, task {type, index}, environment "cloud"), no Service / Job objects.
// rank is this process's GPU rank, nGpus the number of ranks on the box.
func (s *TFReplicaSet) CreateOnGpuRanks(index int32, rank int, nGpus int) (*GpuRank, error) {
	tfConfig := TfConfig{
		Cluster: s.Job.ClusterSpec(),
		Task: TaskSpec{
			Type:  strings.ToLower(string(s.Spec.TfReplicaType)),
			Index: int(index),
		},
		Environment: "cloud",
	}
	tfConfigJson, err := json.Marshal(tfConfig)
	if err != nil {
		return nil, err
	}
	cs := C.CString(string(tfConfigJson))
	defer C.free(unsafe.Pointer(cs))

	var job *C.tfg_job
	if rc := C.tfg_job_create(cs, C.int(rank), C.int(nGpus), &job); rc != C.TFG_OK {
		return nil, lastError(nil, rc)
	}
	if rc := C.tfg_job_bind(job, C.int(rank%nGpus), nil); rc != C.TFG_OK {
		err := lastError(job, rc)
		C.tfg_job_destroy(job)
		return nil, err
	}
	return &GpuRank{job: job, Rank: rank}, nil
}

// BuildMnist registers the variables of the tf_sample CNN under round-robin placement over the
// job's PS tasks and allocates them (balanced = the physical layout that spreads bytes evenly).
func (r *GpuRank) BuildMnist(batchPerReplica int, balanced bool) error {
	// TFG_MATH_TF32X3: the fp32 model on the tensor cores (3xTF32 split products), held to the fp32 bounds
	if rc := C.tfg_model_mnist_cnn(r.job, C.int(batchPerReplica), C.TFG_MATH_TF32X3); rc != C.TFG_OK {
		return lastError(r.job, rc)
	}
	layout := C.int(0)
	if balanced {
		layout = 1
	}
	if rc := C.tfg_vars_finalize(r.job, layout); rc != C.TFG_OK {
		return lastError(r.job, rc)
	}
	return nil
}

// Optimizer selects SGD (0), momentum (1) or Nesterov momentum (2) and its constants; it may be called
// between steps (a learning-rate schedule).
func (r *GpuRank) Optimizer(kind int, lr float32, momentum float32) error {
	if rc := C.tfg_opt_config(r.job, C.int(kind), C.float(lr), C.float(momentum)); rc != C.TFG_OK {
		return lastError(r.job, rc)
	}
	return nil
}

// ExportPeerHandles returns this rank's CUDA-IPC blob.  With nGpus > 1 every rank must hand its blob
// to every other rank (the launcher's rendezvous does the exchange) and call ConnectPeer for each
// before the first Step; without it Step fails with TFG_ESTATE ("GPU rank g is not mapped").
func (r *GpuRank) ExportPeerHandles() ([]byte, error) {
	buf := make([]byte, 8+4*C.TFG_MAX_REPLICAS+C.TFG_IPC_HANDLE_BYTES*(4+C.TFG_MAX_REPLICAS))
	var n C.size_t
	if rc := C.tfg_job_ipc_export(r.job, unsafe.Pointer(&buf[0]), C.size_t(len(buf)), &n); rc != C.TFG_OK {
		return nil, lastError(r.job, rc)
	}
	return buf[:int(n)], nil
}

// ConnectPeer maps the buffers of GPU rank `peer` (its blob from ExportPeerHandles) into this rank.
func (r *GpuRank) ConnectPeer(peer int, blob []byte) error {
	if len(blob) == 0 {
		return errors.New("empty peer blob")
	}
	if rc := C.tfg_job_ipc_open(r.job, C.int(peer), unsafe.Pointer(&blob[0]), C.size_t(len(blob))); rc != C.TFG_OK {
		return lastError(r.job, rc)
	}
	return nil
}

// Step runs one synchronous step for the replicas this rank hosts: images[i] / labels[i] are the
// host batches of local replica i.  Returns the mean loss over all replicas of the job.
func (r *GpuRank) Step(images [][]float32, labels [][]int32) (float32, error) {
	if len(images) == 0 || len(images) != len(labels) {
		return 0, errors.New("one image batch and one label batch per local replica")
	}
	n := len(images)
	// cgo's pointer-passing rules: a Go pointer may be stored in C memory only while the Go object is
	// pinned.  The batches are pinned for the duration of the call (runtime.Pinner, Go 1.21+), the two
	// arrays of pointers live in C memory, and the C side keeps nothing after it returns.
	var pin runtime.Pinner
	defer pin.Unpin()
	ip := (*[1 << 20]*C.float)(C.malloc(C.size_t(n) * C.size_t(unsafe.Sizeof(uintptr(0)))))
	lp := (*[1 << 20]*C.int32_t)(C.malloc(C.size_t(n) * C.size_t(unsafe.Sizeof(uintptr(0)))))
	defer C.free(unsafe.Pointer(ip))
	defer C.free(unsafe.Pointer(lp))
	for i := 0; i < n; i++ {
		if len(images[i]) == 0 || len(labels[i]) == 0 {
			return 0, errors.New("empty batch")
		}
		pin.Pin(&images[i][0])
		pin.Pin(&labels[i][0])
		ip[i] = (*C.float)(unsafe.Pointer(&images[i][0]))
		lp[i] = (*C.int32_t)(unsafe.Pointer(&labels[i][0]))
	}
	var loss C.float
	if rc := C.tfg_step_dense(r.job, &ip[0], &lp[0], &loss); rc != C.TFG_OK {
		return 0, lastError(r.job, rc)
	}
	return float32(loss), nil
}

// Close releases the rank's device memory.
func (r *GpuRank) Close() {
	if r.job != nil {
		C.tfg_job_destroy(r.job)
		r.job = nil
	}
}
