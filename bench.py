#!/usr/bin/env python
"""bench.py -- examples/sec of the synchronous PS step (BASELINE.json metric) on N B200s.

    python bench.py --gpus 1 --steps K --warmup W            # this repo's CUDA path
    torchrun ... bench.py --gpus N --steps K --warmup W       # one rank per GPU
    python bench.py --impl reference ...                      # CPU arm (oracle port, see below)

A step = forward + backward of the sample MNIST CNN on every training replica, gradient
push + mean + optimizer apply on the shard owners, parameter pull.  The
workload is a TfJob document built the way examples/tf_job_gpu.yaml (N = 1: one MASTER, no
PS, no exchange) and BASELINE config 3 (N > 1: N replicas + N PS shards) describe it, run
through this repo's trainer to obtain each rank's TF_CONFIG.

Prints ONE JSON line (rank 0).  See DESIGN.md "Measurement" for what each key means.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

BATCH = 128          # examples per replica per step
N_INPUT_SETS = 8     # distinct resident input batches cycled through
FLOP_PER_EXAMPLE_FWD = 2 * (28 * 28 * 32 * 25 + 14 * 14 * 64 * 800 + 3136 * 1024 + 1024 * 10)


def parse_args(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="graft", choices=["graft", "reference"])
    ap.add_argument("--model", default="mnist", choices=["mnist", "resnet50"],
                    help="mnist = tf_sample CNN (BASELINE configs 1-3, the default headline); "
                         "resnet50 = BASELINE config 4 (synthetic ImageNet 224x224, bf16)")
    ap.add_argument("--batch", type=int, default=0, help="examples per replica per step (default 128 / 256)")
    ap.add_argument("--image", type=int, default=224, help="resnet50: image height = width")
    ap.add_argument("--classes", type=int, default=1000, help="resnet50: number of classes")
    ap.add_argument("--math", default=os.environ.get("TFGRAFT_MATH", "auto"),
                    choices=["auto", "tf32x3", "fp32", "bf16", "tf32"],
                    help="arithmetic of the sample model's contractions; auto = tf32x3 (3xTF32 split products on "
                         "tcgen05, fp32-equivalent: the reference's precision for configs 1-3)")
    ap.add_argument("--layout", default="balanced", choices=["logical", "balanced"])
    ap.add_argument("--opt", default="momentum", choices=["sgd", "momentum", "nesterov"])
    ap.add_argument("--exchange", default="peer", choices=["peer", "nccl"],
                    help="peer = fused push+apply+pull over peer memory (the product); nccl = all-reduce of the flat "
                         "gradient through torch.distributed + full local apply (the baseline it is compared with)")
    ap.add_argument("--ps-params", type=int, default=256 << 20,
                    help="parameters of the PS-update roofline probe (BASELINE config 5)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-resnet", dest="with_resnet", action="store_false",
                    help="skip the ResNet-50 (BASELINE config 4) sub-measurement of the default line")
    a = ap.parse_args(argv)
    if a.batch <= 0:
        a.batch = BATCH if a.model == "mnist" else 256
    return a


def tf_job_doc(n_gpus: int) -> dict:
    """The TfJob document of the benchmark: tf_job_gpu.yaml's shape at N = 1, BASELINE
    config 3's (1 MASTER + N-1 WORKER + N PS) beyond."""
    tmpl = {"spec": {"containers": [{"image": "tf_sample", "name": "tensorflow",
                                     "resources": {"limits": {"nvidia.com/gpu": 1}}}],
                     "restartPolicy": "OnFailure"}}
    specs = [{"tfReplicaType": "MASTER", "template": tmpl}]
    if n_gpus > 1:
        specs.append({"replicas": n_gpus - 1, "tfReplicaType": "WORKER", "template": tmpl})
        specs.append({"replicas": n_gpus, "tfReplicaType": "PS"})
    return {"apiVersion": "tensorflow.org/v1alpha1", "kind": "TfJob",
            "metadata": {"name": "bench-sync-ps"}, "spec": {"RuntimeId": "b200", "replicaSpecs": specs}}


def tf_config_for_rank(n_gpus: int, rank: int) -> str:
    """TF_CONFIG of the training replica hosted by GPU `rank`, produced by the trainer."""
    from k8s_b200 import spec as S
    from k8s_b200 import trainer as T
    be = T.RecordingBackend()
    tj = T.TrainingJob(be, S.tf_job_from_dict(tf_job_doc(n_gpus)))
    tj.reconcile(S.ControllerConfig())
    tj.reconcile(S.ControllerConfig())
    if tj.job.status.phase == S.PHASE_FAILED:
        raise RuntimeError(tj.job.status.reason)
    name = "bench-sync-ps-%s-b200-%d" % (("master", 0) if rank == 0 else ("worker", rank - 1))
    env = [c["env"] for c in be.jobs[name]["spec"]["template"]["spec"]["containers"]][0]
    return [e["value"] for e in env if e["name"] == "TF_CONFIG"][0]


class ClockSampler:
    """SM clock + throttle reasons sampled DURING the timed region: NVML polled every 2 ms from
    a thread of this process (the timed region of a sub-millisecond step is tens of ms long,
    too short for an `nvidia-smi -lms` child to produce samples); nvidia-smi is the fallback."""
    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}

    def __init__(self, index: int):
        self.index, self.sm, self.reasons, self.stop_flag, self.thread = index, [], set(), False, None
        self.max_mhz, self.nvml, self.handle = None, None, None

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[self.index]) if vis and vis.split(",")[self.index].isdigit() else self.index
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
            self.nvml = pynvml
        except Exception:
            self.nvml = None
            return
        self.thread = threading.Thread(target=self._poll, daemon=True)
        self.thread.start()

    def _poll(self):
        n = self.nvml
        while not self.stop_flag:
            try:
                self.sm.append(float(n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM)))
                r = n.nvmlDeviceGetCurrentClocksEventReasons(self.handle)
                for bit, name in self.REASONS.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                break
            time.sleep(0.002)

    def stop(self) -> dict:
        if self.nvml is None:
            return self._smi_once()
        self.stop_flag = True
        self.thread.join(timeout=2)
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.max_mhz,
                "samples": len(self.sm), "reasons": sorted(self.reasons), "source": "nvml, 2 ms poll"}

    def _smi_once(self) -> dict:
        try:
            q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
                 "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
            out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q,
                                  "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=20).stdout
            r = [x.strip() for x in out.strip().split(",")]
            names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
            return {"sm_mhz": float(r[0]), "sm_max_mhz": float(r[1]), "samples": 1,
                    "reasons": [n for n, v in zip(names, r[2:6]) if v.lower().startswith("active")],
                    "source": "nvidia-smi, one sample after the timed region"}
        except Exception:
            return {"sm_mhz": None, "sm_max_mhz": None, "samples": 0, "reasons": ["unavailable"]}


def measured_peaks() -> dict:
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": d["hbm_gbs"], "bf16_tflops": d["bf16_tflops"],
                "bf16_tflops_sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]),
                "source": "measured (MEASURED_PEAKS.json)"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0,
            "source": "fallback (B200_PROFILING.md)"}


LR = 0.001   # SURVEY.md §8d says 0.01; with sigma=0.1 weights and N(0,1) images that overshoots
             # to inf within a few steps, so the timed runs use 0.001 (timing is unaffected)


# ---- CPU arm: the oracle port of the step on the box's host cores ---------------------------
def cpu_step_rate(n_replicas: int, batch: int, opt: str, steps: int, warmup: int):
    """examples/s of the oracle's synchronous step (torch-CPU fp32 forward/backward per
    replica + the C oracle's apply), all host threads.  The only other place besides tests
    and smoke() that executes oracle/."""
    import torch
    from k8s_b200.syncps import mnist_init, synthetic_batch
    from oracle import model_ref as M
    cores = len(os.sched_getaffinity(0))
    torch.set_num_threads(cores)
    params = mnist_init(0)
    accum = [np.zeros_like(p) for p in params]
    sets = [[synthetic_batch(r, batch, s) for r in range(n_replicas)] for s in range(2)]
    for s in range(warmup):
        M.sync_step(params, accum, sets[s % 2], opt, LR, 0.9)
    t0 = time.perf_counter()
    for s in range(steps):
        M.sync_step(params, accum, sets[s % 2], opt, LR, 0.9)
    dt = time.perf_counter() - t0
    return n_replicas * batch * steps / dt, dt / steps * 1e3, cores


def run_reference(a) -> dict:
    """--impl reference.  The reference's own implementation of this path is TensorFlow
    1.3's gRPC PS runtime launched by a Go operator: neither can be built or installed in
    this image (DESIGN.md "Oracle"), so the arm times the oracle port ("kind": "port")."""
    n = a.gpus
    if a.model == "resnet50":
        return run_reference_resnet(a)
    value, ms, cores = cpu_step_rate(n, a.batch, a.opt, a.steps, max(a.warmup, 1))
    return {
        "impl": "reference", "metric": "examples/sec sync-PS step", "value": value, "unit": "examples/s",
        "n_gpus": n, "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": config_dict(a, n, "fp32"),
        "cpu_baseline": {"value": value, "unit": "examples/s", "cores": cores, "kind": "port",
                         "sample": "%d full steps of %d replicas x %d examples (whole workload, "
                                   "oracle/model_ref.py + oracle.c)" % (a.steps, n, a.batch)},
        "e2e": {"value": value, "unit": "examples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }


def cpu_resnet_rate(a, batch: int, steps: int):
    """images/s of the oracle's ResNet-50 forward/backward (fp32 torch-CPU, all host threads) on a
    bounded sample: `steps` steps of `batch` images of the benchmark's shape."""
    import torch
    from oracle import resnet_ref as R
    cores = len(os.sched_getaffinity(0))
    torch.set_num_threads(cores)
    rng = np.random.default_rng(0)
    shapes, cur = {}, {"image": 3, "pool1": 64}
    for name, inp, rk, stride, kout, relu, res in R.units():
        shapes[name + "/weights"] = (rk, rk, cur[inp], kout)
        shapes[name + "/bn/gamma"] = shapes[name + "/bn/beta"] = (kout,)
        cur[name + "/y"] = kout
    shapes["fc/weights"], shapes["fc/biases"] = (2048, a.classes), (a.classes,)
    params = {k: (rng.standard_normal(sh) * (np.sqrt(2.0 / np.prod(sh[:3])) if len(sh) == 4 else 0.01)).astype(np.float32)
              if k.endswith("weights") else (np.ones(sh, np.float32) if k.endswith("gamma") else np.zeros(sh, np.float32))
              for k, sh in shapes.items()}
    x = rng.standard_normal((batch, a.image, a.image, 3)).astype(np.float32)
    y = rng.integers(0, a.classes, batch).astype(np.int32)
    R.forward_backward(params, x, y, emulate=False)
    t0 = time.perf_counter()
    for _ in range(steps):
        R.forward_backward(params, x, y, emulate=False)
    dt = time.perf_counter() - t0
    return batch * steps / dt, dt / steps * 1e3, cores


def run_reference_resnet(a) -> dict:
    batch = 8
    steps = max(1, min(a.steps, 3))
    value, ms, cores = cpu_resnet_rate(a, batch, steps)
    return {
        "impl": "reference", "metric": "images/sec sync-PS step (ResNet-50)", "value": value, "unit": "images/s",
        "n_gpus": a.gpus, "steps": steps, "warmup": 1, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": config_dict(a, a.gpus, "fp32"),
        "cpu_baseline": {"value": value, "unit": "images/s", "cores": cores, "kind": "port",
                         "sample": "%d forward+backward passes of %d images (oracle/resnet_ref.py, fp32; the parameter "
                                   "update is not included)" % (steps, batch)},
        "e2e": {"value": value, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }


MATH_IMPL = {"tf32x3": "3xTF32 split products on tcgen05 (a_lo*b + a*b_lo + a*b; chunked fp32 accumulation, TMEM + registers): "
                       "fp32-equivalent, parity-tested at the fp32 bounds",
             "fp32": "CUDA-core FFMA", "bf16": "tcgen05 kind::f16, bf16 operands, fp32 accumulate",
             "tf32": "tcgen05 kind::tf32, one MMA per product"}


def config_dict(a, n: int, math: str, model: str = None, batch: int = None) -> dict:
    model = model or a.model
    batch = batch or a.batch
    if model == "resnet50":
        return {
            "workload": "tf_cnn_benchmarks-style ResNet-50 v1.5, synthetic ImageNet %dx%d, %d classes, bf16, "
                        "%d replica(s) sync-PS on %d B200 (BASELINE config 4)" % (a.image, a.image, a.classes, n, n),
            "batch_per_replica": batch, "global_batch": batch * n, "replicas": n, "ps_shards": 0 if n == 1 else n,
            "optimizer": a.opt, "lr": LR, "math": "bf16", "layout": a.layout, "params": 25557032,
            "parallelism": "dp%d + %d PS shards" % (n, n),
            "l2": "flushed between timed steps (256 MiB write, untimed)", "exchange": a.exchange,
        }
    return {
        "workload": ("tf_sample MNIST CNN sync-PS step: %s" %
                     ("examples/tf_job_gpu.yaml shape, 1 MASTER on 1 B200, no PS exchange" if n == 1 else
                      "1 MASTER + %d WORKER + %d PS shards on %d B200 (BASELINE config 3 shape)" % (n - 1, n, n))),
        "batch_per_replica": batch, "global_batch": batch * n, "replicas": n, "ps_shards": 0 if n == 1 else n,
        # tf32x3 computes the fp32 model to fp32 accuracy (parity-tested at the fp32 bounds): the config is the
        # reference arm's; HOW the contractions run is the line's "math_impl" key
        "optimizer": a.opt, "lr": LR, "math": "fp32" if math == "tf32x3" else math, "layout": a.layout,
        "params": 3274634, "parallelism": "dp%d + %d PS shards" % (n, n),
        "l2": "flushed between timed steps (256 MiB write, untimed)", "exchange": a.exchange,
    }


# ---- algorithmic work per launch of every kernel in the step (DESIGN.md "Kernels") ----------
X3_PEAK_NOTE = ("3xTF32 ceiling = measured bf16 peak / 6: kind::tf32 runs at half the bf16 rate and every product "
                "takes three MMAs; achieved = fp32-equivalent (algorithmic) FLOPs / time")


def algorithmic_work(name: str, B: int, n_params: int, opt: str, n_replicas: int, math: str = "bf16",
                     peaks: dict | None = None):
    """-> (bound, amount per launch, peak note): FLOPs ("tensor") or bytes ("hbm").  B = examples per
    replica.  A contraction is charged to whichever of its two rooflines takes longer at the measured
    peaks (operands read once + output written once against HBM, 2 M N K against the tensor pipe): the
    thin-K GEMMs of this model (conv2 dgrad: K = 64, an 800-wide output) are HBM-bound.  In tf32x3 math
    the FLOPs are the fp32-equivalent 2 M N K and the tensor ceiling is the bf16 peak / 6 (peak note)."""
    M2 = B * 196
    gemm = {"gemm conv2 fwd": (M2, 64, 800), "gemm conv2 dgrad": (M2, 800, 64), "gemm conv2 wgrad": (800, 64, M2),
            "gemm fc1 fwd": (B, 1024, 3136), "gemm fc1 dgrad": (B, 3136, 1024), "gemm fc1 wgrad": (3136, 1024, B),
            "gemm fc2 fwd": (B, 10, 1024), "gemm fc2 dgrad": (B, 1024, 10), "gemm fc2 wgrad": (1024, 10, B)}
    # the implicit 3xTF32 conv2 kernels: same contraction, operands = the activations themselves (+ low-order planes)
    conv_x3 = {"conv2 fwd x3": (M2 * 32 * 2 + 800 * 64 * 2, M2 * 64), "conv2 dgrad x3": (M2 * 64 * 2 + 800 * 64 * 2 + M2 * 32, M2 * 32),
               "conv2 wgrad x3": (M2 * 32 * 2 + M2 * 64 * 2, 42 * 800 * 64)}
    if name in conv_x3:
        flops = 2.0 * M2 * 64 * 800
        rd, wr = conv_x3[name]
        nbytes = 4.0 * (rd + wr)
        if peaks and nbytes / (peaks["hbm_gbs"] * 1e9) > flops / (peaks["bf16_tflops_sustained"] / 6 * 1e12):
            return "hbm", nbytes, None
        return "tensor", flops, X3_PEAK_NOTE
    name = name[3:] if name.startswith("tc ") else name     # the tensor-path GEMMs carry a "tc " prefix
    if name in gemm:
        m, n, k = gemm[name]
        flops = 2.0 * m * n * k
        esz = 2 if math == "bf16" else 8 if math == "tf32x3" else 4     # tf32x3 reads each operand and its low-order plane
        out = 2 if (math == "bf16" and name == "gemm conv2 dgrad") else 4   # dcol is kept in bf16, the rest fp32
        nbytes = float((m * k + n * k) * esz + m * n * out)
        tensor_peak = peaks["bf16_tflops_sustained"] / (6.0 if math == "tf32x3" else 1.0) if peaks else None
        if peaks and nbytes / (peaks["hbm_gbs"] * 1e9) > flops / (tensor_peak * 1e12):
            return "hbm", nbytes, None
        return "tensor", flops, (X3_PEAK_NOTE if math == "tf32x3" else None)
    f = 4
    hbm = {
        "conv1_pool_relu_kernel": B * 784 * f + B * 6272 * (f + 1),
        "conv1_bwd_kernel": B * 784 * f + B * 6272 * (f + 1) + B * 832 * f,
        "im2col_kernel": B * 6272 * f + M2 * 800 * f,
        "pool_relu_kernel": M2 * 64 * f + B * 3136 * (f + 1),
        "unpool_kernel": B * 3136 * (2 * f + 1) + M2 * 64 * f,
        "relu_mask_kernel": B * 3136 * 3 * f,
        "col2im_mask_kernel": M2 * 800 * f + B * 6272 * 2 * f,
        "softmax_xent_kernel": B * (10 * 2 * f + 2 * f),
        "push_apply_pull_kernel": n_params * f * (n_replicas + 2 + (2 if opt != "sgd" else 0)),
        # three launches per step (conv2 wgrad 21 splits of 800x64, fc1 fwd 7 of Bx1024, fc1 dgrad 3 of Bx3136):
        # partials read once, result written once; the average launch
        "split-K reduce": f * ((21 + 1) * 800 * 64 + (7 + 1) * B * 1024 + (3 + 1) * B * 3136) / 3.0,
    }
    if name in hbm:
        return "hbm", float(hbm[name]), None
    return "hbm", None, None


def measured_traffic(model: str, kernel: str):
    """DRAM bytes per launch of `kernel` from the ncu captures (profiles/r2_traffic.json, else r1's), at the
    default batch sizes; None when that kernel was not captured."""
    for name in ("r2_traffic.json", "r1_traffic.json"):
        try:
            with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", name)) as f:
                v = json.load(f).get("%s:%s" % (model, kernel))
            if v is not None:
                return v
        except (OSError, ValueError):
            pass
    return None


def resnet_work(name: str, B: int, image: int, classes: int):
    """Algorithmic work per STEP of all ResNet-50 launches that share a profile name: FLOPs of the
    convolutions (explicit-operand and implicit-GEMM launches counted apart, as the profile names
    them), bytes of the batch-norm kernels (bf16 activations read / written once, the relu mask one
    bit per element; DESIGN.md "ResNet-50 kernels")."""
    from oracle import resnet_ref as R   # the layer table only; nothing is executed
    hw = {"image": image, "pool1": image // 4}
    cin = {"image": 3, "pool1": 64}
    fl = {k: 0.0 for k in ("rn conv fwd", "rn conv wgrad", "rn conv dgrad", "rn conv fwd (implicit)",
                           "rn conv wgrad (implicit)", "rn conv dgrad (implicit)")}
    by = {k: 0.0 for k in ("rn bn stats", "rn bn apply", "rn bn bwd stats", "rn bn bwd apply")}
    for name_, inp, rk, stride, kout, relu, res in R.units():
        oh = (hw[inp] + 2 * (rk // 2) - rk) // stride + 1
        hw[name_ + "/y"], cin[name_ + "/y"] = oh, kout
        f = 2.0 * B * oh * oh * kout * rk * rk * cin[inp]
        implicit = not (rk == 1 and stride == 1) and cin[inp] % 64 == 0
        tag = " (implicit)" if implicit else ""
        fl["rn conv fwd" + tag] += f
        fl["rn conv wgrad" + tag] += f
        if inp != "image":   # stride-2 data gradients stay on the explicit path (dz x W^T, then col2im)
            fl["rn conv dgrad" + (tag if stride == 1 else "")] += f
        act = 2.0 * B * oh * oh * kout            # one bf16 activation tensor of this unit
        bits = act / 16 if relu else 0.0
        by["rn bn stats"] += act
        by["rn bn apply"] += act * (3 if res else 2) + bits
        by["rn bn bwd stats"] += act * 2 + bits
        by["rn bn bwd apply"] += act * (4 if res else 3) + bits
    if name in fl:
        return "tensor", fl[name]
    if name in by:
        return "hbm", by[name]
    return "hbm", None


def ps_update_probe(lib, L, torch, stream, n_params: int, opt: str, peaks: dict) -> dict:
    """PS-update GB/s of BASELINE config 5's dense shard: one fused apply launch over
    n_params fp32 parameters resident in HBM (>> L2), CUDA events on the launching stream."""
    kind = L.OPT_KINDS[opt]
    with torch.cuda.stream(stream):
        v = torch.randn(n_params, device="cuda")
        g = torch.randn(n_params, device="cuda")
        m = torch.zeros(n_params, device="cuda")
        ga = L.ptr_array([g])
        times = []
        for i in range(8):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            lib.check(lib.dense_apply(kind, v.data_ptr(), m.data_ptr(), ga, 1, 1.0, n_params, LR, 0.9, None, 0,
                                      stream.cuda_stream))
            e1.record(stream)
            stream.synchronize()
            if i >= 3:
                times.append(e0.elapsed_time(e1))
    bpp = 12 if opt == "sgd" else 20
    ms = float(np.mean(times))
    gbs = bpp * n_params / (ms * 1e-3) / 1e9
    return {"kernel": "push_apply_pull_kernel", "params": n_params, "optimizer": opt, "bytes_per_param": bpp,
            "ms": ms, "achieved": gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": gbs / peaks["hbm_gbs"],
            "peak_source": peaks["source"]}


def parity_check(a, n, rank, local, math, torch, dist, stream) -> dict:
    """N > 1, before any timing: three synchronous steps of a small job (16 examples per replica) over
    the real peers.  Step-1 mean loss and the accumulated update after step 3 against the oracle's
    sync_step on the same batches (rank 0; the oracle is the checker here, nothing of it is timed),
    and whether every rank ends with bit-identical parameters (crc32 of the flat buffer, all-gathered)."""
    import zlib
    from k8s_b200.syncps import SyncPSJob, mnist_init, synthetic_batch
    pb = 16
    job = SyncPSJob(tf_config_for_rank(n, rank), rank, n, cuda_device=local, stream=stream.cuda_stream)
    job.build_mnist(pb, math=math, layout=a.layout)
    job.optimizer(a.opt, LR, 0.9)
    job.connect_peers(dist)
    losses = []
    for step in range(3):
        xs, ys = zip(*[synthetic_batch(r, pb, step) for r in job.local_replicas])
        losses.append(job.step(list(xs), list(ys)))
    mine = np.concatenate([job.get_var(v) for v in range(8)])
    crcs = [None] * n
    dist.all_gather_object(crcs, zlib.crc32(mine.tobytes()))
    out = {"layout": a.layout, "math": math, "batch_per_replica": pb, "steps": 3,
           "ranks_bit_identical": all(c == crcs[0] for c in crcs), "loss_step1": losses[0]}
    if rank == 0:
        from oracle import model_ref as M
        params = mnist_init(0)
        init = np.concatenate([p.ravel() for p in params])
        accum = [np.zeros_like(p) for p in params]
        ref_losses = []
        for step in range(3):
            batches = [synthetic_batch(r, pb, step) for r in range(n)]
            ref_losses.append(M.sync_step(params, accum, batches, a.opt, LR, 0.9, math=math)[0])
        ref = np.concatenate([p.ravel() for p in params])
        upd = float(np.linalg.norm((mine - init) - (ref - init)) / np.linalg.norm(ref - init))
        fp32_like = math in ("fp32", "tf32x3")
        out.update(oracle_loss_step1=ref_losses[0], loss_rel_err_step1=abs(losses[0] - ref_losses[0]) / abs(ref_losses[0]),
                   update_rel_l2_after_3_steps=upd, update_tolerance=3e-3 if fp32_like else 0.15,
                   loss_tolerance=1e-5 if fp32_like else 5e-3)
        out["ok"] = bool(out["ranks_bit_identical"] and upd <= out["update_tolerance"] and
                         out["loss_rel_err_step1"] <= out["loss_tolerance"])
    job.close()
    dist.barrier()
    return out


def measure(a, model: str, n: int, rank: int, local: int, lib, L, torch, dist, stream, peaks, steps: int) -> dict:
    """Build `model`'s job on this rank, time `steps` device-resident steps and `steps` end-to-end steps
    (max over ranks), then the per-kernel profile; returns the JSON line's body for that model."""
    from k8s_b200.syncps import SyncPSJob, synthetic_batch, synthetic_images
    resnet = model == "resnet50"
    B = a.batch if model == a.model else (256 if resnet else BATCH)
    math = "bf16" if resnet else ("tf32x3" if a.math == "auto" else a.math)
    job = SyncPSJob(tf_config_for_rank(n, rank), rank, n, cuda_device=local, stream=stream.cuda_stream)
    if resnet:
        job.build_resnet50(B, a.image, a.classes, layout=a.layout)
    else:
        job.build_mnist(B, math=math, layout=a.layout)
    job.optimizer(a.opt, LR, 0.9)
    job.connect_peers(dist if n > 1 else None)
    slots = job.local_replicas

    # synthetic inputs: resident in HBM for `value`, in pinned host memory for `e2e`
    host, dev = [], []
    for s in range(N_INPUT_SETS):
        hx, hy, dx, dy = [], [], [], []
        for r in slots:
            x, y = synthetic_images(r, B, a.image, a.classes, s) if resnet else synthetic_batch(r, B, s)
            px, py = torch.from_numpy(x).pin_memory(), torch.from_numpy(y).pin_memory()
            hx.append(px); hy.append(py)
            dx.append(px.cuda()); dy.append(py.cuda())
        host.append((hx, hy)); dev.append((dx, dy))
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()

    def barrier():
        if n > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, k_steps):
        """K steps, each bracketed by events on the job's stream, L2 flushed (untimed) in
        between; returns the summed device time in ms, max over ranks."""
        evs = []
        barrier()
        for k in range(k_steps):
            with torch.cuda.stream(stream):
                flush.fill_(k & 0xFF)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            fn(k)
            e1.record(stream)
            evs.append((e0, e1))
        barrier()
        per = [e0.elapsed_time(e1) for e0, e1 in evs]
        ms = sum(per)
        if os.environ.get("TFGRAFT_BENCH_DEBUG"):
            print("rank %d per-step ms: %s" % (rank, " ".join("%.3f" % v for v in per)), file=sys.stderr, flush=True)
        if n > 1:
            t = torch.tensor([ms], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    def step_dev(k):
        dx, dy = dev[k % N_INPUT_SETS]
        if a.exchange == "nccl":
            with torch.cuda.stream(stream):     # NCCL work must land on the job's stream
                job.step_nccl(dx, dy, dist)
        else:
            job.step_dev(dx, dy)

    losses = []

    def step_e2e(k):
        hx, hy = host[k % N_INPUT_SETS]
        losses.append(job.step(hx, hy))

    # Everything with a one-time cost comes BEFORE the warm-up: the NVML sampler, and the first
    # dist.barrier() (lazy NCCL set-up takes ~100 ms during which the GPUs idle and drop their
    # clocks; measured at N = 4: the first five "timed" steps took 2.2 / 0.7 / 1.3 / 0.4 / 0.5 ms
    # against 0.26 ms steady state when that barrier sat between warm-up and timing).
    sampler = ClockSampler(local)
    if not os.environ.get("TFGRAFT_BENCH_NO_CLOCKS"):
        sampler.start()
    barrier()
    # The step graph is keyed by its input pointers (eager on first sight, captured on second,
    # replayed after).  Capturing is set-up work (tens of ms of host time for the 560-launch ResNet
    # step, which at N > 1 the other ranks would spend inside their barrier kernel): it is done
    # here, before the W warm-up steps, so warm-up and timed steps are all graph replays.
    for k in range(2 * N_INPUT_SETS):
        step_dev(k)
    job.sync()
    warm = max(a.warmup, 3)
    for k in range(warm):
        step_dev(k)
    job.sync()
    launches0 = job.launches
    ms_dev = timed(step_dev, steps)
    launches_timed = job.launches - launches0
    if a.exchange == "nccl":
        ms_e2e = float("nan")      # the baseline arm is timed device-resident only
    else:
        for k in range(2 * N_INPUT_SETS + 3):      # the host-buffer path stages into its own buffers: its own graph
            step_e2e(k)
        ms_e2e = timed(step_e2e, steps)
    clocks = sampler.stop()

    # per-kernel device time over another K steps (events behind every launch)
    job.profile(True)
    for k in range(steps):
        step_dev(k)
    prof = job.profile_report()
    job.profile(False)
    examples = steps * B * n
    unit = "images/s" if resnet else "examples/s"
    out = {
        "metric": "images/sec sync-PS step (ResNet-50)" if resnet else "examples/sec sync-PS step",
        "value": examples / (ms_dev * 1e-3), "unit": unit, "n_gpus": n,
        "steps": steps, "warmup": warm, "ms_per_step": ms_dev / steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        # tf32x3 = fp32 operands, three tf32 MMAs per product, fp32 accumulate: fp32-equivalent results (held
        # to the fp32 parity bounds in tests/test_gpu_model.py), hence "f32" like the reference arm
        "dtype": {"fp32": "f32", "tf32x3": "f32", "bf16": "bf16", "tf32": "tf32"}[math], "data": "synthetic",
        "config": config_dict(a, n, math, model, B),
        "e2e": {"value": examples / (ms_e2e * 1e-3), "unit": unit, "ms_per_step": ms_e2e / steps,
                "h2d_bytes_per_step": n * B * ((a.image * a.image * 3 if resnet else 784) * 4 + 4),
                "d2h_bytes_per_step": 4 * n, "last_loss": losses[-1] if losses else None},
        "gpu_launches": int(launches_timed),
        "math_impl": MATH_IMPL[math],
        "clocks": clocks,
    }
    if rank == 0:
        total = sum(v["ms"] for v in prof.values()) or 1.0
        # the cross-rank barrier is waiting time (the slowest rank's skew), not work: it is reported as its
        # share of the step, and the roofline kernel is the largest one that moves bytes or does math
        # ... and so is, at N > 1, the exchange kernel's time in this per-kernel pass (it waits inside the kernel
        # for the slowest rank; its bandwidth against the NVLink roofline is scripts/ps_exchange_sweep.py's job)
        waits = ("barrier_kernel",) + (("push_apply_pull_kernel",) if n > 1 else ())
        work = {k_: v for k_, v in prof.items() if k_ not in waits} or prof
        top = max(work, key=lambda k_: work[k_]["ms"])
        bound, amount, peak_note = algorithmic_work(top, B, 3274634, a.opt, n, math, peaks)
        per_launch_ms = prof[top]["ms"] / prof[top]["count"]
        if resnet:   # per-layer launches share a name: work = the sum over the 53 convolutions, per step
            bound, amount = resnet_work(top, B, a.image, a.classes)
            peak_note = None
            per_launch_ms = prof[top]["ms"] / steps
        roof = {"kernel": top, "bound": bound, "share_of_step": prof[top]["ms"] / total,
                "ms_per_launch": per_launch_ms, "traffic": (measured_traffic(model, top)
                            if n == 1 and B == (256 if resnet else 128) else None),   # captured at these sizes only
                "peak_source": peaks["source"]}
        if amount is not None and bound == "tensor":
            ach = amount / (per_launch_ms * 1e-3) / 1e12
            peak = peaks["bf16_tflops_sustained"] / (6.0 if peak_note else 1.0)
            roof.update(achieved=ach, peak=peak, unit="TFLOP/s", frac=ach / peak, algorithmic_flops=amount)
            if peak_note:
                roof["peak_source"] = peaks["source"] + "; " + peak_note
        elif amount is not None:
            ach = amount / (per_launch_ms * 1e-3) / 1e9
            roof.update(achieved=ach, peak=peaks["hbm_gbs"], unit="GB/s", frac=ach / peaks["hbm_gbs"],
                        algorithmic_bytes=amount)
        if "barrier_kernel" in prof:
            roof["barrier_share_of_step"] = prof["barrier_kernel"]["ms"] / total
        if n > 1 and "push_apply_pull_kernel" in prof:
            # bytes one GPU receives (and, mirrored, sends) per step over NVLink: its shard's gradient from the N-1
            # other replicas and the new values (bf16 math: and their shadow) of the N-1 other shards
            n_params = 25557032 if resnet else 3274634
            # (3xTF32: fp32 values only; every rank derives the low-order plane itself at the next forward)
            shadow = 2 if math == "bf16" else 0
            roof["exchange"] = {"share_of_step_incl_wait": prof["push_apply_pull_kernel"]["ms"] / total,
                                "nvlink_bytes_per_gpu_per_direction": (8 + shadow) * n_params * (n - 1) / n,
                                "note": "the kernel waits in-kernel for the slowest rank; see scripts/ps_exchange_sweep.py "
                                        "for its bandwidth against the measured link"}
        out["roofline"] = roof
        out["kernels"] = {k_: {"count": v["count"], "ms_per_launch": v["ms"] / v["count"],
                               "share": v["ms"] / total} for k_, v in sorted(prof.items(), key=lambda kv: -kv[1]["ms"])}
    job.close()
    del dev, flush, host
    torch.cuda.empty_cache()
    return out


def run_graft(a) -> dict:
    import torch
    import torch.distributed as dist
    from k8s_b200 import lib as L

    n = a.gpus
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != n:
        raise SystemExit("--gpus %d but WORLD_SIZE=%d: launch with torchrun --nproc-per-node %d" % (n, world, n))
    lib = L.load()
    if lib.device_count_checked() <= local:
        raise SystemExit("no CUDA device %d: the CUDA path has no CPU fallback" % local)
    torch.cuda.set_device(local)
    if n > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    stream = torch.cuda.Stream()
    peaks = measured_peaks()

    parity = None
    if n > 1 and a.model == "mnist" and a.exchange == "peer":
        parity = parity_check(a, n, rank, local, "tf32x3" if a.math == "auto" else a.math, torch, dist, stream)
    out = measure(a, a.model, n, rank, local, lib, L, torch, dist, stream, peaks, a.steps)
    if parity is not None:
        out["parity"] = parity
    # BASELINE config 4 rides along on the default line so that the driver's clock covers it too
    if a.model == "mnist" and a.with_resnet and a.exchange == "peer":
        r = measure(a, "resnet50", n, rank, local, lib, L, torch, dist, stream, peaks, min(a.steps, 10))
        out["resnet50"] = {k: r[k] for k in ("metric", "value", "unit", "ms_per_step", "steps", "warmup", "dtype", "config",
                                             "e2e", "gpu_launches", "clocks", "roofline") if k in r}
    if rank == 0:
        out["ps_update"] = ps_update_probe(lib, L, torch, stream, a.ps_params, a.opt, peaks)
        if n == 1 and not a.no_cpu_baseline:
            if a.model == "resnet50":
                v, ms, cores = cpu_resnet_rate(a, 8, 2)
                out["cpu_baseline"] = {"value": v, "unit": "images/s", "cores": cores, "kind": "port",
                                       "sample": "2 forward+backward passes of 8 images (oracle/resnet_ref.py, fp32)"}
            else:
                v, ms, cores = cpu_step_rate(1, a.batch, a.opt, 250, 3)
                out["cpu_baseline"] = {"value": v, "unit": "examples/s", "cores": cores, "kind": "port",
                                       "sample": "250 full steps of 1 replica x %d examples (oracle port, "
                                                 "torch-CPU fwd/bwd + C apply)" % a.batch}
    if n > 1:
        dist.barrier()
        dist.destroy_process_group()
    return out


def main(argv=None) -> int:
    a = parse_args(argv)
    rank = int(os.environ.get("RANK", "0"))
    if a.impl == "reference":
        if rank == 0:
            print(json.dumps(run_reference(a)), flush=True)
        return 0
    out = run_graft(a)
    if rank == 0:
        print(json.dumps(out), flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
