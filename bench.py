#!/usr/bin/env python
"""Headline benchmark = BASELINE.json's metric: clips/sec of the whole hot path, raw 1 s 16 kHz PCM ->
40-MFCC (40 mel filters) -> reference DS-CNN (64 channels, 4 blocks) -> logits, 65 536 clips per GPU
per step (`config.workload = "mfcc40_dscnn_s_b65536"`; reference call chain
kws/libs/data_loader.py:102-104 -> kws/libs/training.py:292 -> kws/libs/models.py:160-183).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--clips B]

One process per GPU (torchrun for N>1), batch sharded by rank with no collective on the data path
(clips are independent): weak scaling, every rank runs its own B clips per step.  Prints ONE JSON line
on rank 0.  See DESIGN.md section 6 for every field.  `extra` carries the other BASELINE configs
(MFCC only, batch-256 latency case, DS-CNN-L, log-mel + CRNN, training step) with their own rooflines.
"""
import os

# The CPU legs run the reference's feature path one clip per call in a fork pool of all host cores
# (train.py:108-114 DataLoader workers).  Each worker must be single-threaded: with BLAS/OpenMP threading
# left on, every one of the C workers spawns C threads and the arm loses ~10x.  The
# libraries read these at import time, so they are set before numpy/scipy/torch are imported.
for _v in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS", "NUMEXPR_NUM_THREADS",
           "VECLIB_MAXIMUM_THREADS"):
    os.environ[_v] = "1"

import argparse  # noqa: E402
import json  # noqa: E402
import subprocess  # noqa: E402
import sys  # noqa: E402
import threading  # noqa: E402
import time  # noqa: E402

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_SAMPLES = 16000
NUMCEP = NFILT = 40
NFRAMES = 99
NCLS = 12
METRIC = "clips/sec (1 s 16 kHz, PCM→logits)"
WORKLOAD = "mfcc40_dscnn_s_b65536"

# Algorithmic work per clip (stated in DESIGN.md section 6).
#   bytes: what must cross HBM; f_tc: GEMM-able flops (tensor core); f_fp32: CUDA-core flops.
WORK = {
    "pcm_to_logits": {"bytes": N_SAMPLES * 4 + NCLS * 4, "f_tc": 45.40e6 + 0.32e6, "f_fp32": 4.95e6 + 1.35e6},
    "mfcc40": {"bytes": N_SAMPLES * 4 + NFRAMES * NUMCEP * 4, "f_tc": 0.32e6, "f_fp32": 1.35e6},
    "dscnn_s_model": {"bytes": NFRAMES * NUMCEP * 4 + NCLS * 4, "f_tc": 45.40e6, "f_fp32": 4.95e6},
    "dscnn_l_model": {"bytes": NFRAMES * NUMCEP * 4 + NCLS * 4, "f_tc": 907.5e6, "f_fp32": 28.5e6},
    # conv1 (16 ch) 2.71 + GRU input projection 5.20 + recurrence 1.16 MFLOP; log-mel front end + gates on CUDA cores
    "logmel_crnn": {"bytes": N_SAMPLES * 4 + NCLS * 4, "f_tc": 9.07e6, "f_fp32": 1.03e6 + 0.12e6},
    # training step at 99x10 features: 3x the forward (fwd + dgrad + wgrad), SURVEY 8(d) config 5
    "train_f10": {"bytes": NFRAMES * 10 * 4 + 8, "f_tc": 3 * 11.80e6, "f_fp32": 3 * 1.44e6},
}


def _peaks():
    """hbm GB/s, bf16 burst / sustained TFLOP/s: MEASURED_PEAKS.json (driver-written) or the profiling
    guide's fallback."""
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        j = json.load(open(p))
        return {"hbm_gbs": float(j["hbm_gbs"]), "bf16_tflops": float(j["bf16_tflops"]),
                "bf16_tflops_sustained": float(j.get("bf16_tflops_sustained", j["bf16_tflops"])),
                "source": "measured (MEASURED_PEAKS.json)"}
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1650.0, "bf16_tflops_sustained": 1400.0,
                "source": "fallback (B200_PROFILING.md)"}


def roofline3(work, clips_per_s_per_gpu, peaks, fp32_tflops, kernel, kernel_ms=None, traffic=None, sustained=True):
    """T_roof = max(B_alg/BW_hbm, F_tc/P_tc, F_fp32/P_fp32) per clip; frac = T_roof/T_measured.
    The binding bound is named, and the plain HBM and tensor fractions are always printed beside it.
    P_tc is the driver-measured cuBLAS bf16 rate (sustained for kernels timed inside a long step), the
    figure SURVEY/BASELINE.md quote the tensor bound on; the kernels compute in tf32 (half that rate)."""
    p_tc = peaks["bf16_tflops_sustained" if sustained else "bf16_tflops"]
    t = {"hbm": work["bytes"] / (peaks["hbm_gbs"] * 1e9), "tensor": work["f_tc"] / (p_tc * 1e12),
         "fp32": work["f_fp32"] / (fp32_tflops * 1e12)}
    bound = max(t, key=t.get)
    t_meas = 1.0 / clips_per_s_per_gpu
    if bound == "hbm":
        achieved, peak, unit = work["bytes"] * clips_per_s_per_gpu / 1e9, peaks["hbm_gbs"], "GB/s"
    elif bound == "tensor":
        achieved, peak, unit = work["f_tc"] * clips_per_s_per_gpu / 1e12, p_tc, "TFLOP/s"
    else:
        achieved, peak, unit = work["f_fp32"] * clips_per_s_per_gpu / 1e12, fp32_tflops, "TFLOP/s"
    return {"bound": bound, "achieved": achieved, "peak": peak, "unit": unit, "frac": t[bound] / t_meas,
            "traffic": traffic, "kernel": kernel, "kernel_ms": kernel_ms,
            "hbm_frac": t["hbm"] / t_meas, "tensor_frac": t["tensor"] / t_meas, "fp32_frac": t["fp32"] / t_meas,
            "tensor_frac_of_tf32_peak": 2.0 * t["tensor"] / t_meas,
            "roofline_clips_per_s": 1.0 / t[bound],
            "per_clip": {"bytes": work["bytes"], "f_tc": work["f_tc"], "f_fp32": work["f_fp32"]},
            "peaks": {"hbm_gbs": peaks["hbm_gbs"], "tensor_tflops": p_tc, "fp32_tflops": fp32_tflops,
                      "source": peaks["source"] + "; fp32 = kws_probe_fp32_tflops measured in this run"}}


def _config(world, clips_per_gpu):
    """Identical in both arms (the reference arm runs a bounded sample of the same workload)."""
    return {"workload": WORKLOAD, "clips_per_gpu": clips_per_gpu, "samples": N_SAMPLES, "numcep": NUMCEP,
            "nfilt": NFILT, "nfft": 512, "model": "reference DepthwiseSeparableConv (64 ch, 4 blocks, 12 classes)",
            "parallelism": f"batch-sharded x{world}",
            "l2": "inputs (4.2 GB/step/GPU) larger than L2, no flush needed"}


# ---------------------------------------------------------------------------------------
# CPU arm: the reference's own CPU path for PCM -> logits.
#   features: psf 0.6 restated in oracle/psf_mfcc.py, one clip per call like the reference's DataLoader
#             workers (train.py:108-114), fork pool of all host cores, every worker single-threaded;
#   model:    the reference network forward on CPU (oracle/dscnn_ref.forward = the same ATen conv2d/linear
#             calls as kws/libs/models.py:160-183), torch.no_grad(), all host cores as intra-op threads,
#             batch 256 (BASELINE configs[0]);
#   end to end = the two back to back on the same cores: clips / (t_features + t_model).
# ---------------------------------------------------------------------------------------
def _cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


_WORKER_PCM = None


def _cpu_worker(args):
    """n clips through the reference's per-clip feature call; the synthetic PCM of a worker is generated once
    (outside the timed calls: the reference's workers read wav files, they do not draw random numbers)."""
    global _WORKER_PCM
    seed, n = args
    import numpy as np
    from oracle import psf_mfcc
    if _WORKER_PCM is None:
        try:
            from threadpoolctl import threadpool_limits
            threadpool_limits(1)
        except Exception:
            pass
        rs = np.random.RandomState(os.getpid() & 0xffff)
        _WORKER_PCM = (rs.rand(64, N_SAMPLES).astype(np.float32) * 2.0 - 1.0)
    acc = 0.0
    for i in range(n):
        acc += float(psf_mfcc.extract_features(_WORKER_PCM[(seed + i) % 64], numcep=NUMCEP, nfilt=NFILT)[0, 0])
    return acc


class CpuReference:
    """Fork pool first (before torch spins up intra-op threads), then the model on all cores."""

    def __init__(self, cores):
        import multiprocessing as mp
        self.cores = cores
        self.pool = mp.get_context("fork").Pool(cores)
        self.pool.map(_cpu_worker, [(i, 1) for i in range(cores)])        # imports + warm-up
        import torch
        from oracle import dscnn_ref
        torch.set_num_threads(cores)
        self.torch, self.fwd = torch, dscnn_ref.forward
        self.sd = dscnn_ref.init_state_dict(seed=0)
        g = torch.Generator().manual_seed(0)
        self.x = torch.randn(256, 1, NFRAMES, NUMCEP, generator=g)
        with torch.no_grad():
            self.fwd(self.sd, self.x[:32])

    def features_rate(self, per_worker, seed0=100):
        t0 = time.perf_counter()
        self.pool.map(_cpu_worker, [(seed0 + i, per_worker) for i in range(self.cores)])
        return per_worker * self.cores / (time.perf_counter() - t0)

    def model_rate(self, batches=1):
        t0 = time.perf_counter()
        with self.torch.no_grad():
            for _ in range(batches):
                self.fwd(self.sd, self.x)
        return batches * self.x.shape[0] / (time.perf_counter() - t0)

    def psf_check(self):
        """If the pod ships the real python_speech_features, record the restatement's deviation from it."""
        try:
            import numpy as np
            import python_speech_features as psf
            from oracle import psf_mfcc
            rs = np.random.RandomState(7)
            dev = 0.0
            for _ in range(4):
                x = (rs.rand(N_SAMPLES).astype(np.float32) * 2 - 1)
                a = psf.mfcc(x, 16000, numcep=NUMCEP, nfilt=NFILT, nfft=512, winlen=0.025, winstep=0.01)
                dev = max(dev, float(np.abs(a - psf_mfcc.extract_features(x, numcep=NUMCEP, nfilt=NFILT)).max()))
            return {"present": True, "oracle_max_abs_dev": dev}
        except ImportError:
            return {"present": False}
        except Exception as e:  # pragma: no cover
            return {"present": True, "error": repr(e)}

    def close(self):
        self.pool.close()
        self.pool.join()


def _combine(f, m):
    return 1.0 / (1.0 / f + 1.0 / m)


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = _cores()
    ref = CpuReference(cores)
    per_worker, batches = 16, 1
    for _ in range(max(args.warmup, 1)):
        ref.features_rate(2)
        ref.model_rate(1)
    fr, mr = [], []
    t_all = time.perf_counter()
    for s in range(args.steps):
        fr.append(ref.features_rate(per_worker, seed0=1000 + s * cores))
        mr.append(ref.model_rate(batches))
    total_s = time.perf_counter() - t_all
    f, m = sum(fr) / len(fr), sum(mr) / len(mr)
    value = _combine(f, m)
    psf_pkg = ref.psf_check()
    ref.close()
    import numpy
    import scipy
    import torch
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "clips/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total_s / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64 features / f32 model",
        "data": "synthetic", "config": _config(args.gpus, args.clips),
        "cpu_baseline": {"value": value, "unit": "clips/s", "cores": cores, "kind": "port",
                         "sample": (f"per step: {per_worker * cores} clips through oracle/psf_mfcc.py (psf 0.6 restatement, one clip "
                                    f"per call, {cores} single-threaded fork workers) + {256 * batches} clips through the reference "
                                    f"network on CPU (torch {cores} threads, batch 256); x {args.steps} steps; "
                                    "value = clips / (t_features + t_model)"),
                         "features_clips_per_s": f, "model_clips_per_s": m,
                         "python_speech_features": psf_pkg,
                         "versions": {"numpy": numpy.__version__, "scipy": scipy.__version__, "torch": torch.__version__}},
        "e2e": {"value": value, "unit": "clips/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    _emit(line)


# ---------------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        super().__init__(daemon=True)
        self.gpu, self.rows, self._stop_evt = gpu_index, [], threading.Event()

    def run(self):
        while not self._stop_evt.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                      "-i", str(self.gpu)], capture_output=True, text=True, timeout=5).stdout
                self.rows.append([c.strip() for c in out.strip().split(",")])
            except Exception:
                pass
            self._stop_evt.wait(0.1)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=6)
        sm, mx, reasons = [], 0.0, set()
        for r in self.rows:
            if len(r) < 8:
                continue
            try:
                sm.append(float(r[1])); mx = max(mx, float(r[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None,
                "reasons": sorted(reasons), "samples": len(sm)}


def _bind_to_gpu_numa_node(gpu_index):
    """Multi-GPU host-buffer path: pin this rank to the CPUs NVML lists as local to its GPU before the pinned
    buffers are allocated, so that first-touch places them on the GPU's own NUMA node (the e2e copy otherwise
    crosses the socket link for half of the ranks).  Best effort: returns the CPU count bound to, or None."""
    try:
        import pynvml
        pynvml.nvmlInit()
        try:   # CUDA and NVML may enumerate differently: go through the PCI address
            import torch
            pr = torch.cuda.get_device_properties(gpu_index)
            hnd = pynvml.nvmlDeviceGetHandleByPciBusId(
                ("%08x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)).encode())
        except Exception:
            hnd = pynvml.nvmlDeviceGetHandleByIndex(gpu_index)
        ncpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(hnd, (ncpu + 63) // 64)
        cpus = {64 * wi + b for wi, wd in enumerate(words) for b in range(64) if (wd >> b) & 1}
        cpus &= set(os.sched_getaffinity(0))
        if cpus:
            os.sched_setaffinity(0, cpus)
            return len(cpus)
    except Exception:
        pass
    return None


def _traffic(kernel):
    """dram bytes per clip of `kernel` from the committed ncu --set full capture (profiles/traffic.json)."""
    try:
        j = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        for rec in j.get("kernels", []):
            if rec.get("kernel") == kernel:
                return rec
    except Exception:
        pass
    return None


def run_ours(args):
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    # CPU baseline first (fork before CUDA is initialised), rank 0 at N=1 only: ~10-20 s of CPU work
    cpu_base = None
    if world == 1 and not args.no_cpu_baseline:
        cores = _cores()
        ref = CpuReference(cores)
        per_worker, batches = 48, 8
        ref.features_rate(2); ref.model_rate(1)
        f, m = ref.features_rate(per_worker), ref.model_rate(batches)
        psf_pkg = ref.psf_check()
        ref.close()
        cpu_base = {"value": _combine(f, m), "unit": "clips/s", "cores": cores, "kind": "port",
                    "sample": (f"{per_worker * cores} clips through oracle/psf_mfcc.py (psf 0.6 restatement, one clip per call, "
                               f"{cores} single-threaded fork workers) + {256 * batches} clips through the reference network on "
                               f"CPU (torch {cores} threads, batch 256); value = clips / (t_features + t_model)"),
                    "features_clips_per_s": f, "model_clips_per_s": m, "python_speech_features": psf_pkg}

    import torch
    import torch.distributed as dist
    torch.set_num_threads(max(1, min(8, _cores())))

    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa = _bind_to_gpu_numa_node(local) if world > 1 else None
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    import ctypes as C

    import kws_b200  # noqa: F401
    from kws_b200 import _native
    from kws_b200.audio_processor import AudioConfig, AudioProcessor
    from kws_b200.models import DepthwiseSeparableConv
    from kws_b200.pipeline import KeywordSpotter

    peaks = _peaks()
    a, b = C.c_double(0), C.c_double(0)
    _native.check(_native.lib().kws_probe_fp32_tflops(_native.handle_for(local).ptr, C.byref(a), C.byref(b)))
    fp32_tflops = max(a.value, b.value)

    B = args.clips
    torch.manual_seed(0)
    model = DepthwiseSeparableConv().to(dev)
    with torch.no_grad():
        for prm in model.parameters():
            if prm.dim() == 1:
                prm.normal_(0, 0.1)       # non-zero biases: the padded-pointwise border rule is exercised
    ks = KeywordSpotter(model, AudioConfig(), numcep=NUMCEP, nfilt=NFILT, device=dev)
    proc = AudioProcessor(None, AudioConfig(), device=dev)
    g = torch.Generator(device=dev).manual_seed(1234 + rank)
    pcm = torch.rand((B, N_SAMPLES), generator=g, device=dev, dtype=torch.float32) * 2.0 - 1.0

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def maxr(x):
        t = torch.tensor([x], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def timed(fn, reps, warm=2):
        """ms per call, CUDA events on the launching stream, max over ranks."""
        for _ in range(warm):
            fn()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        barrier()
        return maxr(e0.elapsed_time(e1) / reps)

    # ---- headline: PCM resident in HBM -> logits, K steps ----
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    W = max(args.warmup, 3)
    for _ in range(W):
        out = ks(pcm)
    barrier()
    l0 = model.launch_count()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    barrier()
    ev[0].record()
    for i in range(args.steps):
        out = ks(pcm)
        ev[i + 1].record()
    barrier()
    launches = model.launch_count() - l0
    total_ms = maxr(ev[0].elapsed_time(ev[-1]))
    value = world * B * args.steps / (total_ms * 1e-3)
    step_ms = total_ms / args.steps
    del out

    # ---- the same step as two launches (MFCC kernel, then network kernel, features through HBM), for comparison ----
    two_kernel = None
    try:
        ks.configure_single_launch(min_clips=0)
        ms2 = timed(lambda: ks(pcm), 5)
        two_kernel = {"ms_per_step": ms2, "clips_per_s": world * B / (ms2 * 1e-3),
                      "note": "kws_pcm_to_logits with the single launch switched off: mfcc512_ws_kernel + dscnn_fused_kernel"}
    except Exception as e:
        two_kernel = {"error": repr(e)}
    ks.configure_single_launch(min_clips=1)

    # ---- the stand-alone kernels of the two phases, each alone on the same data (CUDA events) ----
    kern = {}
    try:
        feats = proc.extract_features(pcm, numcep=NUMCEP, nfilt=NFILT).unsqueeze(1)
        ms_mfcc = timed(lambda: proc.extract_features(pcm, numcep=NUMCEP, nfilt=NFILT), 5)
        ms_model = timed(lambda: model(feats), 5)
        del feats
        kern = {"mfcc512_ws_kernel": {"ms": ms_mfcc, "clips_per_s": world * B / (ms_mfcc * 1e-3),
                                      "roofline": roofline3(WORK["mfcc40"], B / (ms_mfcc * 1e-3), peaks, fp32_tflops,
                                                            "mfcc512_ws_kernel", ms_mfcc, _traffic("mfcc512_ws_kernel"))},
                "dscnn_fused_kernel": {"ms": ms_model, "clips_per_s": world * B / (ms_model * 1e-3),
                                       "roofline": roofline3(WORK["dscnn_s_model"], B / (ms_model * 1e-3), peaks, fp32_tflops,
                                                             "dscnn_fused_kernel", ms_model, _traffic("dscnn_fused_kernel"))}}
    except Exception as e:
        kern = {"error": repr(e)}

    # ---- e2e: HOST pinned PCM through the public API, H2D and D2H inside the timed region ----
    Be = min(B, args.e2e_clips)
    host = torch.empty((Be, N_SAMPLES), dtype=torch.float32, pin_memory=True)
    host.copy_(pcm[:Be])
    host_logits = torch.empty((Be, NCLS), dtype=torch.float32, pin_memory=True)

    def timed_host(x, reps):
        for _ in range(2):
            ks.infer_host(x, out=host_logits)
        barrier()
        t0 = time.perf_counter()
        for _ in range(reps):
            ks.infer_host(x, out=host_logits)       # returns after the last D2H has completed
        torch.cuda.synchronize()
        return world * Be * reps / maxr(time.perf_counter() - t0)

    e2e_value = timed_host(host, args.steps)
    host16 = torch.empty((Be, N_SAMPLES), dtype=torch.int16, pin_memory=True)
    host16.copy_((pcm[:Be] * 32767.0).round().to(torch.int16))
    e2e_i16 = timed_host(host16, args.steps)
    del host, host16

    extra = {}
    if not args.no_extra:
        extra = _extras(args, torch, dist, dev, rank, world, pcm, model, ks, proc, timed, barrier, maxr, peaks, fp32_tflops)
    clocks = sampler.stop() if sampler else None

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": "clips/s", "n_gpus": world, "steps": args.steps,
            "warmup": W, "ms_per_step": step_ms, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "tf32 GEMMs (fp32 accumulate) + fp32", "data": "synthetic",
            "config": _config(world, B),
            "roofline": roofline3(WORK["pcm_to_logits"], B / (step_ms * 1e-3), peaks, fp32_tflops,
                                  "pcm_logits_fused_kernel (one launch per step: MFCC phase + network phase per CTA and chunk)",
                                  step_ms, _traffic("pcm_logits_fused_kernel")),
            "two_kernel_path": two_kernel,
            "kernels": kern,
            "cpu_baseline": cpu_base,
            "e2e": {"value": e2e_value, "unit": "clips/s", "h2d_bytes_per_step": Be * N_SAMPLES * 4,
                    "d2h_bytes_per_step": Be * NCLS * 4, "clips_per_step": Be, "input": "float32 pinned host PCM (librosa.load's dtype)",
                    "int16_value": e2e_i16, "int16_h2d_bytes_per_step": Be * N_SAMPLES * 2,
                    "int16_note": "same call fed 16-bit PCM, the dataset's wav sample format (audio_processor.py:145)"},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "numa_bound_cpus": numa,
            "extra": extra,
        }
        _emit(line)
    if world > 1:
        dist.destroy_process_group()


def _extras(args, torch, dist, dev, rank, world, pcm, model, ks, proc, timed, barrier, maxr, peaks, fp32_tflops):
    """The other BASELINE configs, each a small timed leg with its own roofline block.  Failures are recorded,
    never fatal: the headline line must print."""
    from kws_b200.audio_processor import AudioConfig
    from kws_b200.models import CRNN, DepthwiseSeparableConv
    from kws_b200.pipeline import KeywordSpotter
    from kws_b200.training import KWSTrainStep, TrainingConfig
    B = pcm.shape[0]
    ex = {}

    def leg(name, fn):
        try:
            ex[name] = fn()
        except Exception as e:
            ex[name] = {"error": repr(e)}

    def cfg2():   # BASELINE configs[1]: MFCC only, 65 536 clips
        ms = timed(lambda: proc.extract_features(pcm, numcep=NUMCEP, nfilt=NFILT), 5)
        cps = B / (ms * 1e-3)
        return {"workload": "mfcc40_b%d" % B, "clips_per_s": world * cps, "ms": ms,
                "roofline": roofline3(WORK["mfcc40"], cps, peaks, fp32_tflops, "mfcc512_ws_kernel", ms)}

    def cfg1():   # BASELINE configs[0]: batch 256 (1.7 clips per SM: one wave, latency-bound by construction)
        x = pcm[:256]
        ms = timed(lambda: ks(x), 50, warm=5)
        return {"workload": "mfcc40_dscnn_s_b256", "ms_per_batch": ms, "clips_per_s": world * 256 / (ms * 1e-3),
                "note": "whole batch = one wave of 128 clip pairs on 148 SMs; latency, not throughput"}

    def cfg_default10():   # the reference's default feature shape: 10 coefficients of 26 filters
        ks10 = KeywordSpotter(model, AudioConfig(), numcep=10, nfilt=26, device=dev)
        Bm = min(B, 16384)
        ms = timed(lambda: ks10(pcm[:Bm]), 5)
        ks._bind()
        return {"workload": "mfcc10_dscnn_s_b%d" % Bm, "clips_per_s": world * Bm / (ms * 1e-3), "ms": ms}

    def cfg3():   # BASELINE configs[2]: DS-CNN-L (276 ch, 5 blocks), 8192 clips per GPU = 65 536 over 8 GPUs
        Bl = min(B, 8192)
        torch.manual_seed(1)
        big = DepthwiseSeparableConv(channels=276, num_blocks=5).to(dev)
        feats = proc.extract_features(pcm[:Bl], numcep=NUMCEP, nfilt=NFILT).unsqueeze(1)
        ms = timed(lambda: big(feats), 3, warm=1)
        cps = Bl / (ms * 1e-3)
        # what the LAYERED design moves through HBM per clip (fp32 NHWC activations, 288-float pixel pitch, every
        # activation written once and read once; the last block's output is pooled on chip): its own floor, next to
        # the algorithmic roofline of a fully fused network that roofline3 reports
        h1, w1 = (NFRAMES + 4 - 10) // 2 + 1, (NUMCEP + 4 - 10) // 2 + 1
        px = h1 * w1 + sum((h1 + 2 * l) * (w1 + 2 * l) for l in range(5)) + sum((h1 + 2 * l + 2) * (w1 + 2 * l + 2) for l in range(4))
        layered = px * 288 * 4
        return {"workload": "mfcc40 features -> DS-CNN-L (276 ch, 5 blocks) forward, %d clips/GPU" % Bl,
                "clips_per_s": world * cps, "ms": ms,
                "roofline": roofline3(WORK["dscnn_l_model"], cps, peaks, fp32_tflops, "dscnn_conv1_kernel + dscnn_block_ws_kernel (x5)", ms),
                "layered_hbm": {"bytes_per_clip": layered, "achieved_gbs": cps * layered / 1e9, "peak_gbs": peaks["hbm_gbs"],
                                "frac": cps * layered / 1e9 / peaks["hbm_gbs"]}}

    def cfg4():   # BASELINE configs[3]: log-mel + CRNN/GRU, batch 16 384
        Bc = min(B, 16384)
        torch.manual_seed(2)
        crnn = CRNN(n_features=40).to(dev)
        x = pcm[:Bc]
        ms = timed(lambda: crnn(proc.extract_logfbank(x, nfilt=40).unsqueeze(1)), 5)
        cps = Bc / (ms * 1e-3)
        return {"workload": "logmel40 + CRNN(conv 16 ch, GRU 64) PCM -> logits, %d clips/GPU" % Bc,
                "clips_per_s": world * cps, "ms": ms,
                "roofline": roofline3(WORK["logmel_crnn"], cps, peaks, fp32_tflops,
                                      "mfcc512_ws_kernel(logmel) + conv1_tc + dense_gemm_tf32x3 + gru_recurrent_tc", ms)}

    def cfg5():   # BASELINE configs[4]: training step, NCCL allreduce of the flat gradient buffer when world > 1
        res = {}
        for Bt in (1028, 8192):
            torch.manual_seed(3)
            ts = KWSTrainStep(DepthwiseSeparableConv(), TrainingConfig(batch_size=Bt), device=dev)
            gg = torch.Generator(device=dev).manual_seed(rank)
            x = torch.randn(Bt, 1, NFRAMES, 10, device=dev, generator=gg)
            y = torch.randint(0, NCLS, (Bt,), device=dev, generator=gg)
            ms = timed(lambda: ts.step(x, y), 10, warm=3)
            e = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
            e[0].record(); ts.forward_backward(x, y)
            e[1].record(); ts.allreduce_gradients()
            e[2].record(); ts.optimizer_step()
            e[3].record()
            barrier()
            sps = Bt / (ms * 1e-3)
            res["per_gpu_batch_%d" % Bt] = {
                "samples_per_s": world * sps, "ms_per_step": ms,
                "last_step_us": {"fwd_bwd": 1e3 * maxr(e[0].elapsed_time(e[1])), "allreduce": 1e3 * maxr(e[1].elapsed_time(e[2])),
                                 "adam": 1e3 * maxr(e[2].elapsed_time(e[3]))},
                "allreduce_bytes": ts.grads.numel() * 4, "world": world,
                "roofline": roofline3(WORK["train_f10"], sps, peaks, fp32_tflops, "train step kernels (csrc/train.cu)", ms)}
            del ts
        res["workload"] = "training step, reference network, features 99x10, 12 classes, fwd+bwd+allreduce+Adam"
        return res

    leg("config2_mfcc_only", cfg2)
    leg("config1_batch256", cfg1)
    leg("default10_pcm_to_logits", cfg_default10)
    leg("config3_dscnn_l", cfg3)
    leg("config4_logmel_crnn", cfg4)
    leg("config5_training", cfg5)
    return ex


_JSON_FD = None


def _claim_stdout():
    """The contract is ONE JSON line on stdout.  Libraries write there too (NCCL prints its version line to
    stdout when NCCL_DEBUG is set in the environment), so everything this process and its children write to
    fd 1 goes to stderr and the JSON line goes to the original stdout."""
    global _JSON_FD
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)


def _emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


def main():
    _claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--clips", type=int, default=65536, help="clips per GPU per step")
    ap.add_argument("--e2e-clips", type=int, default=32768, help="clips per step of the host-buffer path")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extra", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
