#!/usr/bin/env python
"""Headline benchmark: clips/sec of batched MFCC extraction (1 s, 16 kHz, 40 coefficients /
40 mel filters, PCM -> MFCC) -- BASELINE.json config[1], "MFCC feature extraction only,
batch=65536 synthetic clips, 1/2/4/8 B200".

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--clips B]

One process per GPU (torchrun for N>1), batch sharded by rank with no collective on the data
path (clips are independent): weak scaling, every rank extracts its own B clips per step.
Prints ONE JSON line on rank 0.  See DESIGN.md "Measurement" for every field.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_SAMPLES = 16000
NUMCEP = NFILT = 40
NFRAMES = 99
BYTES_PER_CLIP = N_SAMPLES * 4 + NFRAMES * NUMCEP * 4      # 79 840 B algorithmic
METRIC = "clips/sec (1 s 16 kHz PCM -> 40-MFCC)"


def _peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


# ---------------------------------------------------------------------------------------
# CPU arm: the reference's feature path (psf 0.6 restated in oracle/psf_mfcc.py), one clip per
# call like the reference's DataLoader workers (train.py:108-114), on all host cores.
# ---------------------------------------------------------------------------------------
def _cpu_worker(args):
    seed, n = args
    import numpy as np
    from oracle import psf_mfcc
    rs = np.random.RandomState(seed)
    pcm = (rs.rand(n, N_SAMPLES).astype(np.float32) * 2.0 - 1.0)
    t0 = time.perf_counter()
    acc = 0.0
    for x in pcm:
        acc += float(psf_mfcc.extract_features(x, numcep=NUMCEP, nfilt=NFILT)[0, 0])
    return time.perf_counter() - t0, acc


def cpu_clips_per_sec(total_clips, workers):
    """clips/s of the oracle port over `total_clips` clips split across `workers` processes."""
    import multiprocessing as mp
    per = max(1, total_clips // workers)
    ctx = mp.get_context("fork")
    with ctx.Pool(workers) as pool:
        pool.map(_cpu_worker, [(i, 2) for i in range(workers)])        # warm-up / imports
        t0 = time.perf_counter()
        pool.map(_cpu_worker, [(100 + i, per) for i in range(workers)])
        dt = time.perf_counter() - t0
    return per * workers / dt, per * workers


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    sample = max(cores * 40, 256)
    vals = []
    for _ in range(args.warmup):
        cpu_clips_per_sec(max(cores * 4, 32), cores)
    t_all = time.perf_counter()
    for _ in range(args.steps):
        v, n = cpu_clips_per_sec(sample, cores)
        vals.append(v)
    total_s = time.perf_counter() - t_all
    value = sum(vals) / len(vals)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "clips/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total_s / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": {"workload": "mfcc40_b65536", "numcep": NUMCEP, "nfilt": NFILT, "samples": N_SAMPLES,
                   "note": "psf-0.6 restatement (oracle/psf_mfcc.py), one clip per call, process pool"},
        "cpu_baseline": {"value": value, "unit": "clips/s", "cores": cores, "kind": "port",
                         "sample": f"{sample} clips per step x {args.steps} steps"},
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


def run_ours(args):
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    # CPU baseline first (fork before CUDA is initialised), rank 0 at N=1 only
    cpu_base = None
    if world == 1 and not args.no_cpu_baseline:
        cores = os.cpu_count() or 1
        sample = max(cores * 60, 480)
        v, n = cpu_clips_per_sec(sample, cores)
        cpu_base = {"value": v, "unit": "clips/s", "cores": cores, "kind": "port",
                    "sample": f"{n} clips, oracle/psf_mfcc.py (psf 0.6 restatement), {cores}-process pool"}

    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa = _bind_to_gpu_numa_node(local) if world > 1 else None
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    import kws_b200  # noqa: F401
    from kws_b200.audio_processor import AudioConfig, AudioProcessor

    B = args.clips
    proc = AudioProcessor(None, AudioConfig(), device=dev)
    g = torch.Generator(device=dev).manual_seed(1234 + rank)
    pcm = torch.rand((B, N_SAMPLES), generator=g, device=dev, dtype=torch.float32) * 2.0 - 1.0

    def step():
        return proc.extract_features(pcm, numcep=NUMCEP, nfilt=NFILT)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    for _ in range(max(args.warmup, 3)):
        out = step()
    barrier()
    l0 = proc.launch_count()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    barrier()
    ev[0].record()
    for i in range(args.steps):
        out = step()
        ev[i + 1].record()
    barrier()
    launches = proc.launch_count() - l0
    total_ms = ev[0].elapsed_time(ev[-1])
    per_launch_ms = [ev[i].elapsed_time(ev[i + 1]) for i in range(args.steps)]
    t = torch.tensor([total_ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t.item())
    value = world * B * args.steps / (total_ms * 1e-3)
    del out

    # ---- e2e: HOST buffers through the public API, H2D and D2H inside the timed region ----
    Be = min(B, args.e2e_clips)
    host = torch.empty((Be, N_SAMPLES), dtype=torch.float32, pin_memory=True)
    host.copy_(pcm[:Be])
    host_out = torch.empty((Be, NFRAMES, NUMCEP), dtype=torch.float32, pin_memory=True)
    for _ in range(2):
        proc.extract_features_host(host, out=host_out, numcep=NUMCEP, nfilt=NFILT)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        proc.extract_features_host(host, out=host_out, numcep=NUMCEP, nfilt=NFILT)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    te = torch.tensor([e2e_s], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_value = world * Be * args.steps / float(te.item())
    # ---- secondary: the same host-buffer call fed 16-bit PCM (the wav sample format), half the H2D bytes ----
    e2e_i16 = None
    try:
        host16 = torch.empty((Be, N_SAMPLES), dtype=torch.int16, pin_memory=True)
        host16.copy_((pcm[:Be] * 32767.0).to(torch.int16))
        for _ in range(2):
            proc.extract_features_host(host16, out=host_out, numcep=NUMCEP, nfilt=NFILT)
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            proc.extract_features_host(host16, out=host_out, numcep=NUMCEP, nfilt=NFILT)
        torch.cuda.synchronize()
        t16 = torch.tensor([time.perf_counter() - t0], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t16, op=dist.ReduceOp.MAX)
        e2e_i16 = world * Be * args.steps / float(t16.item())
        del host16
    except Exception as e:
        e2e_i16 = repr(e)
    # ---- secondary (not the headline): BASELINE configs[0]/[2] shape, PCM -> MFCC -> DS-CNN logits ----
    extra = None
    try:
        from kws_b200.models import DepthwiseSeparableConv
        from kws_b200.pipeline import KeywordSpotter
        torch.manual_seed(0)
        model = DepthwiseSeparableConv().to(dev)
        with torch.no_grad():
            for prm in model.parameters():
                if prm.dim() == 1:
                    prm.normal_(0, 0.1)
        ks = KeywordSpotter(model, AudioConfig(), numcep=NUMCEP, nfilt=NFILT, device=dev)
        Bm = min(B, 16384)
        feats = step()[:Bm].unsqueeze(1).contiguous()

        def timed(fn, reps=5):
            for _ in range(2):
                fn()
            barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps):
                fn()
            e1.record()
            barrier()
            t_ms = torch.tensor([e0.elapsed_time(e1) / reps], device=dev, dtype=torch.float64)
            if world > 1:
                dist.all_reduce(t_ms, op=dist.ReduceOp.MAX)
            return float(t_ms.item())

        l1 = model.launch_count()
        ms_model = timed(lambda: model(feats))
        ms_e2l = timed(lambda: ks(pcm[:Bm]))
        tf32_peak = None
        try:
            tf32_peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops"]) / 2.0
        except Exception:
            pass
        model_cps = world * Bm / (ms_model * 1e-3)
        # the reference's default feature shape (10 coefficients from 26 filters -> 99x10 map, two clips per slot)
        ks10 = KeywordSpotter(model, AudioConfig(), numcep=10, nfilt=26, device=dev)
        feats10 = proc.extract_features(pcm[:Bm], numcep=10, nfilt=26).unsqueeze(1).contiguous()
        ms_model10 = timed(lambda: model(feats10))
        ms_e2l10 = timed(lambda: ks10(pcm[:Bm]))
        # host buffers -> logits through kws_pcm_to_logits_host(_i16): PCM in over PCIe, 48 B of logits back
        host_f32 = pcm[:Bm].cpu().pin_memory()
        host_i16 = (pcm[:Bm] * 32767.0).round().to(torch.int16).cpu().pin_memory()
        host_logits = torch.empty((Bm, model.num_classes), dtype=torch.float32).pin_memory()

        def timed_host(x, reps=4):
            ks.infer_host(x, out=host_logits)
            barrier()
            t0 = time.perf_counter()
            for _ in range(reps):
                ks.infer_host(x, out=host_logits)
            torch.cuda.synchronize()
            t = torch.tensor([(time.perf_counter() - t0) / reps], device=dev, dtype=torch.float64)
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return world * Bm / float(t.item())

        e2l_host_f32 = timed_host(host_f32)
        e2l_host_i16 = timed_host(host_i16)
        del host_f32, host_i16
        extra = {"workload": "mfcc40 + reference DS-CNN (64 ch, 4 blocks), %d clips/GPU" % Bm,
                 "e2e_pcm_to_logits_host_f32_clips_per_s": e2l_host_f32,
                 "e2e_pcm_to_logits_host_int16_clips_per_s": e2l_host_i16,
                 "model_forward_clips_per_s": model_cps,
                 "pcm_to_logits_clips_per_s": world * Bm / (ms_e2l * 1e-3),
                 "model_kernel": "dscnn_fused_kernel (tcgen05 kind::tf32)",
                 "model_gemm_flop_per_clip": 45.4e6,
                 "model_tensor_frac_of_tf32_peak": (model_cps / world * 45.4e6 / (tf32_peak * 1e12)) if tf32_peak else None,
                 "tf32_peak_tflops": tf32_peak, "tf32_peak_source": "MEASURED_PEAKS.json bf16_tflops / 2",
                 "default10_model_forward_clips_per_s": world * Bm / (ms_model10 * 1e-3),
                 "default10_pcm_to_logits_clips_per_s": world * Bm / (ms_e2l10 * 1e-3),
                 "model_launches": int(model.launch_count() - l1),
                 "e2e_int16_host_clips_per_s": e2e_i16}
    except Exception as e:  # the headline line must still print
        extra = {"error": repr(e)}
    clocks = sampler.stop() if sampler else None

    if rank == 0:
        peak, peak_src = _peaks()
        kern_ms = sum(per_launch_ms) / len(per_launch_ms)
        achieved = BYTES_PER_CLIP * B / (kern_ms * 1e-3) / 1e9
        traffic = None
        tp = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tp):
            try:
                traffic = json.load(open(tp)).get("dram_bytes_per_launch")
            except Exception:
                traffic = None
        line = {
            "metric": METRIC, "value": value, "unit": "clips/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": total_ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "mfcc40_b65536", "clips_per_gpu": B, "samples": N_SAMPLES,
                       "numcep": NUMCEP, "nfilt": NFILT, "nfft": 512, "parallelism": f"batch-sharded x{world}",
                       "l2": "inputs (4.2 GB/step) larger than L2, no flush needed",
                       "e2e_clips_per_step": Be, "numa_bound_cpus": numa},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": BYTES_PER_CLIP * B, "kernel": "mfcc512_ws_kernel",
                         "kernel_ms": kern_ms},
            "cpu_baseline": cpu_base,
            "e2e": {"value": e2e_value, "unit": "clips/s", "h2d_bytes_per_step": Be * N_SAMPLES * 4,
                    "d2h_bytes_per_step": Be * NFRAMES * NUMCEP * 4},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "extra": extra,
        }
        _emit(line)
    if world > 1:
        dist.destroy_process_group()


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
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
