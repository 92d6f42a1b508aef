#!/usr/bin/env python
"""Benchmark of the inference hot path (BASELINE.json metric: images/sec ResNet-18 bs=256).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A step is one pass of the whole operator (normalise -> ResNet-18 -> softmax -> top-5) over one
batch of 256 synthetic 224x224 u8 images per GPU (weak scaling: per-GPU work fixed).  For N > 1 the
driver launches one rank per GPU with torchrun; ranks share nothing on the data path (the path has
no collective), torch.distributed supplies only the barrier and the max-over-ranks reduction.

`value`  : whole-job images/s with inputs already in HBM, timed with CUDA events on the launching
           stream, max over ranks.
`e2e`    : the same metric through the reference-facing C-ABI call (b2i_predict_u8) on HOST pinned
           buffers: H2D of the u8 batch and D2H of top-5 inside the timed region.
`--impl reference` times the CPU statement of the same path (oracle/: torch-CPU restatement of the
tch calls; the reference's own Rust binary cannot be built here) on the box's host cores.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

BATCH = 256
NCU_DRAM_BYTES_PER_CONV_LAUNCH = 97420146   # see roofline["traffic"] below
K_TOP = 5
MODEL = "resnet18"
WORKLOAD = "ResNet-18 (resnet18.ot layout, seeded synthetic weights) synthetic 224x224 u8, batch 256 per GPU, bf16"
ROTATE = 8  # distinct input batches cycled through, 8 x 38.5 MB = 308 MB > 126 MB L2


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": d["hbm_gbs"], "tflops": d.get("bf16_tflops_sustained", d["bf16_tflops"]),
                "tflops_burst": d["bf16_tflops"], "source": "measured (MEASURED_PEAKS.json, sustained bf16)"}
    return {"hbm_gbs": 6650.0, "tflops": 1400.0, "tflops_burst": 1590.0, "source": "fallback (B200_PROFILING.md)"}


class ClockSampler:
    """SM clock and throttle reasons polled every 20 ms through NVML while the timed regions run
    (falls back to `nvidia-smi -lms 100` when the NVML binding is missing)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.sm, self.max_mhz, self.reasons, self.stop_flag = [], None, set(), False
        self.proc = self.thread = None
        try:
            import pynvml
            pynvml.nvmlInit()
            h = pynvml.nvmlDeviceGetHandleByIndex(gpu_index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))
            names = {"hw_slowdown": getattr(pynvml, "nvmlClocksEventReasonHwSlowdown", 0x8),
                     "hw_thermal_slowdown": getattr(pynvml, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
                     "sw_thermal_slowdown": getattr(pynvml, "nvmlClocksEventReasonSwThermalSlowdown", 0x20),
                     "sw_power_cap": getattr(pynvml, "nvmlClocksEventReasonSwPowerCap", 0x4)}

            def poll():
                while not self.stop_flag:
                    try:
                        self.sm.append(float(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)))
                        get = getattr(pynvml, "nvmlDeviceGetCurrentClocksEventReasons", None) or \
                            pynvml.nvmlDeviceGetCurrentClocksThrottleReasons
                        mask = int(get(h))
                        for k, bit in names.items():
                            if mask & bit:
                                self.reasons.add(k)
                    except Exception:
                        pass
                    time.sleep(0.02)

            self.thread = threading.Thread(target=poll, daemon=True)
            self.thread.start()
        except Exception:
            self.thread = None
            try:
                self.proc = subprocess.Popen(["nvidia-smi", "-i", str(gpu_index), "--query-gpu=" + self.Q,
                                              "--format=csv,noheader,nounits", "-lms", "100"],
                                             stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
                self.rows = []
                self.reader = threading.Thread(target=lambda: [self.rows.append([c.strip() for c in ln.split(",")])
                                                               for ln in self.proc.stdout], daemon=True)
                self.reader.start()
            except OSError:
                self.proc = None

    def stop(self):
        if self.thread:
            self.stop_flag = True
            self.thread.join(timeout=2)
            return {"sm_mhz": statistics.median(self.sm) if self.sm else None, "sm_max_mhz": self.max_mhz,
                    "reasons": sorted(self.reasons), "samples": len(self.sm), "source": "nvml, 20 ms"}
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no NVML and no nvidia-smi"], "samples": 0}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm = [float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 6 for i in range(4) if r[2 + i].lower() == "active"})
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm), "source": "nvidia-smi, 100 ms"}


def cpu_oracle_rate(budget_s, batch=64):
    """Oracle (torch CPU fp32, all host threads) on a bounded sample of the same workload."""
    import torch
    from oracle import synth, tch_oracle
    torch.set_num_threads(os.cpu_count() or 1)
    w = synth.resnet18_weights()
    imgs = synth.synthetic_images(batch, seed=0)
    tch_oracle.predict_u8(MODEL, w, imgs[:8], k=K_TOP)  # warm-up
    done, t0 = 0, time.perf_counter()
    while True:
        tch_oracle.predict_u8(MODEL, w, imgs, k=K_TOP, batch=batch)
        done += batch
        el = time.perf_counter() - t0
        if el >= budget_s or done >= 4096:
            break
    return done / el, torch.get_num_threads(), "%d images in batches of %d (%.1f s)" % (done, batch, el)


def run_reference(args, rank, world):
    """--impl reference: the CPU restatement of the path, each step a 64-image sample of the batch."""
    if rank != 0:
        return
    import torch
    from oracle import synth, tch_oracle
    torch.set_num_threads(os.cpu_count() or 1)
    sample = 64
    w = synth.resnet18_weights()
    imgs = synth.synthetic_images(sample, seed=0)
    for _ in range(max(1, min(args.warmup, 2))):
        tch_oracle.predict_u8(MODEL, w, imgs, k=K_TOP, batch=sample)
    steps = min(args.steps, 40)
    t0 = time.perf_counter()
    for _ in range(steps):
        tch_oracle.predict_u8(MODEL, w, imgs, k=K_TOP, batch=sample)
    el = time.perf_counter() - t0
    val = steps * sample / el
    desc = "%d steps x %d-image sample of the 256-image batch, torch CPU fp32 oracle" % (steps, sample)
    print(json.dumps({
        "impl": "reference", "metric": "images/sec ResNet-18 bs=256", "value": val, "unit": "images/s",
        "n_gpus": args.gpus, "steps": steps, "warmup": args.warmup, "ms_per_step": 1000.0 * el / steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "batch_per_gpu": BATCH, "sample_per_step": sample},
        "cpu_baseline": {"value": val, "unit": "images/s", "cores": torch.get_num_threads(), "kind": "port", "sample": desc},
        "e2e": {"value": val, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


def bind_to_gpu_numa_node(gpu_index):
    """Pin this process to the CPUs NVML reports as local to its GPU before any pinned host memory is
    touched: with eight ranks streaming 38.5 MB per step each, buffers that land on the other socket
    make the end-to-end number a measurement of the inter-socket link.  Returns the CPU count or None."""
    try:
        import pynvml
        pynvml.nvmlInit()
        pynvml.nvmlDeviceSetCpuAffinity(pynvml.nvmlDeviceGetHandleByIndex(gpu_index))
        return len(os.sched_getaffinity(0))
    except Exception:
        return None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cpu-budget", type=float, default=12.0, help="seconds of oracle work for cpu_baseline")
    ap.add_argument("--tile-n-max", type=int, default=0)
    ap.add_argument("--conv-mode", type=int, default=-1)
    ap.add_argument("--micro-batch", type=int, default=0)
    ap.add_argument("--profile-out", default="")
    ap.add_argument("--patch-conv", type=int, default=None)
    ap.add_argument("--pair-mma", type=int, default=None)
    ap.add_argument("--fuse-downsample", type=int, default=None)
    ap.add_argument("--precision", default="bf16", choices=["bf16", "fp16"],
                    help="16-bit storage format of activations and weights (fp32 accumulate either way)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import numpy as np
    import torch
    import torch.distributed as dist
    import dmlc_b200 as b2
    from oracle import synth  # synthetic weights/images only; the oracle is never on the timed path

    torch.cuda.set_device(local)
    numa = bind_to_gpu_numa_node(local) if world > 1 else None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ot = synth.ensure_ot(MODEL, os.path.join(ROOT, "tests", "_cache", "resnet18.ot")) if rank == 0 else None
    if world > 1:
        dist.barrier()
        ot = os.path.join(ROOT, "tests", "_cache", "resnet18.ot")

    ctx = b2.Context((local,))
    model = b2.Model(ctx, MODEL, ot, precision=b2.PREC_FP16 if args.precision == "fp16" else b2.PREC_BF16, max_batch=BATCH, micro_batch=args.micro_batch,
                     tile_n_max=args.tile_n_max, conv_mode=args.conv_mode, patch_conv=args.patch_conv, pair_mma=args.pair_mma,
                     fuse_downsample=args.fuse_downsample)
    n_micro = (BATCH + (args.micro_batch or BATCH) - 1) // (args.micro_batch or BATCH)

    host = synth.synthetic_images(BATCH, seed=rank)
    pin = b2.PinnedBuffer(ROTATE * BATCH * b2.IMAGE_BYTES)
    hostbufs = pin.array.reshape(ROTATE, BATCH, 224, 224, 3)
    for r in range(ROTATE):
        hostbufs[r] = np.roll(host, r, axis=0)
    dev_in = [torch.from_numpy(hostbufs[r]).cuda() for r in range(ROTATE)]
    d_idx = torch.empty((BATCH, K_TOP), dtype=torch.int32, device="cuda")
    d_prob = torch.empty((BATCH, K_TOP), dtype=torch.float32, device="cuda")
    stream = torch.cuda.current_stream()

    def step_device(i):
        model.forward_device(dev_in[i % ROTATE].data_ptr(), BATCH, K_TOP, d_idx.data_ptr(), d_prob.data_ptr(),
                             None, stream.cuda_stream)

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    # ---------------- device-resident timing (value)
    sampler = ClockSampler(local) if rank == 0 else None   # polls through the device-timed AND the e2e regions
    for i in range(args.warmup):
        step_device(i)
    sync_all()
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    evs[0].record(stream)
    for i in range(args.steps):
        step_device(i)
        evs[i + 1].record(stream)
    sync_all()
    step_ms = [evs[i].elapsed_time(evs[i + 1]) for i in range(args.steps)]
    total_ms = evs[0].elapsed_time(evs[-1])
    if world > 1:
        t = torch.tensor([total_ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms = float(t.item())
    value = world * BATCH * args.steps / (total_ms / 1000.0)

    # ---------------- end to end through the C ABI on host buffers (e2e)
    # (a) blocking calls, one batch at a time: the batch latency a single caller sees
    for i in range(args.warmup):
        model.predict_u8(hostbufs[i % ROTATE], k=K_TOP)
    sync_all()
    e2e_ms = []
    for i in range(args.steps):
        t0 = time.perf_counter()
        idx, prob = model.predict_u8(hostbufs[i % ROTATE], k=K_TOP)
        e2e_ms.append((time.perf_counter() - t0) * 1000.0)
    blocking_total = sum(e2e_ms) / 1000.0
    # (b) two requests in flight (b2i_submit_u8 / b2i_wait): H2D of batch i+1 overlaps the forward
    # of batch i. Every step still copies its 38.5 MB in from pinned host memory and its top-5 out.
    out_idx = [np.empty((BATCH, K_TOP), np.int32) for _ in range(2)]
    out_prob = [np.empty((BATCH, K_TOP), np.float32) for _ in range(2)]
    sync_all()
    t_all = time.perf_counter()
    prev = None
    for i in range(args.steps):
        t = model.submit_u8(hostbufs[i % ROTATE], out_idx[i & 1], out_prob[i & 1])
        if prev is not None:
            model.wait(prev)
        prev = t
    model.wait(prev)
    e2e_total = time.perf_counter() - t_all
    if world > 1:
        t = torch.tensor([e2e_total, blocking_total], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_total, blocking_total = float(t[0].item()), float(t[1].item())
    e2e_value = world * BATCH * args.steps / e2e_total
    clocks = sampler.stop() if sampler else None
    e2e_blocking_value = world * BATCH * args.steps / blocking_total

    # ---------------- per-launch share (rank 0): CUDA events around every launch of one step
    prof = None
    if rank == 0:
        mb = args.micro_batch or BATCH
        acc = {}
        reps = 5
        for r in range(reps + 1):
            rows = model.profile(dev_in[r % ROTATE].data_ptr(), mb, K_TOP, d_idx.data_ptr(), d_prob.data_ptr(),
                                 stream.cuda_stream)
            if r == 0:
                continue
            for j, (name, ms) in enumerate(rows):
                acc.setdefault((j, name), []).append(ms)
        prof = [(name, statistics.mean(v)) for (j, name), v in sorted(acc.items())]
        if args.profile_out:
            with open(args.profile_out, "w") as f:
                f.write("launch,ms_mean_of_%d,images\n" % reps)
                for name, ms in prof:
                    f.write("%s,%.5f,%d\n" % (name, ms, mb))

    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return

    pk = peaks()
    flops_img = model.flops_per_image
    mb = args.micro_batch or BATCH
    conv_ms = sum(ms for name, ms in prof if name.startswith(("conv:", "fc:")))
    conv_launches = sum(1 for name, _ in prof if name.startswith(("conv:", "fc:")))
    all_ms = sum(ms for _, ms in prof)
    achieved = flops_img * mb / (conv_ms / 1000.0) / 1e12 if conv_ms > 0 else 0.0
    roofline = {
        "kernel": "tcgen05 implicit-GEMM family (stem_kernel, conv3x3_c64_patch_kernel, conv3x3_flat_pair_kernel, "
                  "conv_umma_pair_kernel): all %d conv/FC launches of a step" % conv_launches,
        "bound": "tensor", "achieved": achieved, "peak": pk["tflops"], "unit": "TFLOP/s", "frac": achieved / pk["tflops"],
        # dram__bytes_read.sum + dram__bytes_write.sum averaged over the 18 conv/FC launches of one batch-256 step
        # (ncu, profiles/r02_ncu_dram_tensor_per_launch.csv); only meaningful for the default workload
        "traffic": NCU_DRAM_BYTES_PER_CONV_LAUNCH if (mb == BATCH and conv_launches == 18) else None,
        "peak_source": pk["source"],
        "flops_per_launch_avg": flops_img * mb / max(1, conv_launches),
        "ms_per_launch_avg": conv_ms / max(1, conv_launches),
        "share_of_step": conv_ms / all_ms if all_ms else None,
        "whole_network_frac_of_tensor_peak": value / world * flops_img / 1e12 / pk["tflops"],
    }
    cpu_val, cores, sample = cpu_oracle_rate(args.cpu_budget)
    srt = sorted(e2e_ms)
    dsrt = sorted(step_ms)
    pct = lambda a, q: a[min(len(a) - 1, int(q * len(a)))]
    out = {
        "metric": "images/sec ResNet-18 bs=256", "value": value, "unit": "images/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": total_ms / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": args.precision, "data": "synthetic",
        "config": {"workload": WORKLOAD.replace(", bf16", ", " + args.precision), "batch_per_gpu": BATCH, "micro_batch": mb, "top_k": K_TOP,
                   "l2": "inputs rotate over %d distinct batches (%d MB) > 126 MB L2; activations (> 1 GB per step) never fit" %
                         (ROTATE, ROTATE * BATCH * b2.IMAGE_BYTES // 2 ** 20),
                   "parallelism": "independent batch shards, one process per GPU, no collective",
                   "host_cpus_bound_to_gpu_numa_node": numa},
        "e2e": {"value": e2e_value, "unit": "images/s", "h2d_bytes_per_step": BATCH * b2.IMAGE_BYTES,
                "d2h_bytes_per_step": BATCH * K_TOP * 8,
                "api": "b2i_submit_u8 / b2i_wait, two requests in flight, pinned host buffers",
                "blocking": {"value": e2e_blocking_value, "unit": "images/s", "api": "b2i_predict_u8",
                             "batch_latency_ms": {"p50": pct(srt, 0.50), "p99": pct(srt, 0.99)}}},
        "device_batch_latency_ms": {"p50": pct(dsrt, 0.50), "p99": pct(dsrt, 0.99)},
        "gpu_launches": model.launches_per_microbatch * n_micro * args.steps,
        "clocks": clocks,
        "roofline": roofline,
        "cpu_baseline": {"value": cpu_val, "unit": "images/s", "cores": cores, "kind": "port", "sample": sample},
        "launch_ms": {name: round(ms, 4) for name, ms in prof},
    }
    print(json.dumps(out))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
