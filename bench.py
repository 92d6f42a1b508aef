#!/usr/bin/env python
"""Benchmark of the inference hot path (BASELINE.json metric: images/sec ResNet-18 bs=256).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--model resnet18|alexnet] [--batch B]

A step is one pass of the whole operator (normalise -> network -> softmax -> top-5) over one batch of
synthetic 224x224 u8 images per GPU (weak scaling: per-GPU work fixed).  For N > 1 the driver launches one
rank per GPU with torchrun; ranks share nothing on the data path (the path has no collective),
torch.distributed supplies only the barrier and the max-over-ranks reduction.

`value`     whole-job images/s with inputs already in HBM, CUDA events on the launching stream, max over ranks.
`e2e`       the same metric through the reference-facing C-ABI call on HOST pinned buffers: H2D of the u8
            batch and D2H of top-5 inside the timed region; `e2e.h2d_ceiling` is a copy-only measurement of
            the same buffers on the same ranks (what the host side can feed at all).
`sustained` a >= 5 s back-to-back device-resident segment after the K-step regions, with SM clock, power and
            throttle reasons sampled through NVML: the K-step region is a 20 ms burst, this is not.
`roofline`  conv/FC launches of a step (CUDA-event pair per launch) against BOTH measured bf16 peaks
            (MEASURED_PEAKS.json: burst = best of 10 GEMMs, sustained = 4 s of GEMMs, power-capped); `frac`
            uses the burst peak while this workload's own sustained segment holds the full SM clock, else the
            sustained one.
`parity`    8 images through the timed model against the fp32 CPU restatement (absolute logit gap).
`fp16`      device-resident rate of the IEEE-half mode (the one that meets a flat 2e-2 on the logits).
`inproc`    (one process, >= 2 visible GPUs) the in-process sharder b2i_submit_u8 over all GPUs and the dual-job
            split, ResNet-18 on the first half of the GPUs and AlexNet on the second (reference
            src/services.rs:146-151,199-211), each against its solo rate.
`--impl reference` times the CPU statement of the same path (oracle/: torch-CPU restatement of the tch calls;
the reference's own Rust binary cannot be built here) on the box's host cores.
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

K_TOP = 5
ROTATE_BYTES = 300 * 2 ** 20   # distinct input batches cycled through: > 126 MB L2
NAMES = {"resnet18": "ResNet-18", "alexnet": "AlexNet"}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": d["hbm_gbs"], "tflops_sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]),
                "tflops_burst": d["bf16_tflops"], "source": "measured (MEASURED_PEAKS.json)"}
    return {"hbm_gbs": 6650.0, "tflops_sustained": 1400.0, "tflops_burst": 1590.0, "source": "fallback (B200_PROFILING.md)"}


def metric_name(model, batch):
    return "images/sec %s bs=%d" % (NAMES[model], batch)


def make_config(model, batch, micro_batch, weights_kind):
    """The workload description both arms print (the driver compares the dicts)."""
    rotate = max(2, -(-ROTATE_BYTES // (batch * 224 * 224 * 3)))
    return {"workload": "%s (%s.ot layout, %s weights) synthetic 224x224 u8, batch %d per GPU" %
                        (NAMES[model], model, weights_kind, batch),
            "batch_per_gpu": batch, "micro_batch": micro_batch or batch, "top_k": K_TOP,
            "l2": "inputs rotate over %d distinct batches (%d MB) > 126 MB L2" % (rotate, rotate * batch * 224 * 224 * 3 // 2 ** 20),
            "parallelism": "independent batch shards, one process per GPU, no collective"}


class ClockSampler:
    """SM clock, board power and throttle reasons polled every 20 ms through NVML while a timed region runs
    (falls back to `nvidia-smi -lms 100` when the NVML binding is missing)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.sm, self.power, self.max_mhz, self.reasons, self.stop_flag = [], [], None, set(), False
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
                        self.power.append(pynvml.nvmlDeviceGetPowerUsage(h) / 1000.0)
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
            return {"sm_mhz": statistics.median(self.sm) if self.sm else None, "sm_mhz_min": min(self.sm) if self.sm else None,
                    "sm_max_mhz": self.max_mhz, "power_w_max": max(self.power) if self.power else None,
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


def bench_weights(model):
    """(name -> tensor dict, path of the .ot, kind): the reference's pretrained file when one that authenticates
    against the LFS pointer is present, else the seeded synthetic set."""
    from oracle import ot_format, synth
    real = synth.real_weights(model)
    if real:
        return ot_format.read_ot(real), real, "pretrained"
    return synth.weights(model), None, "seeded synthetic"


def cpu_oracle_rate(model, budget_s, batch=64):
    """Oracle (torch CPU fp32, all host threads) on a bounded sample of the same workload."""
    import torch
    from oracle import synth, tch_oracle
    torch.set_num_threads(os.cpu_count() or 1)
    w, _, _ = bench_weights(model)
    imgs = synth.synthetic_images(batch, seed=0)
    tch_oracle.predict_u8(model, w, imgs[:8], k=K_TOP)  # warm-up
    done, t0 = 0, time.perf_counter()
    while True:
        tch_oracle.predict_u8(model, w, imgs, k=K_TOP, batch=batch)
        done += batch
        el = time.perf_counter() - t0
        if el >= budget_s or done >= 4096:
            break
    return done / el, torch.get_num_threads(), "%d images in batches of %d (%.1f s)" % (done, batch, el)


def cpu_reference_mode(model):
    """The two other CPU figures SURVEY.md 8(d) asks for beside the batched one: the restatement run the way the reference
    runs it (one image per forward, services.rs:421,493) and its preprocessing (resize + crop of a 500 x 375 image)."""
    import numpy as np
    from oracle import stb_resize, synth, tch_oracle
    w, _, _ = bench_weights(model)
    imgs = synth.synthetic_images(12, seed=3)
    tch_oracle.predict_u8(model, w, imgs[:1], k=1, batch=1)
    t0 = time.perf_counter()
    for i in range(len(imgs)):
        tch_oracle.predict_u8(model, w, imgs[i:i + 1], k=1, batch=1)
    b1 = len(imgs) / (time.perf_counter() - t0)
    raw = np.random.default_rng(5).integers(0, 256, (375, 500, 3), dtype=np.uint8)
    stb_resize.resize_crop_224(raw)
    t0 = time.perf_counter()
    for _ in range(8):
        stb_resize.resize_crop_224(raw)
    return {"batch1_images_per_s": b1, "preprocess_ms_per_image": 1000.0 * (time.perf_counter() - t0) / 8,
            "note": "one image per forward as the reference does (torch CPU, all host threads); resize + crop of a 500 x 375 image by the numpy restatement of stb_image_resize, one thread (test infrastructure, not the reference's C code)"}


def run_reference(args, rank, world):
    """--impl reference: the CPU restatement of the path, each step a bounded sample of the batch."""
    if rank != 0:
        return
    import torch
    from oracle import synth, tch_oracle
    torch.set_num_threads(os.cpu_count() or 1)
    sample = min(64, args.batch)
    w, _, kind = bench_weights(args.model)
    imgs = synth.synthetic_images(sample, seed=0)
    for _ in range(max(1, min(args.warmup, 2))):
        tch_oracle.predict_u8(args.model, w, imgs, k=K_TOP, batch=sample)
    steps = min(args.steps, 40)
    t0 = time.perf_counter()
    for _ in range(steps):
        tch_oracle.predict_u8(args.model, w, imgs, k=K_TOP, batch=sample)
    el = time.perf_counter() - t0
    val = steps * sample / el
    desc = "%d steps x %d-image sample of the %d-image batch, torch CPU fp32 restatement of the tch path" % (steps, sample, args.batch)
    print(json.dumps({
        "impl": "reference", "metric": metric_name(args.model, args.batch), "value": val, "unit": "images/s",
        "n_gpus": args.gpus, "steps": steps, "warmup": args.warmup, "ms_per_step": 1000.0 * el / steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": make_config(args.model, args.batch, args.micro_batch, kind),
        "cpu_baseline": {"value": val, "unit": "images/s", "cores": torch.get_num_threads(), "kind": "port", "sample": desc},
        "e2e": {"value": val, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


def from_files_record(b2, ctx, ot, prec):
    """File names -> answers through the reference-facing interface (Leader.run_job, services.rs:404-433 / 476-496) over the
    committed reference JPEGs, cycled: images/s with one PIL decode at a time on the calling thread (the reference's
    arrangement) and with the library's decode stage (nvJPEG on host threads + GPU where it loads, else a PIL pool)."""
    import shutil
    import tempfile
    root = os.path.join(ROOT, "tests", "golden", "imagenet_1k", "train")
    labels = b2.load_labels(os.path.join(ROOT, "tests", "golden", "synset_words.txt"))
    have = set(os.listdir(root))
    lab = [(w, name) for w, name in labels if w in have]
    if not lab:
        return None
    rec = {"api": "Leader.run_job(batch=256, in_flight=2) from image ids", "jpeg_files": len(lab), "unit": "images/s"}
    wd = tempfile.mkdtemp()
    os.symlink(os.path.abspath(ot), os.path.join(wd, "resnet18.ot"))
    try:
        for key, decoder, threads, ids, passes in (("pil_on_calling_thread", "pil", 1, lab[:128], 2),
                                                   ("decode_stage", "auto", None, lab * max(1, 1024 // len(lab)), 3)):
            member = b2.MemberState(ctx, labels, weight_dir=wd, image_root=root, precision=prec, max_batch=256,
                                    models=("resnet18",), decoder=decoder, decode_threads=threads)
            try:
                best = 0.0
                for rep in range(passes):   # the first pass warms the lanes and the decoder up
                    leader = b2.Leader(member, ids)
                    leader.resnet18.assigned_member_ids = [0]
                    t0 = time.perf_counter()
                    leader.run_job(leader.resnet18, batch=256)
                    dt = time.perf_counter() - t0
                    if rep:
                        best = max(best, leader.resnet18.finished_prediction_count / dt)
                rec[key] = {"value": best, "images": len(ids), "decoder": member.decoder, "host_threads": member.pool._max_workers}
                if member.decoder == "nvjpeg":
                    rec[key]["backend"] = ctx.jpeg_backend()
            finally:
                member.close()
        if rec["pil_on_calling_thread"]["value"] > 0:
            rec["ratio"] = rec["decode_stage"]["value"] / rec["pil_on_calling_thread"]["value"]
    finally:
        shutil.rmtree(wd, ignore_errors=True)
    return rec


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


def inproc_record(b2, synth, ot_paths, n_gpus, steps, seconds=2.0):
    """One process driving all visible GPUs through the library's own sharder (b2i_submit_u8: per-GPU copy and
    compute streams, pinned staging, two requests in flight), then the dual-job split."""
    import numpy as np
    out = {"gpus": n_gpus}
    ctx = b2.Context(tuple(range(n_gpus)))
    try:
        def e2e_rate(model, first, count, batch_per_gpu, secs):
            n = batch_per_gpu * count
            pin = b2.PinnedBuffer(2 * n * b2.IMAGE_BYTES)
            bufs = pin.array.reshape(2, n, 224, 224, 3)
            bufs[0] = np.resize(synth.synthetic_images(min(n, 512), seed=3), bufs[0].shape)
            bufs[1] = np.roll(bufs[0], 1, axis=0)
            oi = [np.empty((n, K_TOP), np.int32) for _ in range(2)]
            op = [np.empty((n, K_TOP), np.float32) for _ in range(2)]
            for i in range(3):
                model.predict_u8(bufs[i & 1], k=K_TOP, devices=(first, count))
            done, t0 = 0, time.perf_counter()
            while time.perf_counter() - t0 < secs:
                for i in range(4):
                    model.predict_u8(bufs[i & 1], k=K_TOP, devices=(first, count))
                done += 4 * n
            dt = time.perf_counter() - t0
            pin.close()
            return done / dt

        # (a) all GPUs, one ResNet-18 job, asynchronous submit/wait with two requests in flight
        with b2.Model(ctx, "resnet18", ot_paths["resnet18"], precision=b2.PREC_BF16, max_batch=256) as m:
            n = 256 * n_gpus
            pin = b2.PinnedBuffer(2 * n * b2.IMAGE_BYTES)
            bufs = pin.array.reshape(2, n, 224, 224, 3)
            bufs[0] = np.resize(synth.synthetic_images(512, seed=3), bufs[0].shape)
            bufs[1] = np.roll(bufs[0], 1, axis=0)
            oi = [np.empty((n, K_TOP), np.int32) for _ in range(2)]
            op = [np.empty((n, K_TOP), np.float32) for _ in range(2)]
            for i in range(3):
                m.predict_u8(bufs[i & 1], k=K_TOP)
            prev, t0 = None, time.perf_counter()
            k = max(steps, 20)
            for i in range(k):
                t = m.submit_u8(bufs[i & 1], oi[i & 1], op[i & 1])
                if prev is not None:
                    m.wait(prev)
                prev = t
            m.wait(prev)
            dt = time.perf_counter() - t0
            out["e2e"] = {"value": k * n / dt, "unit": "images/s", "api": "b2i_submit_u8 / b2i_wait over %d GPUs in one process" % n_gpus,
                          "batch": n, "steps": k, "h2d_bytes_per_step": n * b2.IMAGE_BYTES}
            pin.close()
            # (b) dual job: ResNet-18 on the first half, AlexNet on the second, solo then together
            half = n_gpus // 2
            with b2.Model(ctx, "alexnet", ot_paths["alexnet"], precision=b2.PREC_BF16, max_batch=256) as a:
                solo_r = e2e_rate(m, 0, half, 256, seconds)
                solo_a = e2e_rate(a, half, n_gpus - half, 256, seconds)
                both = {}
                th = [threading.Thread(target=lambda: both.__setitem__("resnet18", e2e_rate(m, 0, half, 256, seconds))),
                      threading.Thread(target=lambda: both.__setitem__("alexnet", e2e_rate(a, half, n_gpus - half, 256, seconds)))]
                for t in th:
                    t.start()
                for t in th:
                    t.join()
                out["dual_job"] = {"split": "resnet18 on GPUs 0-%d, alexnet on GPUs %d-%d (b2i_predict_u8_on)" % (half - 1, half, n_gpus - 1),
                                   "resnet18": {"solo": solo_r, "together": both.get("resnet18"),
                                                "ratio": both.get("resnet18", 0) / solo_r if solo_r else None},
                                   "alexnet": {"solo": solo_a, "together": both.get("alexnet"),
                                               "ratio": both.get("alexnet", 0) / solo_a if solo_a else None},
                                   "unit": "images/s end to end, batch 256 per GPU, blocking calls"}
    finally:
        ctx.close()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--model", default="resnet18", choices=["resnet18", "alexnet"])
    ap.add_argument("--batch", type=int, default=256, help="images per GPU per step")
    ap.add_argument("--cpu-budget", type=float, default=12.0, help="seconds of oracle work for cpu_baseline")
    ap.add_argument("--sustained-seconds", type=float, default=5.0)
    ap.add_argument("--tile-n-max", type=int, default=0)
    ap.add_argument("--conv-mode", type=int, default=-1)
    ap.add_argument("--micro-batch", type=int, default=0)
    ap.add_argument("--profile-out", default="")
    ap.add_argument("--patch-conv", type=int, default=None)
    ap.add_argument("--pair-mma", type=int, default=None)
    ap.add_argument("--fuse-downsample", type=int, default=None)
    ap.add_argument("--no-extras", action="store_true", help="skip sustained / fp16 / inproc / parity (A/B timing runs)")
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

    MODEL, BATCH = args.model, args.batch
    torch.cuda.set_device(local)
    numa = bind_to_gpu_numa_node(local) if world > 1 else None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    cache = os.path.join(ROOT, "tests", "_cache")
    weights_kind = "pretrained" if synth.real_weights(MODEL) else "seeded synthetic"
    ot = None
    if rank == 0:
        ot = synth.real_weights(MODEL) or synth.ensure_ot(MODEL, os.path.join(cache, MODEL + ".ot"))
    if world > 1:
        dist.barrier()
        ot = synth.real_weights(MODEL) or os.path.join(cache, MODEL + ".ot")

    prec = b2.PREC_FP16 if args.precision == "fp16" else b2.PREC_BF16
    ctx = b2.Context((local,))
    model = b2.Model(ctx, MODEL, ot, precision=prec, max_batch=BATCH, micro_batch=args.micro_batch,
                     tile_n_max=args.tile_n_max, conv_mode=args.conv_mode, patch_conv=args.patch_conv, pair_mma=args.pair_mma,
                     fuse_downsample=args.fuse_downsample)
    n_micro = (BATCH + (args.micro_batch or BATCH) - 1) // (args.micro_batch or BATCH)
    config = make_config(MODEL, BATCH, args.micro_batch, weights_kind)
    ROTATE = max(2, -(-ROTATE_BYTES // (BATCH * b2.IMAGE_BYTES)))

    host = synth.synthetic_images(min(BATCH, 256), seed=rank)
    host = np.resize(host, (BATCH, 224, 224, 3))
    pin = b2.PinnedBuffer(ROTATE * BATCH * b2.IMAGE_BYTES)
    hostbufs = pin.array.reshape(ROTATE, BATCH, 224, 224, 3)
    for r in range(ROTATE):
        hostbufs[r] = np.roll(host, r, axis=0)
    dev_in = [torch.from_numpy(hostbufs[r]).cuda() for r in range(ROTATE)]
    d_idx = torch.empty((BATCH, K_TOP), dtype=torch.int32, device="cuda")
    d_prob = torch.empty((BATCH, K_TOP), dtype=torch.float32, device="cuda")
    stream = torch.cuda.current_stream()

    def step_device(i, mdl=None):
        (mdl or model).forward_device(dev_in[i % ROTATE].data_ptr(), BATCH, K_TOP, d_idx.data_ptr(), d_prob.data_ptr(),
                                      None, stream.cuda_stream)

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def max_over_ranks(vals):
        if world == 1:
            return list(vals)
        t = torch.tensor(list(vals), device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return [float(v) for v in t.tolist()]

    def timed_device(steps, mdl=None):
        for i in range(args.warmup):
            step_device(i, mdl)
        sync_all()
        evs = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
        evs[0].record(stream)
        for i in range(steps):
            step_device(i, mdl)
            evs[i + 1].record(stream)
        sync_all()
        return [evs[i].elapsed_time(evs[i + 1]) for i in range(steps)], evs[0].elapsed_time(evs[-1])

    # ---------------- device-resident timing (value)
    sampler = ClockSampler(local) if rank == 0 else None   # polls through the device-timed AND the e2e regions
    step_ms, total_ms = timed_device(args.steps)
    total_ms, = max_over_ranks([total_ms])
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
    # of batch i. Every step still copies its batch in from pinned host memory and its top-5 out.
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
    e2e_total, blocking_total = max_over_ranks([e2e_total, blocking_total])
    e2e_value = world * BATCH * args.steps / e2e_total
    clocks = sampler.stop() if sampler else None
    e2e_blocking_value = world * BATCH * args.steps / blocking_total

    # ---------------- copy-only ceiling of the host side: the same pinned buffers, plain asynchronous copies,
    # all ranks at once -- what the box can feed whatever runs on the GPUs
    h2d = None
    if not args.no_extras:
        src = [torch.from_numpy(hostbufs[r]) for r in range(ROTATE)]
        dst = torch.empty_like(dev_in[0])
        pinned_seen = bool(src[0].is_pinned())
        for r in range(3):
            dst.copy_(src[r % ROTATE], non_blocking=True)
        sync_all()
        reps = max(20, int(1.0 / (BATCH * b2.IMAGE_BYTES / 25e9)))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for r in range(reps):
            dst.copy_(src[r % ROTATE], non_blocking=True)
        e1.record(stream)
        sync_all()
        copy_s, = max_over_ranks([e0.elapsed_time(e1) / 1000.0])
        gbs = reps * BATCH * b2.IMAGE_BYTES / copy_s / 1e9
        h2d = {"gbs_per_gpu": gbs, "gbs_aggregate": gbs * world, "images_per_s_bound": world * reps * BATCH / copy_s,
               "bytes_per_copy": BATCH * b2.IMAGE_BYTES, "copies": reps, "source_is_pinned": pinned_seen,
               "how": "cudaMemcpyAsync of the e2e buffers alone, all %d rank(s) at once, max over ranks" % world}
        del dst

    # ---------------- sustained segment: >= 5 s of back-to-back device-resident steps, clock and power sampled
    sustained = None
    if not args.no_extras and args.sustained_seconds > 0:
        n_sus = max(args.steps, int(args.sustained_seconds * 1.05 / (total_ms / args.steps / 1000.0)))
        sync_all()
        s2 = ClockSampler(local) if rank == 0 else None
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for i in range(n_sus):
            step_device(i)          # asynchronous launches; the driver's queue depth bounds how far the host runs ahead
        e1.record(stream)
        sync_all()
        sus_s, = max_over_ranks([e0.elapsed_time(e1) / 1000.0])
        sc = s2.stop() if s2 else {}
        sustained = {"seconds": sus_s, "steps": n_sus, "images_per_s": world * BATCH * n_sus / sus_s,
                     "sm_mhz_median": sc.get("sm_mhz"), "sm_mhz_min": sc.get("sm_mhz_min"), "sm_max_mhz": sc.get("sm_max_mhz"),
                     "power_w_max": sc.get("power_w_max"), "reasons": sc.get("reasons"), "samples": sc.get("samples")}

    # ---------------- the IEEE-half mode, device-resident (same kernels, the mode that meets a flat 2e-2)
    other = None
    if not args.no_extras:
        oprec, oname = (b2.PREC_BF16, "bf16") if args.precision == "fp16" else (b2.PREC_FP16, "fp16")
        with b2.Model(ctx, MODEL, ot, precision=oprec, max_batch=BATCH, micro_batch=args.micro_batch) as m2:
            _, o_ms = timed_device(args.steps, m2)
        o_ms, = max_over_ranks([o_ms])
        other = {"precision": oname, "value": world * BATCH * args.steps / (o_ms / 1000.0), "unit": "images/s",
                 "ms_per_step": o_ms / args.steps}

    # ---------------- per-launch share (rank 0): CUDA events around every launch of one step
    prof = None
    if rank == 0:
        mb = args.micro_batch or BATCH
        acc = {}
        reps = 5
        for r in range(reps + 1):
            rows = model.profile(dev_in[r % ROTATE].data_ptr(), min(mb, BATCH), K_TOP, d_idx.data_ptr(), d_prob.data_ptr(),
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

    # ---------------- parity of the timed model on a few images (outside every timed region)
    parity = None
    if not args.no_extras:
        from oracle import tch_oracle
        w, _, _ = bench_weights(MODEL)
        imgs = synth.synthetic_images(8, seed=123)
        _, _, got = model.predict_u8(imgs, k=K_TOP, want_logits=True)
        _, _, ref = tch_oracle.predict_u8(MODEL, w, imgs, k=K_TOP)
        gap = float(np.abs(got - ref).max())
        parity = {"images": 8, "precision": args.precision, "against": "fp32 CPU restatement of the tch path (oracle/)",
                  "max_abs_logit_gap": gap, "logit_std": float(ref.std()), "gap_over_std": gap / float(ref.std()),
                  "top1_identical": bool(np.array_equal(np.argmax(got, 1), np.argmax(ref, 1))),
                  "full_record": "profiles/: parity_1000 (all 1000 reference images, every precision)"}

    pk = peaks()
    flops_img = model.flops_per_image
    mb = min(args.micro_batch or BATCH, BATCH)
    conv_ms = sum(ms for name, ms in prof if name.startswith(("conv:", "fc:")))
    conv_launches = sum(1 for name, _ in prof if name.startswith(("conv:", "fc:")))
    all_ms = sum(ms for _, ms in prof)
    achieved = flops_img * mb / (conv_ms / 1000.0) / 1e12 if conv_ms > 0 else 0.0
    # which measured peak matches this run: cuBLAS' sustained figure was taken power-capped near 1.3 GHz; if this
    # workload keeps the full SM clock through its own >= 5 s segment, the burst figure is the matching one
    full_clock = bool(sustained and sustained.get("sm_mhz_median") and sustained.get("sm_max_mhz") and
                      sustained["sm_mhz_median"] >= 0.97 * sustained["sm_max_mhz"] and
                      not set(sustained.get("reasons") or []) & {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"})
    use = "burst" if (full_clock or sustained is None) else "sustained"
    peak = pk["tflops_burst"] if use == "burst" else pk["tflops_sustained"]
    traffic, traffic_src = None, "not captured for this build; see profiles/ for the ncu tables"
    tp = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if os.path.exists(tp):
        tj = json.load(open(tp))
        if tj.get("model") == MODEL and tj.get("batch") == mb and tj.get("conv_fc_launches") == conv_launches:
            traffic, traffic_src = tj.get("dram_bytes_per_conv_fc_launch"), tj.get("source")
    whole_tflops = value / world * flops_img / 1e12
    sus_tflops = sustained["images_per_s"] / world * flops_img / 1e12 if sustained else None
    roofline = {
        "kernel": "tcgen05 implicit-GEMM family: all %d conv/FC launches of a step" % conv_launches,
        "bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak,
        "peak_used": use + (" (this workload holds %.0f MHz through its %.1f s segment)" %
                            (sustained["sm_mhz_median"], sustained["seconds"]) if sustained and sustained.get("sm_mhz_median") else ""),
        "frac_burst": achieved / pk["tflops_burst"], "frac_sustained": achieved / pk["tflops_sustained"],
        "peak_burst": pk["tflops_burst"], "peak_sustained": pk["tflops_sustained"], "peak_source": pk["source"],
        "traffic": traffic, "traffic_source": traffic_src,
        "flops_per_launch_avg": flops_img * mb / max(1, conv_launches),
        "ms_per_launch_avg": conv_ms / max(1, conv_launches),
        "share_of_step": conv_ms / all_ms if all_ms else None,
        "whole_step": {"tflops": whole_tflops, "frac_burst": whole_tflops / pk["tflops_burst"],
                       "frac_sustained": whole_tflops / pk["tflops_sustained"]},
        "whole_step_sustained_segment": None if sus_tflops is None else {
            "tflops": sus_tflops, "frac_burst": sus_tflops / pk["tflops_burst"], "frac_sustained": sus_tflops / pk["tflops_sustained"]},
    }
    cpu_baseline = None
    if world == 1:   # rank 0 at N = 1 only: with more ranks the other processes compete for the host cores
        cpu_val, cores, sample = cpu_oracle_rate(MODEL, args.cpu_budget)
        cpu_baseline = {"value": cpu_val, "unit": "images/s", "cores": cores, "kind": "port", "sample": sample}
        if not args.no_extras:
            try:
                cpu_baseline["reference_mode"] = cpu_reference_mode(MODEL)
            except Exception as ex:   # an extra record must not cost the bench line
                cpu_baseline["reference_mode"] = {"error": str(ex)}
    else:
        cpu_baseline = {"value": None, "unit": "images/s", "cores": None, "kind": "port",
                        "sample": "taken at N = 1 only (the --impl reference arm measures it at every N)"}
    gpu_launches = model.launches_per_microbatch * n_micro * args.steps
    from_files = None
    if world == 1 and not args.no_extras and MODEL == "resnet18":
        try:
            from_files = from_files_record(b2, ctx, ot, prec)
        except Exception as ex:   # an extra record must not cost the bench line
            from_files = {"error": str(ex)}
    inproc = None
    if world == 1 and not args.no_extras and MODEL == "resnet18":
        try:
            ng = b2.device_count()
            if ng >= 2:
                model.close()
                model = None
                ots = {"resnet18": ot, "alexnet": synth.real_weights("alexnet") or synth.ensure_ot("alexnet", os.path.join(cache, "alexnet.ot"))}
                inproc = inproc_record(b2, synth, ots, ng, args.steps)
        except Exception as ex:   # an extra record must not cost the bench line
            inproc = {"error": str(ex)}
    srt = sorted(e2e_ms)
    dsrt = sorted(step_ms)
    pct = lambda a, q: a[min(len(a) - 1, int(q * len(a)))]
    e2e = {"value": e2e_value, "unit": "images/s", "h2d_bytes_per_step": BATCH * b2.IMAGE_BYTES,
           "d2h_bytes_per_step": BATCH * K_TOP * 8,
           "api": "b2i_submit_u8 / b2i_wait, two requests in flight, pinned host buffers",
           "blocking": {"value": e2e_blocking_value, "unit": "images/s", "api": "b2i_predict_u8",
                        "batch_latency_ms": {"p50": pct(srt, 0.50), "p99": pct(srt, 0.99)}}}
    if h2d:
        e2e["h2d_ceiling"] = h2d
        e2e["h2d_ceiling_gbs"] = h2d["gbs_aggregate"]
        e2e["fraction_of_h2d_ceiling"] = e2e_value / h2d["images_per_s_bound"]
    out = {
        "metric": metric_name(MODEL, BATCH), "value": value, "unit": "images/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": total_ms / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": args.precision, "data": "synthetic",
        "config": config,
        "host_cpus_bound_to_gpu_numa_node": numa,
        "e2e": e2e,
        "device_batch_latency_ms": {"p50": pct(dsrt, 0.50), "p99": pct(dsrt, 0.99)},
        "gpu_launches": gpu_launches,
        "clocks": clocks,
        "sustained": sustained,
        "roofline": roofline,
        "cpu_baseline": cpu_baseline,
        "parity": parity,
        ("fp16" if args.precision == "bf16" else "bf16"): other,
        "inproc": inproc,
        "from_files": from_files,
        "launch_ms": {name: round(ms, 4) for name, ms in prof},
    }
    print(json.dumps(out))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
