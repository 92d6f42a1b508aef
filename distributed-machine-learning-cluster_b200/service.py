"""Host-side mirror of the reference's operator interface for the inference path, over the C ABI.

Same names, argument meaning and error behaviour as the Rust it stands in for (the Rust toolchain
is absent here, see INTEGRATION.md for the cgo-style binding a maintainer would add):

  load_labels          reference src/services.rs:170-184   (synset_words.txt -> [(wnid, label)])
  sample_predict       reference src/model_utils.rs:7-22
  load_model           reference src/model_utils.rs:24-36  (fails loudly instead of swallowing errors)
  predict_image        reference src/model_utils.rs:38-49
  MemberState.predict  reference src/services.rs:475-497   (Member::predict RPC handler body)
  MemberState.train    reference src/main.rs:231-256       (member side: reload the copied .ot)
  Job                  reference src/services.rs:54-81
  Leader.assign_members reference src/services.rs:199-211  (dual-job fair share; members = GPUs)
  Leader.run_job       reference src/services.rs:407-433   (batched: many ids per predict call)
  jobs_report          reference src/main.rs:281-309       (ms percentiles + accuracy)

The ABI takes bytes, not paths.  Decode is off the calling thread:
MemberState reads the files and hands the encoded streams to b2i_predict_jpeg (nvJPEG on the GPU); where
nvJPEG is not installed, or for the few streams it refuses, a host thread pool decodes with PIL (which
releases the GIL) and the pixels go through b2i_predict_raw_u8_status.
"""
import math
import os
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np

from . import _lib
from .infer import ALEXNET_TCH, PREC_BF16, Context, Model

IMAGE_ROOT = os.path.join("test_files", "imagenet_1k", "train")       # services.rs:476
WEIGHT_DIR = "pretrained_models"                                       # services.rs:516,522
MODEL_NAMES = ("resnet18", "alexnet")


def load_labels(path="synset_words.txt"):
    """[(wnid, label)] in class-index order; label = remaining words joined by single spaces."""
    labels = []
    with open(path) as f:
        for line in f:
            parts = line.split()
            if parts:
                labels.append((parts[0], " ".join(parts[1:])))
    return labels


def decode_image(path):
    """u8 [H,W,3] RGB; None when the file cannot be decoded (tch's `.ok()?`, services.rs:492)."""
    try:
        from PIL import Image
        with Image.open(path) as im:
            return np.asarray(im.convert("RGB"))
    except Exception:
        return None


def decode_bytes(blob):
    """Host decode of one encoded image held in memory; None when it cannot be decoded."""
    if not blob:
        return None
    try:
        import io
        from PIL import Image
        with Image.open(io.BytesIO(blob)) as im:
            return np.asarray(im.convert("RGB"))
    except Exception:
        return None


class MissingDirectory(Exception):
    """`read_dir(folder/id)` failed: the reference unwraps and the handler panics (services.rs:485)."""


SKIPPED = "skipped"      # the id's directory is empty: the reference `continue`s (services.rs:490)
UNDECODABLE = "undecodable"   # load_image_and_resize failed: the reference's RPC returns None (services.rs:492)
MISSING = "missing"      # no such directory: the reference's handler panics (services.rs:485)


def sample_predict(labels, ctx=None, weight_dir=WEIGHT_DIR,
                   image_path=os.path.join(IMAGE_ROOT, "n01491361", "ILSVRC2012_00034413.JPEG")):
    """The reference's demo (model_utils.rs:7-22): AlexNet top-1 label of one fixed image."""
    ctx = ctx or Context((0,))
    img = decode_image(image_path)
    if img is None:
        raise IOError("cannot decode %s" % image_path)           # `?` on the TchError in the reference
    with Model(ctx, "alexnet", os.path.join(weight_dir, "alexnet.ot"), max_batch=1) as model:
        idx, _ = model.predict_u8(ctx.resize_crop(img)[None], k=1)
    return labels[int(idx[0, 0])][1]


def load_model(model_name, ctx=None, weight_dir=WEIGHT_DIR, **kw):
    """`load_model(name)`: build the named network and load `<weight_dir>/<name>.ot`.  Unknown names
    raise (the reference panics, model_utils.rs:30); unlike the reference's dead-code twin a missing
    or invalid archive is an error, not silently random weights."""
    if model_name not in MODEL_NAMES:
        raise ValueError("Model not found: %r" % (model_name,))
    ctx = ctx or Context((0,))
    return Model(ctx, model_name, os.path.join(weight_dir, model_name + ".ot"), **kw)


def predict_image(model, image_path, label, labels, ctx=None):
    """Some(top1_label == label) / None on decode failure (model_utils.rs:38-49)."""
    img = decode_image(image_path)
    if img is None:
        return None
    x = (ctx or model.ctx).resize_crop(img)[None]
    idx, _ = model.predict_u8(x, k=1)
    return labels[int(idx[0, 0])][1] == label


class MemberState:
    """The model half of the reference's `MemberState` (services.rs:450-457, 513-524): one ResNet-18
    and one AlexNet resident on this member's GPUs, and the body of `Member::predict`.

    decoder: "nvjpeg" = encoded streams go to the GPU decoder (b2i_predict_jpeg); "pil" = host thread
    pool; "auto" (default) = nvjpeg when the library can be loaded on this box, else pil."""

    def __init__(self, ctx, labels, weight_dir=WEIGHT_DIR, image_root=IMAGE_ROOT, precision=PREC_BF16,
                 max_batch=256, alexnet_geometry=ALEXNET_TCH, models=MODEL_NAMES, decoder="auto",
                 decode_threads=None):
        self.ctx = ctx
        self.labels = labels
        self.image_root = image_root
        self.weight_dir = weight_dir
        # services.rs:516,522 `.unwrap()`: a missing archive is fatal at start-up
        self.models = {
            name: Model(ctx, name, os.path.join(weight_dir, name + ".ot"), precision=precision,
                        max_batch=max_batch, alexnet_geometry=alexnet_geometry)
            for name in models
        }
        if decoder == "auto":
            try:
                ctx.jpeg_backend()
                decoder = "nvjpeg"
            except _lib.B2IError:
                decoder = "pil"
        if decoder not in ("nvjpeg", "pil"):
            raise ValueError("decoder must be auto, nvjpeg or pil")
        self.decoder = decoder
        self.pool = ThreadPoolExecutor(max_workers=decode_threads or min(32, os.cpu_count() or 1))

    def close(self):
        self.pool.shutdown(wait=True)
        for m in self.models.values():
            m.close()

    def train(self, model_name, weight_dir=None):
        """Member side of the `train <sdfs_file> <model>` command (main.rs:231-256): the leader has
        just copied a new `<model>.ot` to every member; re-read it and swap the weights in on every
        GPU.  The reference stops at the copy (its models keep the weights they started with)."""
        model = self.models.get(model_name)
        if model is None:
            raise ValueError("Model not found: %r" % (model_name,))
        model.reload(os.path.join(weight_dir or self.weight_dir, model_name + ".ot"))

    # ---- file side ---------------------------------------------------------------------------
    def _read_one(self, input_id):
        """First file of test_files/.../<id>/ (services.rs:485-490) as bytes; MISSING / SKIPPED markers."""
        folder = os.path.join(self.image_root, input_id)
        try:
            entries = sorted(os.listdir(folder))
        except OSError:
            return MISSING
        if not entries:
            return SKIPPED
        try:
            with open(os.path.join(folder, entries[0]), "rb") as f:
                return f.read()
        except OSError:
            return b""

    def read_ids(self, input_ids):
        """Encoded file contents for every id, read by the thread pool, aligned with input_ids."""
        return list(self.pool.map(self._read_one, input_ids))

    # ---- the handler -------------------------------------------------------------------------
    def predict_each(self, model_name, input_ids, k=1, devices=None, blobs=None):
        """One entry per input id, in order: (prob, label) -- or a list of k such pairs when k > 1 -- for
        an id that was answered, else one of the markers SKIPPED / UNDECODABLE / MISSING.  This is what a
        batched caller needs to keep the reference's per-query accounting: there, one id travels per RPC
        (services.rs:421), so a bad file costs its own query and nothing else.  None for an unknown
        model (services.rs:480).  `blobs` = the result of read_ids(input_ids) when the caller has
        already fetched the files (run_job reads the next batch while this one is on the GPU)."""
        model = self.models.get(model_name)
        if model is None:
            return None
        if blobs is None:
            blobs = self.read_ids(input_ids)
        out = list(blobs)
        live = [i for i, b in enumerate(blobs) if not isinstance(b, str)]
        if not live:
            return out
        idx = prob = None
        status = np.full(len(live), _lib.EDECODE, np.int32)
        if self.decoder == "nvjpeg":
            idx, prob, status = model.predict_jpeg([blobs[i] for i in live], k=k, devices=devices)
        retry = [j for j in range(len(live)) if status[j] != _lib.OK]
        if retry:
            # host decode for what the GPU decoder did not take (all of it when nvJPEG is absent)
            pixels = list(self.pool.map(decode_bytes, [blobs[live[j]] for j in retry]))
            if any(p is not None for p in pixels):
                ridx, rprob, rstatus = model.predict_raw_status(pixels, k=k, devices=devices)
                if idx is None:
                    idx = np.full((len(live), k), -1, np.int32)
                    prob = np.zeros((len(live), k), np.float32)
                for j, jj in enumerate(retry):
                    idx[jj], prob[jj], status[jj] = ridx[j], rprob[j], rstatus[j]
        for j, i in enumerate(live):
            if status[j] != _lib.OK:
                out[i] = UNDECODABLE
            elif k == 1:
                out[i] = (float(prob[j, 0]), self.labels[int(idx[j, 0])][1])
            else:
                out[i] = [(float(pp), self.labels[int(ii)][1]) for ii, pp in zip(idx[j], prob[j])]
        return out

    def predict(self, model_name, input_ids, k=1, devices=None):
        """Option<Vec<(f64, String)>> exactly as the reference's handler returns it: None for an unknown
        model (services.rs:480) or when any image fails to decode (services.rs:492); ids whose directory
        is empty are skipped (services.rs:490); a missing directory raises (the reference unwraps,
        services.rs:485).  With k > 1 each entry is a list of k (prob, label) pairs instead of one pair."""
        each = self.predict_each(model_name, input_ids, k=k, devices=devices)
        if each is None:
            return None
        if any(e == MISSING for e in each):
            bad = input_ids[[e == MISSING for e in each].index(True)]
            raise FileNotFoundError(os.path.join(self.image_root, bad))
        if any(e == UNDECODABLE for e in each):
            return None
        return [e for e in each if e != SKIPPED]


class Job:
    """services.rs:54-81."""

    def __init__(self, model_name):
        self.model_name = model_name
        self.finished_prediction_count = 0
        self.correct_prediction_count = 0
        self.query_durations = []          # seconds, one per query (image)
        self.assigned_member_ids = []

    def add_query_result(self, correct, duration):
        self.finished_prediction_count += 1
        if correct:
            self.correct_prediction_count += 1
        self.query_durations.append(duration)


def assign_members(active_ids, job_resnet18, job_alexnet):
    """Resource manager tick (services.rs:199-211): sorted ids, first floor(n/2) to ResNet-18, the
    rest to AlexNet.  Here an id is a GPU index of the member's context."""
    ids = sorted(set(active_ids))
    job_resnet18.assigned_member_ids = ids[:len(ids) // 2]
    job_alexnet.assigned_member_ids = ids[len(ids) // 2:]


class Leader:
    """The job half of `LeaderState` (services.rs:83-95, 146-155, 407-433) driving one MemberState."""

    def __init__(self, member, labels):
        self.member = member
        self.labels = labels
        self.resnet18 = Job("resnet18")
        self.alexnet = Job("alexnet")

    def jobs(self):
        return [self.resnet18, self.alexnet]

    def run_job(self, job, batch=32, limit=None, log=None, in_flight=2):
        """Batched `run_job`: resumes at labels[finished_prediction_count..] (services.rs:410-411),
        sends `batch` ids per predict call to the job's assigned GPUs, skips the labels silently when
        nobody is assigned (services.rs:416), records one duration per query = the wall time of the
        batch that carried it, and prints the reference's result lines (services.rs:425-429).  Every
        answer is scored against ITS OWN label; an id that produced no answer (empty or missing
        directory, undecodable file) is a lost query, as in the reference, and costs nothing else.
        `in_flight` batches are in progress at once (the reference spawns one task per query and never
        waits for the previous one, services.rs:418): while one batch is on the GPU the next one's files
        are read and decoded.  Results are accounted in label order."""
        start_at = job.finished_prediction_count
        todo = self.labels[start_at:] if limit is None else self.labels[start_at:start_at + limit]
        chunks = [todo[i:i + batch] for i in range(0, len(todo), batch)]

        def one(chunk):
            ids = list(job.assigned_member_ids)
            if not ids:
                return None, 0.0
            devices = (min(ids), max(ids) - min(ids) + 1)
            t0 = time.perf_counter()
            wn = [wnid for wnid, _ in chunk]
            res = self.member.predict_each(job.model_name, wn, devices=devices, blobs=self.member.read_ids(wn))
            return res, time.perf_counter() - t0

        workers = ThreadPoolExecutor(max_workers=max(1, in_flight))
        try:
            pending = []
            for ci in range(len(chunks) + in_flight):
                if ci < len(chunks):
                    pending.append((chunks[ci], workers.submit(one, chunks[ci])))
                if ci >= in_flight - 1 and pending:
                    chunk, fut = pending.pop(0)
                    res, dt = fut.result()
                    if res is None:
                        continue
                    for (wnid, true_label), r in zip(chunk, res):
                        if isinstance(r, str):          # SKIPPED / UNDECODABLE / MISSING: no answer for this id
                            continue
                        prob, pred = r
                        job.add_query_result(pred == true_label, dt)
                        if log:
                            if pred == true_label:
                                log("%s - %s: %s (%.2f%%)" % (job.model_name, wnid, pred, prob * 100.0))
                            else:
                                log("%s - %s: %s (%.2f%%) (should be %s)" % (job.model_name, wnid, pred, prob * 100.0, true_label))
        finally:
            workers.shutdown(wait=True)

    def predict(self, batch=32, limit=None, log=None):
        """`Leader::predict` (services.rs:146-151) joins both jobs; they run on disjoint GPU halves."""
        import threading
        th = [threading.Thread(target=self.run_job, args=(j, batch, limit, log)) for j in self.jobs()]
        for t in th:
            t.start()
        for t in th:
            t.join()


def jobs_report(job):
    """Numbers the CLI `jobs` command prints (main.rs:281-309): accuracy and query-latency stats in
    integer milliseconds.  Percentiles are exact order statistics (the reference's bucketed
    histogram is an approximation of these)."""
    ms = sorted(int(d * 1000.0) for d in job.query_durations)
    n = len(ms)
    if n == 0:
        return {"model": job.model_name, "finished": 0, "correct": 0, "accuracy": 0.0, "total": 0}
    mean = sum(ms) / n
    std = math.sqrt(sum((v - mean) ** 2 for v in ms) / n)
    pct = lambda q: ms[min(n - 1, max(0, int(math.ceil(q / 100.0 * n)) - 1))]
    return {
        "model": job.model_name,
        "finished": job.finished_prediction_count,
        "correct": job.correct_prediction_count,
        "accuracy": 100.0 * job.correct_prediction_count / max(1, job.finished_prediction_count),
        "total": n, "avg_ms": mean, "std_ms": std, "median_ms": pct(50), "p90_ms": pct(90), "p95_ms": pct(95),
        "p99_ms": pct(99),
    }
