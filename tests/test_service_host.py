"""Host logic mirrored from the reference (labels, Job accounting, the resource-manager split, the
batched run_job and the `jobs` report), driven with a stand-in member so no GPU is needed."""
import os

import pytest

from conftest import GOLDEN


class FakeMember:
    """Answers like MemberState.predict_each: one entry per id, (prob, label) with the true label for ids
    whose wnid ends in an even digit, or a marker for the ids listed in `skipped` / `undecodable` /
    `missing`."""

    def __init__(self, labels, skipped=(), undecodable=(), missing=()):
        self.by_wnid = dict(labels)
        self.calls = []
        self.marks = {}
        import importlib
        svc = importlib.import_module("distributed-machine-learning-cluster_b200.service")
        for ids, mark in ((skipped, svc.SKIPPED), (undecodable, svc.UNDECODABLE), (missing, svc.MISSING)):
            for w in ids:
                self.marks[w] = mark

    def read_ids(self, input_ids):
        return [self.marks.get(w, b"jpeg bytes of " + w.encode()) for w in input_ids]

    def predict_each(self, model_name, input_ids, k=1, devices=None, blobs=None):
        self.calls.append((model_name, list(input_ids), devices))
        assert blobs is not None and len(blobs) == len(input_ids)      # run_job fetched the files itself
        if model_name not in ("resnet18", "alexnet"):
            return None
        return [b if isinstance(b, str) else (0.5, self.by_wnid[w] if int(w[-1]) % 2 == 0 else "wrong")
                for w, b in zip(input_ids, blobs)]


def test_load_labels_follows_reference_parsing(pkg):
    labels = pkg.load_labels(os.path.join(GOLDEN, "synset_words.txt"))
    assert len(labels) == 1000
    assert labels[0] == ("n01440764", "tench, Tinca tinca")           # services.rs:176-180: rest of line, single spaces
    assert labels[999][0] == "n15075141"
    assert all(" " not in w and lbl for w, lbl in labels)


def test_job_accounting(pkg):
    job = pkg.Job("resnet18")
    job.add_query_result(True, 0.010)
    job.add_query_result(False, 0.030)
    assert (job.finished_prediction_count, job.correct_prediction_count) == (2, 1)
    assert job.query_durations == [0.010, 0.030]


def test_assign_members_split(pkg):
    a, b = pkg.Job("resnet18"), pkg.Job("alexnet")
    pkg.assign_members([3, 1, 0, 2, 7, 6, 5, 4], a, b)                # sorted, first half / second half (services.rs:205-209)
    assert a.assigned_member_ids == [0, 1, 2, 3] and b.assigned_member_ids == [4, 5, 6, 7]
    pkg.assign_members([0], a, b)                                     # n = 1: resnet18 gets nobody (services.rs:207)
    assert a.assigned_member_ids == [] and b.assigned_member_ids == [0]
    pkg.assign_members([0, 1, 2], a, b)
    assert a.assigned_member_ids == [0] and b.assigned_member_ids == [1, 2]


def test_run_job_batches_resumes_and_reports(pkg):
    labels = pkg.load_labels(os.path.join(GOLDEN, "synset_words.txt"))[:50]
    member = FakeMember(labels)
    leader = pkg.Leader(member, labels)
    pkg.assign_members([0, 1, 2, 3], leader.resnet18, leader.alexnet)
    lines = []
    leader.run_job(leader.resnet18, batch=16, limit=20, log=lines.append, in_flight=1)
    assert [len(c[1]) for c in member.calls] == [16, 4]               # many ids per call
    assert member.calls[0][2] == (0, 2)                               # device range of the job's share
    assert leader.resnet18.finished_prediction_count == 20 and len(lines) == 20
    assert any("(should be " in ln for ln in lines) and any("(should be " not in ln for ln in lines)
    leader.run_job(leader.resnet18, batch=16)                         # resume at finished_prediction_count (services.rs:410-411)
    assert sorted(c[1][0] for c in member.calls[2:]) == sorted([labels[20][0], labels[36][0]])   # two batches in flight
    assert leader.resnet18.finished_prediction_count == 50
    rep = pkg.jobs_report(leader.resnet18)
    want_correct = sum(1 for w, _ in labels if int(w[-1]) % 2 == 0)
    assert rep["finished"] == 50 and rep["correct"] == want_correct
    assert abs(rep["accuracy"] - 100.0 * want_correct / 50) < 1e-9
    assert rep["median_ms"] <= rep["p90_ms"] <= rep["p95_ms"] <= rep["p99_ms"]
    empty = pkg.jobs_report(pkg.Job("alexnet"))
    assert empty["total"] == 0


def test_run_job_skips_when_nobody_is_assigned(pkg):
    labels = pkg.load_labels(os.path.join(GOLDEN, "synset_words.txt"))[:8]
    member = FakeMember(labels)
    leader = pkg.Leader(member, labels)
    leader.run_job(leader.alexnet, batch=4)                           # services.rs:416: `continue`
    assert member.calls == [] and leader.alexnet.finished_prediction_count == 0


def test_leader_predict_runs_both_jobs(pkg):
    labels = pkg.load_labels(os.path.join(GOLDEN, "synset_words.txt"))[:12]
    member = FakeMember(labels)
    leader = pkg.Leader(member, labels)
    pkg.assign_members(range(8), leader.resnet18, leader.alexnet)
    leader.predict(batch=5)
    assert leader.resnet18.finished_prediction_count == 12 and leader.alexnet.finished_prediction_count == 12
    assert {c[2] for c in member.calls} == {(0, 4), (4, 4)}           # disjoint GPU halves


def test_decode_failure_and_unknown_model(pkg, tmp_path):
    bad = tmp_path / "x.JPEG"
    bad.write_bytes(b"junk")
    assert pkg.decode_image(str(bad)) is None
    with pytest.raises(ValueError):
        pkg.load_model("vgg16")


def test_run_job_scores_every_answer_against_its_own_label(pkg):
    """An empty directory, an undecodable file and a missing directory in the middle of a batch: each is a
    lost query, as in the reference (one id per RPC, services.rs:421,485-492); the ids after them are still
    compared with their own labels and the rest of the batch is counted."""
    labels = pkg.load_labels(os.path.join(GOLDEN, "synset_words.txt"))[:32]
    lost = [labels[3][0], labels[10][0], labels[17][0]]
    member = FakeMember(labels, skipped=[lost[0]], undecodable=[lost[1]], missing=[lost[2]])
    leader = pkg.Leader(member, labels)
    pkg.assign_members([0, 1], leader.resnet18, leader.alexnet)
    lines = []
    leader.run_job(leader.alexnet, batch=32, log=lines.append)
    assert leader.alexnet.finished_prediction_count == 29
    answered = [(w, l) for w, l in labels if w not in lost]
    assert leader.alexnet.correct_prediction_count == sum(1 for w, _ in answered if int(w[-1]) % 2 == 0)
    # every printed line names the id it belongs to and, when wrong, that id's own label
    by = dict(labels)
    for ln in lines:
        wnid = ln.split(" - ")[1].split(":")[0]
        assert wnid not in lost
        if "(should be " in ln:
            assert ln.endswith("(should be %s)" % by[wnid])


def test_member_file_side_markers(pkg, tmp_path):
    """read_ids: first file of the id's directory as bytes; markers for an empty and a missing directory."""
    import importlib
    svc = importlib.import_module("distributed-machine-learning-cluster_b200.service")
    (tmp_path / "a").mkdir()
    (tmp_path / "a" / "2.JPEG").write_bytes(b"second")
    (tmp_path / "a" / "1.JPEG").write_bytes(b"first")
    (tmp_path / "b").mkdir()
    m = svc.MemberState.__new__(svc.MemberState)          # file side only: no GPU, no models
    m.image_root = str(tmp_path)
    from concurrent.futures import ThreadPoolExecutor
    m.pool = ThreadPoolExecutor(max_workers=2)
    assert m.read_ids(["a", "b", "c"]) == [b"first", svc.SKIPPED, svc.MISSING]
    assert svc.decode_bytes(b"junk") is None and svc.decode_bytes(b"") is None
    img = svc.decode_bytes(open(os.path.join(GOLDEN, "imagenet_1k", "train", sorted(os.listdir(os.path.join(GOLDEN, "imagenet_1k", "train")))[0],
                                             sorted(os.listdir(os.path.join(GOLDEN, "imagenet_1k", "train", sorted(os.listdir(os.path.join(GOLDEN, "imagenet_1k", "train")))[0])))[0]), "rb").read())
    assert img.ndim == 3 and img.shape[2] == 3
    m.pool.shutdown()
