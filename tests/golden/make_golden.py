"""Regenerates tests/golden/golden_vectors.npz from the oracle (run from the repo root):

    python tests/golden/make_golden.py

The reference cannot be executed here (Rust + tch, no toolchain, weights are LFS pointers), so
these vectors freeze the ORACLE's outputs on the committed fixture images (the subset
select_fixtures.py picks from the reference's test_files/imagenet_1k) with the seeded synthetic
weights: they pin the oracle, the synthetic-weight generator and the preprocessing restatement
against drift, and give the GPU tests a fixed target.  Parity with the reference itself stays
unpinned.
"""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle import stb_resize, synth, tch_oracle  # noqa: E402

FIXTURES = os.path.join(ROOT, "tests", "golden", "imagenet_1k", "train")


def fixture_paths():
    out = []
    for wnid in sorted(os.listdir(FIXTURES)):
        files = sorted(os.listdir(os.path.join(FIXTURES, wnid)))
        out.append((wnid, os.path.join(FIXTURES, wnid, files[0])))
    return out


def main():
    paths = fixture_paths()
    from concurrent.futures import ProcessPoolExecutor
    with ProcessPoolExecutor() as ex:
        crops = np.stack(list(ex.map(stb_resize.load_image_and_resize, [p for _, p in paths])))
    data = {
        "wnids": np.array([w for w, _ in paths]),
        "crop_sha256": np.array([hashlib.sha256(c.tobytes()).hexdigest() for c in crops]),
        # two full crops (an up-sampled and a down-sampled image) for a bit-exact pixel check
        "crop_n03825788": crops[[w for w, _ in paths].index("n03825788")],
        "crop_n02840245": crops[[w for w, _ in paths].index("n02840245")],
    }
    rw = synth.resnet18_weights()
    aw = synth.alexnet_weights()
    _, _, data["resnet18_logits"] = tch_oracle.predict_u8("resnet18", rw, crops)
    _, _, data["alexnet_tch_logits"] = tch_oracle.predict_u8("alexnet", aw, crops, alexnet_geometry="tch")
    _, _, data["alexnet_torchvision_logits"] = tch_oracle.predict_u8("alexnet", aw, crops,
                                                                    alexnet_geometry="torchvision")
    # the same ResNet-18 with logits of pretrained scale (std ~2.5): flat absolute tolerances are checked on it
    _, _, data["resnet18_ps_logits"] = tch_oracle.predict_u8("resnet18", synth.weights("resnet18_ps"), crops)
    syn = synth.synthetic_images(4, seed=0)
    _, _, data["resnet18_logits_synthetic4"] = tch_oracle.predict_u8("resnet18", rw, syn)
    data["synthetic4_sha256"] = np.array(hashlib.sha256(syn.tobytes()).hexdigest())
    data["resnet18_weight_checksum"] = np.array([float(sum(v.double().sum() for v in rw.values()))])
    data["alexnet_weight_checksum"] = np.array([float(sum(v.double().sum() for v in aw.values()))])
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "golden_vectors.npz"), **data)
    print("wrote golden_vectors.npz:", {k: getattr(v, "shape", None) for k, v in data.items()})


if __name__ == "__main__":
    main()
