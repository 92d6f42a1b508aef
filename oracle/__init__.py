"""CPU oracle for the inference hot path.  TEST INFRASTRUCTURE ONLY.

A restatement, on torch CPU fp32 / numpy, of what the reference computes through tch 0.4.1 under
`Member::predict` (reference src/services.rs:475-497).  Only `tests/`, `__graft_entry__.smoke()`
and `bench.py`'s cpu-baseline / `--impl reference` legs may import this package; the product
(`distributed-machine-learning-cluster_b200/`) never does.

PARITY UNPINNED: the reference has no test, golden vector or published accuracy for this path
(SURVEY.md 8c), tch's own source is not vendored in the reference and the pretrained weights are
Git-LFS pointers.  Every function below says which reference call site it follows and which parts
are recollection of tch 0.4.1 / stb_image_resize behaviour ([tch-recall]).
"""
