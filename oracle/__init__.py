"""CPU oracle for the rtseg hot path.  TEST INFRASTRUCTURE ONLY.

Nothing under oracle/ is imported by the product package (realtime-semantic-segmentation-pytorch_b200/).
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import it,
and only as the checker or the timed CPU baseline.

What it is: a plain-PyTorch fp32 (CPU) restatement of the reference's arithmetic on this path, written
functionally over a state dict (so it shares no code with the product's nn.Modules), plus a numpy
restatement of the torchmetrics confusion matrix the reference delegates to (torchmetrics is a
third-party dependency of the reference — requirements.txt:3, unpinned, `task=` keyword => >= 0.11 —
and is not installed in this image).

Parity pinning: the reference ships no tests or golden vectors.  The oracle is pinned
against outputs of the reference itself, imported in the build container from /root/reference by
tests/golden/make_golden.py; the vectors are committed under tests/golden/ and checked by
tests/test_oracle_golden.py on every run (and live against /root/reference when it exists).
torchmetrics could not be run, so the confusion-matrix restatement is pinned only to its published
semantics ("parity unpinned" for that one function; the int64 matrix itself is cross-checked against
an independent pure-Python loop).
"""
