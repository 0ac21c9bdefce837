"""TEST INFRASTRUCTURE ONLY — CPU restatement of the reference's conv-BN-ReLU path.

Nothing under oracle/ is imported by the product package (deep-vision_b200/).  Only tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline / --impl reference legs may use it, and only as the checker or the timed CPU baseline.

Parity status: the reference ships no tests, golden vectors or known-answer fixtures for this path
and its arithmetic lives in an unpinned third-party dependency, `torch` (ResNet/pytorch/requirements.in:2; installed
here: 2.11.0+cu128).  The oracle is therefore pinned against OUTPUTS OF THE REFERENCE ITSELF: oracle/make_golden.py
imports the unmodified model files from /root/reference in the build container, runs them on seeded inputs and freezes
the results under tests/golden/; tests/test_oracle.py checks both halves of the oracle (ref_ops: numpy, ref_models:
torch-functional) against those fixtures.
"""
