"""oracle/ — CPU restatement of the reference's hot path.  TEST INFRASTRUCTURE ONLY.

Nothing in the product package (convnets_b200/) imports this.  It is used by tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs as the checker
(and as the timed CPU baseline), never as the thing shipped.

The reference (milan-pavlovic-ai/ConvNets) is pure Python over PyTorch: its arithmetic lives in
the third-party dependency torch (requirements.txt:20 pins torch==1.6.0+cu101; this image has
torch 2.11.0+cu128).  The oracle therefore has two layers:

* oracle/ops.py  — per-op semantics: the torch CPU op the reference calls (fp64 where a
  noise-free answer is wanted) plus independent numpy restatements of the index-exact ops
  (max-pool argmax, channel shuffle) that pin torch's tie-breaking rules.
* oracle/zoo.py  — the twelve model families restated functionally over a reference
  state_dict (same keys as the reference's nn.Module tree), each citing the reference file:line.

Pinning: the reference has no tests of its own.  The oracle is pinned against
the reference itself, imported in the build container from /root/reference by
oracle/make_golden.py, which writes small fixtures to tests/golden/ (parameter counts from
data/results.xlsx, state_dict key lists, logits / loss / gradient norms of one training step on
seeded inputs).  tests/test_oracle_golden.py checks the oracle against those fixtures on CPU.
"""
