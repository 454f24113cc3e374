"""CPU oracle for the attention + KV-compression hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under compactor-vllm_b200/ imports this
package; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
`--impl reference` legs may.  It is a plain restatement (torch fp32 on CPU for
the floating-point parts, numpy / python ints for index work) of the reference
algorithms, each function citing the reference file:line it follows (paths
relative to /root/reference/src/compactor_vllm/).

Parity pin: the reference ships no golden vectors.  The oracle is pinned
against the reference's OWN Triton kernels executed on CPU under
TRITON_INTERPRET=1 (oracle/gen_golden.py, run in the build container where
/root/reference exists); the resulting vectors are committed under
tests/golden/ and checked by tests/test_oracle_golden.py.  The one function
the interpreter cannot run (approximate_leverage_scores: `.cuda()` +
cuSOLVER gesvda, compression/compactor.py:157,177) is pinned against the
reference's torch code path re-executed on CPU with torch.linalg.svd
substituted for the gesvda driver — see gen_golden.py; that substitution means
leverage parity is "pinned to the reference formula, not to cuSOLVER's bits".
"""
from . import allocator, attention, prepath, scoring, selection, store  # noqa: F401
