"""CPU oracle for the MEANT encoder hot path.  TEST INFRASTRUCTURE ONLY.

This package is a plain-PyTorch, fp32, CPU restatement of the reference algorithm
(biirving/meant, top-level tree; `src/` tree for the TimeSformer path).  Each function
cites the reference file:line it follows.

Who may import it: `tests/`, `__graft_entry__.smoke()` and the `cpu_baseline` /
`--impl reference` legs of `bench.py` -- as the checker or the timed CPU baseline,
never as the product.  Nothing under `meant_b200/` imports it; the product path
raises when the CUDA library is missing instead of falling back to this code.

Parity pinning: the reference has no tests, golden vectors or fixtures,
so the oracle is pinned against *outputs of the reference itself run in the build
container*: `oracle/gen_golden.py` imports the unmodified reference modules from
/root/reference, feeds them hashed (RNG-free) weights and inputs and commits the
outputs / gradient digests under `tests/golden/`; `tests/test_oracle_golden.py`
checks every oracle function against those files.  Third-party arithmetic that
cannot run here (`flash-attn==2.5.2`, CUDA only) is restated from its published
definition and is marked "parity unpinned" in `flash_variants` below and in DESIGN.md.
"""
