"""CPU oracle for the instance-segmentation post-processing path.

TEST INFRASTRUCTURE ONLY.  Nothing in the product package
(`instancesegmentation-_b200/`, imported as `isb`) may import this package.
Allowed importers: `tests/`, `__graft_entry__.smoke()`, and the
`cpu_baseline` / `--impl reference` legs of `bench.py`.

The modules here are numpy restatements of the reference's torch code
(`/root/reference/inst_model/...`); every function cites the reference
file:line it follows.  Parity pin: the reference itself ships no tests or
golden vectors, so the pin is made by running the real
reference functions in the build container on seeded synthetic inputs
(`oracle/gen_golden.py`) and committing their outputs under `tests/golden/`;
`tests/test_oracle_golden.py` checks these restatements against those
fixtures.
"""
