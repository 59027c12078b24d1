"""CPU oracle for the MOSS-TTS synthesis hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under `oracle/` is imported by the product
package `moss_tts_b200`; only `tests/`, `__graft_entry__.smoke()` and the
`cpu_baseline` / `--impl reference` legs of `bench.py` use it, and only as the
checker or as the timed CPU baseline.

Each function restates, in plain torch-CPU / numpy, the algorithm of the cited
reference file:line.  Pinning: the reference ships no golden vectors of its own
(SURVEY.md §4), so the restatement is pinned against outputs of the reference
itself executed in the build container (`oracle/make_golden.py`, fixtures under
`tests/golden/`, checked by `tests/test_oracle_golden.py`).
"""
