"""oracle/ - TEST INFRASTRUCTURE ONLY.

CPU restatement (numpy) of the reference's feature hot path plus the harness that
executes the reference's own files under stand-in modules.  Only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg may import
anything from here; the product package (audio-analysis_b200/) never does.

Parity status: the reference ships NO tests, golden vectors or fixtures
(SURVEY.md section 4), so the oracle is pinned against outputs of the reference's own
files executed in the build container (tests/golden/make_golden.py, committed with the
fixtures it made).  The librosa functions underneath are a restatement of librosa 0.11
semantics (librosa is not installable here): that boundary is "parity unpinned".
"""
