"""CPU oracle for the active-learning acquisition pass.

TEST INFRASTRUCTURE ONLY.  Nothing in the product package
(`medical-image-analysis_b200/`) imports this directory; only `tests/`,
`__graft_entry__.smoke()` and `bench.py`'s CPU-baseline legs may.  The product
path fails loudly when its CUDA library is missing instead of routing here.

Contents
--------
ref_import   import the UNMODIFIED reference (/root/reference/src) in this
             container with unused third-party modules stubbed.  The reference
             cannot travel to the GPU box; it is used to pin the restatement
             and to generate tests/golden/* (script: oracle/make_golden.py).
unet_ref     torch-fp32 functional restatement of the reference UNet forward
             (src/models/unet/unet.py, blocks.py) with the same state_dict
             key layout; seeded weight builder.
select_ref   restatement of the selectors' score / rank / k-center logic
             (src/activelearning/*.py) and of the scikit-learn distance
             formulas the reference calls (scikit-learn 1.9.0).
synth        seeded synthetic FUGC-/BUSI-shaped pools and mirror dataset types.

Parity pinning: the reference has no tests, golden vectors or known-answer
fixtures of its own.  The restatement is pinned against outputs
of the reference itself run here (tests/test_oracle_vs_reference.py when
/root/reference exists; frozen in tests/golden/ by oracle/make_golden.py).
MC-dropout and hybrid acquisition do not exist in the reference; their oracle
is a composition of reference pieces -> "parity unpinned by the reference" for
those two paths.
"""
