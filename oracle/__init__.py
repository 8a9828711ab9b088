"""Oracle = TEST INFRASTRUCTURE ONLY.

Everything under ``oracle/`` is a CPU restatement of the reference's algorithm
(or a loader for the unmodified reference itself) used to CHECK the CUDA path.
Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline /
``--impl reference`` legs may import it.  The product package
(``lit_gpt_b200``) never imports ``oracle`` and has no CPU fallback.

Parity status:
  * transformer math (RMSNorm, RoPE, KV-cache attention, SwiGLU MLP, generate,
    tensor-parallel sharding): PINNED — ``oracle/model_ref.py`` is checked
    against golden vectors produced by the unmodified reference
    (``tests/golden/make_golden.py`` imports /root/reference behind a stub).
  * nf4 / fp4 / nf4-dq / int8 quantisation arithmetic: PARITY UNPINNED — the
    arithmetic lives in bitsandbytes==0.41.0 (requirements-all.txt:3), which
    is not vendored in /root/reference and not installable here; the
    restatement follows the published algorithm of that version.
  * int4 group-128: this project's own format (the reference has no gptq/int4
    code at all); PARITY UNPINNED by construction.
"""
