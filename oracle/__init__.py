"""CPU oracle for the CIDEr-D reward path.  TEST INFRASTRUCTURE ONLY.

Nothing in the product package (``cst_captioning_b200``) imports this package.
Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline /
``--impl reference`` legs may use it, and only as the checker or the timed CPU
baseline, never as the thing shipped.
"""
