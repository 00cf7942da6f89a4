"""CPU oracle for the FiT denoising hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``fit_b200/`` may import this package;
only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` do, and there only as the checker or
as the timed CPU baseline -- never as the product path.

The oracle is a plain fp32 torch / float64 numpy restatement of the reference
algorithm (jlnk03/FiT), function by function, each citing the reference
file:line it follows.  Parity is pinned: ``oracle/make_goldens.py`` imports the
unmodified reference from ``/root/reference`` (CPU, with a stub ``timm``) and
writes ``tests/golden/*.npz``; ``tests/test_oracle_golden.py`` checks every
oracle function against those fixtures.
"""
from . import fit_oracle, diffusion_oracle, pos_embed_oracle  # noqa: F401
