"""CPU oracle for the OnePose++ 2D-3D / LoFTR 2D-2D matching hot path.

TEST INFRASTRUCTURE ONLY.  Nothing in ``onepose_st_b200`` imports this package.
Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline /
``--impl reference`` legs may import it, and there only as the checker or as
the timed CPU baseline -- never as the shipped compute path.

What it is: a plain torch-CPU *functional* restatement of the reference's
algorithm (mizeller/OnePose_ST @ 320dac4), one function per reference
operator, each citing the reference file:line it follows.  It takes a flat
``state_dict`` whose key names are the reference's own parameter names, so the
same weights drive the reference modules, this oracle and the CUDA product.

Parity pinning: the reference has no tests / golden vectors of its own
(SURVEY.md section 4).  The oracle is therefore pinned against outputs of the
reference's own modules, imported unmodified from /root/reference in the build
container by ``tests/golden/make_golden.py`` and committed as fixtures under
``tests/golden/*.pt`` (``tests/test_oracle_golden.py`` checks them on CPU).
The one third-party dependency on the path, kornia==0.4.1 (``dsnt.
spatial_expectation2d`` / ``create_meshgrid``), is absent from the image and is
restated from its published semantics in ``ref_ops.create_meshgrid`` /
``ref_ops.spatial_expectation2d``; that boundary is pinned by closed-form
tests only ("parity unpinned" for kornia itself).
"""
from . import ref_ops  # noqa: F401
from .ref_ops import *  # noqa: F401,F403
