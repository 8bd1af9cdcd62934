"""CPU oracle for the ImVoxelNet image->voxel path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is product code: only
``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline /
``--impl reference`` legs may import it, and only as the checker or as the
timed CPU baseline -- never as the thing shipped.  The product
(``imvoxelnet_b200``) never imports this package and raises if its CUDA
library is missing.

Parity status: PINNED.  The restatements here are checked, bit for bit on
pixel indices / valid masks and to 1e-6 on volumes, against the outputs of the
reference's own functions (``/root/reference/mmdet3d/models/detectors/
imvoxelnet.py`` and ``.../necks/imvoxelnet.py``) imported verbatim in the
build container by ``oracle/ref_loader.py``; those outputs are committed under
``tests/golden/`` together with ``tests/golden/make_golden.py`` which made
them.  The reference's own test-suite holds no golden vectors for this path
(SURVEY.md section 4), so the fixtures generated from the reference itself
are the pin.
"""
