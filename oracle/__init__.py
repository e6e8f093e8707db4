"""CPU oracle for the tmat segmentation hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is part of the shipped
product: it may be imported only by ``tests/``, ``__graft_entry__.smoke()`` and
the ``cpu_baseline`` / ``--impl reference`` legs of ``bench.py``, and there only
as the checker or as the timed CPU arm -- never as the thing that produces a
result for a user.  ``tmat_b200`` never imports it.

Each module restates one stage of the reference path in numpy / scipy / cv2 /
torch-CPU and cites the reference ``file:line`` it follows (paths relative to
``fl_tissue_model_tools/`` of fogg-lab/tissue-model-analysis-tools @ d1c494b).

Pinning status (see DESIGN.md "Oracle"):

* ``stp``, ``zstacks``, ``thresholds``     -- PINNED: checked bit-for-bit against the
  reference's own functions imported in the build container; the outputs are
  frozen in ``tests/golden/*.npz`` by ``tests/golden/make_golden.py``.
* ``morphology.filter_branch_seg_mask``   -- composition PINNED against the
  reference's ``transforms.py`` executed with scikit-image primitives replaced by
  this oracle's primitives; the primitives themselves (median, label,
  perimeter, Zhang-Suen thinning) restate scikit-image 0.22, which is not
  installable here: **parity unpinned** at that third-party boundary.
* ``wellmask``                            -- restates scikit-image calls with
  scipy.ndimage: **parity unpinned** at the scikit-image boundary.
* ``unet``                                -- restates Keras 2.14 layer semantics in
  torch-CPU fp32; TensorFlow is not installable here: **parity unpinned**.
"""
