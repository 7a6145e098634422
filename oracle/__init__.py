"""CPU oracle for the optical-flow hot path -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference`` legs of
``bench.py`` may import anything from this package, and there only as the checker or as the timed CPU
baseline.  The product path (``cardioflow_b200``) never imports it and has no CPU fallback.

What is restated here (reference = /root/reference, nicolas1805961/Cardiac-Segmentation-Optical-flow):

* ``warp_oracle``   -- ``SpatialTransformer`` / ``SpatialTransformerContour`` / ``VecInt`` of
                       ``nnunet/network_architecture/integration.py:5-99`` (torch CPU fp32).
* ``corr_oracle``   -- ``CorrBlock`` / ``coords_grid``.  The reference imports them from
                       ``nnunet/lib/raft_initial.py`` (``SegFlowGaussian.py:42``) but never committed that
                       file, so the algorithm lives in a third-party dependency: the public RAFT
                       ``core/corr.py`` + ``core/utils/utils.py`` whose names/signatures the call sites use
                       (``SegFlowGaussian.py:929,935,839``), and torchvision's
                       ``models/optical_flow/raft.py:337-431`` which the reference calls through
                       ``raft_large`` (``nnMTLTrainerV2Raft.py:97-118``; torchvision version not pinned by
                       the reference's ``setup.py``, 0.26.0 in this image).
* ``taps``          -- numpy restatement of the exact fp32 coordinate sequence (normalise ->
                       un-normalise -> floor -> weights) used to check tap indices / weights bit-exactly,
                       plus a grid_sample-free lookup and warp built from those taps.
* ``cf_oracle.c``   -- the same algorithms in plain C (gcc, -ffp-contract=off), built by ``oracle/Makefile``
                       into ``oracle/_build/libcf_oracle.so`` and wrapped by ``c_oracle.py``.

Pinning status: the reference ships no tests / golden vectors for this path.
``SpatialTransformer`` is pinned by executable source: ``oracle/make_golden.py`` imports the reference's
``integration.py`` by file path and freezes its outputs under ``tests/golden/``.  ``CorrBlock`` is pinned at the
reference's in-tree third-party call site (torchvision ``CorrBlock``), also frozen under ``tests/golden/``;
the reference's *own* ``CorrBlock`` file is absent, so for that symbol parity is "unpinned by the
reference's own source" and anchored on its call-site contract.
"""
