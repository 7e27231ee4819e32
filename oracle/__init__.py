"""CPU oracle for the `shared.processing` hot path of matthewhoung/inference-arena.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is part of the product:
only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline /
``--impl reference`` legs may import it, and there only as the checker or as
the timed CPU arm.  The product path (``inference-arena_b200/``) never imports
it and has no CPU fallback.

Parity pinning: the reference holds no golden vectors for resized pixels, NMS
indices or decoded boxes, so the oracle is pinned against
outputs of the reference itself: ``tests/golden/make_golden.py`` imports the
unmodified reference from ``/root/reference`` in the build container, runs it
on seeded inputs and commits the results under ``tests/golden/``;
``tests/test_oracle_golden.py`` checks every oracle function against them.

Modules
    cv_resize   numpy restatement of OpenCV's 8-bit INTER_LINEAR resize
                (third-party arithmetic the reference calls at
                transforms.py:160-164 and mobilenet_preprocess.py:139-143)
    pipeline    numpy restatement of the reference functions a1..a9
    synth       shim for workloads/synth.py (seeded synthetic inputs C1..C5)
"""
