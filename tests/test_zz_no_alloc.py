"""ia_decode_nms_static — the entry point for callers that cannot allocate device memory
(a process whose HBM belongs to another allocator, or a thread that must stay out of
cudaMalloc/cudaFree): it takes no workspace and keeps its candidate lists in a fixed array
inside the library.

Checked here, through the C-ABI: its outputs equal ia_decode_nms's (kept indices, counts and
status exactly; rows bit for bit, the tolerance every GPU test in this suite uses) on two
calls in a row with different inputs — the second call finds the first call's leftovers in
the shared array and must not be affected by them — and sizes beyond the fixed array are
refused.  The file sorts last so a `-x` run reaches it only after every other test.
"""

import ctypes

import numpy as np
import pytest
import torch

from oracle import pipeline as O
from workloads import synth

pytestmark = pytest.mark.gpu

CONF, IOU, MAX_DET = 0.25, 0.45, 1024
IA_ERR_CAPACITY = -4
STATIC_MAX_IMAGES, STATIC_MAX_CANDIDATES = 64, 1024


@pytest.fixture(scope="module")
def native():
    from shared.processing import _native
    _native.lib()
    return _native


def _outputs(n, dev):
    return (torch.full((n, MAX_DET, 6), -7.0, dtype=torch.float32, device=dev),
            torch.full((n, MAX_DET), -7, dtype=torch.int32, device=dev),
            torch.full((n,), -7, dtype=torch.int32, device=dev),
            torch.full((n,), -7, dtype=torch.int32, device=dev))


def _call_workspace(native, head, max_candidates):
    """The existing entry point, with a caller-allocated workspace."""
    n, anchors, row = head.shape
    lib = native.lib()
    dets, keep, count, status = _outputs(n, head.device)
    ws = torch.empty(lib.ia_decode_nms_workspace_bytes(n, anchors, max_candidates), dtype=torch.uint8,
                     device=head.device)
    rc = lib.ia_decode_nms(head.data_ptr(), native.IA_HEAD_LAYOUT_V5, n, anchors, row - 5, CONF, IOU,
                           max_candidates, MAX_DET, dets.data_ptr(), keep.data_ptr(), count.data_ptr(),
                           status.data_ptr(), ws.data_ptr(), ws.numel(),
                           torch.cuda.current_stream().cuda_stream)
    assert rc == 0, native.lib().ia_error_string(rc)
    torch.cuda.synchronize()
    return dets.cpu().numpy(), keep.cpu().numpy(), count.cpu().numpy(), status.cpu().numpy()


def _call_static(native, head, max_candidates, expect_rc=0):
    """The allocation-free entry point: inputs and outputs only, no workspace."""
    n, anchors, row = head.shape
    dets, keep, count, status = _outputs(n, head.device)
    rc = native.lib().ia_decode_nms_static(head.data_ptr(), native.IA_HEAD_LAYOUT_V5, n, anchors, row - 5,
                                           CONF, IOU, max_candidates, MAX_DET, dets.data_ptr(),
                                           keep.data_ptr(), count.data_ptr(), status.data_ptr(),
                                           torch.cuda.current_stream().cuda_stream)
    assert rc == expect_rc, (rc, native.lib().ia_error_string(rc))
    torch.cuda.synchronize()
    return dets.cpu().numpy(), keep.cpu().numpy(), count.cpu().numpy(), status.cpu().numpy()


def _assert_same(got, want):
    g_dets, g_keep, g_count, g_status = got
    w_dets, w_keep, w_count, w_status = want
    assert np.array_equal(g_count, w_count) and np.array_equal(g_status, w_status)
    for i, k in enumerate(w_count):
        assert np.array_equal(g_keep[i, :k], w_keep[i, :k])
        assert np.array_equal(g_dets[i, :k], w_dets[i, :k])  # bit-exact float32 rows


def _assert_matches_oracle(got, head_np):
    """The CPU oracle's kept indices and rows (oracle/pipeline.py::detect_v5, pinned to the
    reference's own apply_nms by tests/test_oracle_golden.py), image by image."""
    g_dets, g_keep, g_count, g_status = got
    for i in range(head_np.shape[0]):
        want_dets, want_keep = O.detect_v5(head_np[i], CONF, IOU)
        k = int(g_count[i])
        assert g_status[i] == 0 and k == len(want_keep)
        assert g_keep[i, :k].tolist() == list(want_keep)
        assert np.array_equal(g_dets[i, :k], want_dets)


def test_two_calls_in_a_row_match_the_workspace_entry_point(native):
    # call 1: three images, ~100 candidates each; call 2: two images with MORE candidates per
    # image and different objects, so stale keys/counters from call 1 would show up at once
    first_np = synth.head_v5(3, seed=91)
    second_np = synth.head_v5_dense(2, seed=92, candidates=600, anchors=5000)
    first, second = torch.from_numpy(first_np).cuda(), torch.from_numpy(second_np).cuda()
    want_first = _call_workspace(native, first, STATIC_MAX_CANDIDATES)
    want_second = _call_workspace(native, second, STATIC_MAX_CANDIDATES)
    assert want_first[2].min() >= 3 and want_second[2].min() > want_first[2].max()

    got_first = _call_static(native, first, STATIC_MAX_CANDIDATES)
    _assert_same(got_first, want_first)
    got_second = _call_static(native, second, STATIC_MAX_CANDIDATES)
    _assert_same(got_second, want_second)
    # and back again: fewer candidates than the leftovers now in the array
    _assert_same(_call_static(native, first, STATIC_MAX_CANDIDATES), want_first)
    # not only equal to the other entry point: equal to the CPU oracle of the reference's decode + NMS
    _assert_matches_oracle(got_first, first_np)
    _assert_matches_oracle(got_second, second_np)


def test_candidate_budget_overflow_is_reported_identically(native):
    head = torch.from_numpy(synth.head_v5_dense(1, seed=93, candidates=1500, anchors=3000)).cuda()
    want = _call_workspace(native, head, STATIC_MAX_CANDIDATES)
    got = _call_static(native, head, STATIC_MAX_CANDIDATES)
    assert want[3][0] & 1 and np.array_equal(got[3], want[3])  # bit0: more than 1024 passed


def test_sizes_beyond_the_fixed_array_are_refused(native):
    lib = native.lib()
    small = torch.from_numpy(synth.head_v5(1, seed=94, anchors=2000, dup=10)).cuda()
    # too many candidates per image (explicit, and implicit via "use every anchor")
    out = _call_static(native, small, STATIC_MAX_CANDIDATES + 1, expect_rc=IA_ERR_CAPACITY)
    assert (out[2] == -7).all()  # nothing was launched: outputs untouched
    _call_static(native, small, 0, expect_rc=IA_ERR_CAPACITY)
    # too many images in one call
    many = torch.zeros((STATIC_MAX_IMAGES + 1, 64, 85), dtype=torch.float32, device="cuda")
    _call_static(native, many, 64, expect_rc=IA_ERR_CAPACITY)
    assert b"capacity" in lib.ia_error_string(IA_ERR_CAPACITY)
    # the largest accepted shape works: 64 images, "every anchor" with anchors <= 1024
    full = torch.from_numpy(synth.head_v5_dense(STATIC_MAX_IMAGES, seed=95, candidates=50, anchors=1000)).cuda()
    _assert_same(_call_static(native, full, 0), _call_workspace(native, full, 0))
