"""N4 parity: softmax + top-k / arg-max of classifier logits on the GPU vs the oracle and the
vectors the reference's own classify() produced.  Class indices: exact.  Probabilities:
1e-5 relative (measured ~1e-7 — CUDA's expf and numpy's differ in the last bit)."""

import numpy as np
import pytest
import torch

from oracle import pipeline as O
from workloads import synth

pytestmark = pytest.mark.gpu

RTOL = 1e-5


@pytest.fixture(scope="module")
def D():
    from shared.processing import device
    return device


def _run(D, logits, k, **kw):
    ids, vals, status = D.classify_topk_cuda(torch.from_numpy(logits).cuda(), k, **kw)
    torch.cuda.synchronize()
    return ids.cpu().numpy(), vals.cpu().numpy(), status.cpu().numpy()


@pytest.mark.parametrize("name,seed,classes,k,rows", [("cls", 400, 1000, 5, 16), ("cls_small", 401, 10, 3, 4),
                                                      ("cls_k1", 400, 1000, 1, 4)])
def test_matches_reference_vectors(D, golden, name, seed, classes, k, rows):
    logits = synth.logits(16 if name != "cls_small" else 4, seed=seed, classes=classes)[:rows]
    ids, vals, status = _run(D, logits, k)
    assert np.array_equal(ids, golden[f"{name}_ids"]) and not status.any()
    np.testing.assert_allclose(vals.astype(np.float64), golden[f"{name}_probs"], rtol=RTOL, atol=0)


@pytest.mark.parametrize("rows,classes,k", [(1, 1000, 5), (1035, 1000, 5), (7, 1, 1), (5, 33, 33), (3, 1001, 7),
                                            (2, 21843, 5), (2, 40000, 3), (300, 100, 10), (4, 1024, 5),
                                            (4, 1025, 5), (3, 64, 40), (9, 31, 6)])
def test_softmax_topk_vs_oracle(D, rows, classes, k):
    logits = synth.logits(rows, seed=500 + classes, classes=classes)
    ids, vals, status = _run(D, logits, k)
    for r in range(rows):
        want_ids, want_probs = O.softmax_topk(logits[r], k)
        top = min(k, 5, classes)  # the generator separates the first five
        assert ids[r, :top].tolist() == want_ids[:top].tolist()
        np.testing.assert_allclose(vals[r, :top], want_probs[:top], rtol=RTOL, atol=0)
        # beyond the separated five: same multiset of probabilities, order up to rounding
        np.testing.assert_allclose(vals[r], want_probs, rtol=RTOL, atol=1e-30)
        assert len(set(ids[r].tolist())) == k
    assert not status.any()


@pytest.mark.parametrize("rows,classes,k", [(12000, 1000, 5), (6500, 1024, 5), (7000, 896, 3), (9000, 128, 5),
                                            (8000, 4, 2), (6001, 900, 32)])
def test_many_rows(D, rows, classes, k):
    """More rows than the GPU holds warps at once (5 920 on a B200), several waves of CTAs.  Every row
    against numpy; rows with NaN, infinities, nothing but -inf, ties and constants are mixed in —
    those leave the fast path inside the same launch."""
    logits = synth.logits(rows, seed=700 + classes, classes=classes)
    special = {}
    if classes >= 128:
        logits[17, 3] = np.nan
        logits[rows - 1, 7] = np.inf
        logits[4000, :] = -np.inf
        logits[4001, 5] = logits[4001, 77] = 60.0
        logits[5999, :] = 1.5
        special = {17, rows - 1, 4000, 4001, 5999}
    ids, vals, status = _run(D, logits, k)
    with np.errstate(invalid="ignore"):
        e = np.exp(logits - logits.max(axis=1, keepdims=True))
        probs = e / e.sum(axis=1, keepdims=True)
    want_ids = np.argsort(probs, axis=1, kind="stable")[:, ::-1][:, :k]
    top = min(k, 5, classes)  # the generator separates the first five
    plain = np.array([r not in special for r in range(rows)])
    assert np.array_equal(ids[plain][:, :top], want_ids[plain][:, :top])
    want_vals = np.take_along_axis(probs, want_ids, axis=1)
    np.testing.assert_allclose(vals[plain], want_vals[plain], rtol=RTOL, atol=1e-30)
    assert all(len(set(row.tolist())) == k for row in ids[::97])
    for r in special:
        o_ids, o_probs = O.softmax_topk(logits[r], k)
        if np.isnan(o_probs).all():
            assert np.isnan(vals[r]).all() and ids[r].tolist() == o_ids.tolist(), r
        else:
            np.testing.assert_allclose(vals[r], o_probs, rtol=RTOL, atol=1e-30)
            assert ids[r, :2].tolist() == o_ids[:2].tolist(), r
    assert set(np.nonzero(status)[0].tolist()) == ({17} if special else set())


def test_unaligned_rows_take_the_scalar_loads(D):
    logits = synth.logits(9, seed=640)
    flat = torch.empty(logits.size + 1, dtype=torch.float32, device="cuda")
    view = flat[1:].view(logits.shape)
    view.copy_(torch.from_numpy(logits))
    ids, vals, _ = D.classify_topk_cuda(view, 5)
    ids2, vals2, _ = D.classify_topk_cuda(torch.from_numpy(logits).cuda(), 5)
    # the two load patterns sum the row in different orders: last-bit differences allowed
    assert torch.equal(ids, ids2) and torch.allclose(vals, vals2, rtol=1e-6, atol=0)
    for r in range(9):
        assert ids[r].tolist() == O.softmax_topk(logits[r], 5)[0].tolist()


def test_argmax_raw_matches_np_argmax(D):
    logits = synth.logits(64, seed=600)
    logits[3, 10] = logits[3, 700] = 50.0  # tie: np.argmax takes the first
    logits[4, 5] = np.nan  # np.argmax returns the first NaN
    logits[4, 900] = np.nan
    logits[5, :] = -0.0
    logits[5, 17] = 0.0  # -0.0 == +0.0: still the first index
    ids, vals, status = _run(D, logits, 1, softmax=False, low_index_first=True)
    for r in range(64):
        cid, conf = O.top1(logits[r][None])
        assert ids[r, 0] == cid
        assert vals[r, 0] == conf or (np.isnan(conf) and np.isnan(vals[r, 0]))
    assert status[4] == 1 and status.sum() == 1


def test_ties_and_special_values(D):
    x = synth.logits(6, seed=601, classes=50)
    x[0, 7] = x[0, 30] = 40.0  # equal probabilities: larger index first (stable sort reversed)
    x[1, 11] = np.nan  # every probability NaN
    x[2, 3] = np.inf  # inf - inf -> NaN in the sum -> every probability NaN
    x[3, :] = -np.inf
    x[4, :] = 3.0  # uniform
    ids, vals, status = _run(D, x, 4)
    assert ids[0, :2].tolist() == [30, 7] and vals[0, 0] == vals[0, 1]
    for r in (1, 2, 3):
        assert np.isnan(vals[r]).all() and ids[r].tolist() == [49, 48, 47, 46]
    assert status.tolist() == [0, 1, 0, 0, 0, 0]
    assert ids[4].tolist() == [49, 48, 47, 46]
    np.testing.assert_allclose(vals[4], np.float32(1.0) / np.float32(50.0), rtol=1e-6)
    want_ids, want_probs = O.softmax_topk(x[5], 4)
    assert ids[5].tolist() == want_ids.tolist()


def test_large_magnitudes_are_stable(D):
    x = synth.logits(4, seed=602) * np.float32(50.0)  # exp() of the raw values would overflow
    ids, vals, _ = _run(D, x, 5)
    for r in range(4):
        want_ids, want_probs = O.softmax_topk(x[r], 5)
        assert ids[r].tolist() == want_ids.tolist()
        np.testing.assert_allclose(vals[r], want_probs, rtol=RTOL, atol=1e-38)


def test_argument_checks(D):
    x = torch.zeros((2, 10), device="cuda")
    with pytest.raises(RuntimeError, match="invalid"):
        D.classify_topk_cuda(x, 11)
    with pytest.raises(RuntimeError):
        D.classify_topk_cuda(torch.zeros((1, 60000), device="cuda"), 1)
    ids, vals, status = D.classify_topk_cuda(torch.zeros((0, 10), device="cuda"), 3)
    assert ids.shape == (0, 3)
    with pytest.raises(RuntimeError, match="CUDA tensor"):
        D.classify_topk_cuda(torch.zeros((2, 10)), 1)


def test_numpy_wrappers(golden):
    from shared.processing.classify import softmax_top_k, top1
    row = synth.logits(16, seed=400)[2]
    got = softmax_top_k(row, 5)
    assert [g[0] for g in got] == golden["cls_ids"][2].tolist()
    np.testing.assert_allclose([g[1] for g in got], golden["cls_probs"][2], rtol=RTOL)
    assert top1(row[None]) == O.top1(row[None])
    assert len(softmax_top_k(row[:3], 5)) == 3
    with pytest.raises(ValueError):
        softmax_top_k(np.zeros((2, 3)))
