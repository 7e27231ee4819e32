"""The whole chain through DevicePipeline (device-resident) and HostPipeline (pinned host
buffers in and out, chunked over streams, asynchronous submits) vs the oracle: the same
five calls the reference makes per request (monolithic/app/inference.py:161-193)."""

import numpy as np
import pytest
import torch

from oracle import pipeline as O
from workloads import synth

pytestmark = pytest.mark.gpu


def _oracle(frames, heads):
    return [O.run_image(f, h, 0.25, 0.45) for f, h in zip(frames, heads)]


def _check_device(out, want):
    offs = out.roi_offsets.cpu().numpy()
    for i, (tensor, dets, keep, _, crops) in enumerate(want):
        k = int(out.count[i])
        assert int(out.status[i]) == 0 and k == len(keep)
        assert out.keep[i, :k].cpu().tolist() == keep
        assert np.array_equal(out.tensor[i : i + 1].cpu().numpy(), tensor)
        assert np.array_equal(out.dets[i, :k].cpu().numpy(), dets)
        assert np.array_equal(out.crops[offs[i] : offs[i + 1]].cpu().numpy(), crops)


@pytest.mark.parametrize("h,w", [(480, 640), (1080, 1920)])
def test_device_pipeline(h, w):
    from shared.processing.pipeline import DevicePipeline

    n = 5
    frames, heads = synth.frames(n, h, w, seed=60), synth.head_v5(n, seed=61)
    pipe = DevicePipeline(n, h, w, "cuda")
    out = pipe.run(torch.from_numpy(frames).cuda(), torch.from_numpy(heads).cuda())
    torch.cuda.synchronize()
    _check_device(out, _oracle(frames, heads))


def test_device_pipeline_full_c5_batch():
    """BASELINE config C5 at the bench's own size: 256 x 1080p frames + (256, 25200, 85) heads in one
    pass; every image's status / count / ROI offsets checked, 32 sampled images compared bit for
    bit with the oracle (tensor, kept indices, rows, crops)."""
    from shared.processing.pipeline import DevicePipeline

    n, h, w = 256, 1080, 1920
    heads = synth.head_v5(n, seed=71)
    frames = synth.frames(n, h, w, seed=70)
    pipe = DevicePipeline(n, h, w, "cuda", max_det=32, max_candidates=4096, rois_per_image=8)
    out = pipe.run(torch.from_numpy(frames).cuda(), torch.from_numpy(heads).cuda())
    torch.cuda.synchronize()
    count = out.count.cpu().numpy()
    offs = out.roi_offsets.cpu().numpy()
    assert int(out.status.max()) == 0 and count.min() >= 1
    assert np.array_equal(offs, np.concatenate([[0], np.cumsum(count)]))
    rng = np.random.default_rng(9)
    for i in sorted(rng.choice(n, size=32, replace=False).tolist()):
        tensor, dets, keep, _, crops = O.run_image(frames[i], heads[i], 0.25, 0.45)
        k = int(count[i])
        assert k == len(keep) and out.keep[i, :k].cpu().tolist() == keep, i
        assert np.array_equal(out.dets[i, :k].cpu().numpy(), dets), i
        assert np.array_equal(out.tensor[i : i + 1].cpu().numpy(), tensor), i
        assert np.array_equal(out.crops[offs[i] : offs[i + 1]].cpu().numpy(), crops), i


def _check_host(res, want, chunk, rpi):
    for i, (tensor, dets, keep, _, crops) in enumerate(want):
        k = int(res["count"][i])
        assert int(res["status"][i]) == 0 and k == len(keep)
        assert res["keep"][i, :k].tolist() == keep
        assert np.array_equal(res["tensor"][i : i + 1].numpy(), tensor)
        assert np.array_equal(res["dets"][i, :k].numpy(), dets)
        ci, li = divmod(i, chunk)
        offs = res["roi_offsets"][ci].numpy()
        base = ci * chunk * rpi
        got = res["crops"][base + offs[li] : base + offs[li + 1]].numpy()
        assert np.array_equal(got, crops)


def test_host_pipeline_sync_and_async():
    from shared.processing.pipeline import HostPipeline

    h, w, n, chunk = 480, 640, 7, 3  # 3 chunks, the last one ragged
    hp = HostPipeline(n, h, w, "cuda", chunk=chunk, slots=2)
    batches = []
    for seed in (70, 72, 74):
        frames, heads = synth.frames(n, h, w, seed=seed), synth.head_v5(n, seed=seed + 1)
        batches.append((torch.from_numpy(frames).pin_memory(), torch.from_numpy(heads).pin_memory(),
                        _oracle(frames, heads)))
    # synchronous
    res = hp.run(batches[0][0], batches[0][1])
    _check_host(res, batches[0][2], chunk, hp.rpi)
    # asynchronous: two batches in flight, result sets rotate
    t1 = hp.submit(batches[1][0], batches[1][1])
    t2 = hp.submit(batches[2][0], batches[2][1])
    _check_host(t1.wait(), batches[1][2], chunk, hp.rpi)
    t3 = hp.submit(batches[0][0], batches[0][1])  # re-uses t1's result set (already waited)
    _check_host(t2.wait(), batches[2][2], chunk, hp.rpi)
    _check_host(t3.wait(), batches[0][2], chunk, hp.rpi)
    hp.drain()
    assert hp.d2h_bytes(n) > 0 and hp.h2d_bytes_per_image == h * w * 3 + 25200 * 85 * 4


def test_host_pipeline_from_jpeg_files():
    """N3 in the chain: the frames arrive as JPEG files, are decoded on the device into the
    chunk buffers and run through K1..K4; the oracle sees the same files through the oracle
    decoder (== cv2.imdecode, tests/test_oracle_jpeg.py)."""
    from oracle import jpeg as OJ
    from shared.processing.pipeline import HostPipeline
    from workloads import jpeg_cases as W

    h, w, n, chunk = 240, 320, 7, 3
    files = [W.encode(W.photo(h, w, seed=80 + i), 85, "420", restart=(i % 2) * 4) for i in range(n)]
    frames = np.stack([OJ.load_image_from_bytes(f) for f in files])
    heads = synth.head_v5(n, seed=81)
    hp = HostPipeline(n, h, w, "cuda", chunk=chunk, slots=2)
    want = _oracle(frames, heads)
    h_heads = torch.from_numpy(heads).pin_memory()
    _check_host(hp.run(files, h_heads), want, chunk, hp.rpi)
    t1 = hp.submit(files, h_heads)
    t2 = hp.submit(torch.from_numpy(frames).pin_memory(), h_heads)  # raw frames through the same pipeline
    _check_host(t1.wait(), want, chunk, hp.rpi)
    _check_host(t2.wait(), want, chunk, hp.rpi)
    assert hp.last_h2d_bytes == frames.nbytes + heads.nbytes
    with pytest.raises(ValueError):
        hp.run([W.encode(W.photo(100, 100, seed=1), 85)] * n, h_heads)  # wrong frame size
    hp.drain()
    # both models on the GPU: the head is already a device tensor, the model inputs stay there
    res = hp.run(files, torch.from_numpy(heads).cuda(), tensors_to_host=False)
    for i, (_, dets, keep, _, _) in enumerate(want):
        k = int(res["count"][i])
        assert int(res["status"][i]) == 0 and res["keep"][i, :k].tolist() == keep
        assert np.array_equal(res["dets"][i, :k].numpy(), dets)
    assert hp.last_h2d_bytes == sum(len(f) for f in files)
    hp.drain()


def test_graphed_pipeline_matches_eager():
    """The chain captured in a CUDA graph replays to the same bits, for changing inputs."""
    from shared.processing.pipeline import DevicePipeline, GraphedPipeline

    n, h, w = 2, 480, 640
    gp = GraphedPipeline(DevicePipeline(n, h, w, "cuda"), n)
    for seed in (90, 92):
        frames, heads = synth.frames(n, h, w, seed=seed), synth.head_v5(n, seed=seed + 1)
        out = gp.run(torch.from_numpy(frames).cuda(), torch.from_numpy(heads).cuda())
        torch.cuda.synchronize()
        _check_device(out, _oracle(frames, heads))


def test_device_pipeline_overlapped_streams():
    from shared.processing.pipeline import DevicePipeline

    n, h, w = 4, 480, 640
    frames, heads = synth.frames(n, h, w, seed=94), synth.head_v5(n, seed=95)
    pipe = DevicePipeline(n, h, w, "cuda")
    for _ in range(2):
        out = pipe.run(torch.from_numpy(frames).cuda(), torch.from_numpy(heads).cuda(), overlap=True)
    torch.cuda.synchronize()
    _check_device(out, _oracle(frames, heads))


def test_ragged_pipeline_from_mixed_jpeg_files():
    """A service-like batch: JPEG files of different sizes and samplings -> device decode ->
    K1 (one launch for all shapes) -> K2+K3 -> K4, against the oracle applied per request."""
    pytest.importorskip("cv2")
    from shared.processing.pipeline import RaggedPipeline
    from workloads import jpeg_cases as W

    shapes = [(480, 640), (427, 640), (375, 500), (640, 480), (333, 500), (480, 640), (612, 612), (240, 320)]
    files = [W.encode(W.photo(h, w, seed=90 + i), 88, W.SAMPLINGS[i % 3]) for i, (h, w) in enumerate(shapes)]
    # cv2.imdecode is the call load_image_from_bytes makes; oracle/jpeg.py is pinned equal to it
    # (tests/test_oracle_jpeg.py) but is a pure-Python loop, too slow for frames of this size
    frames = [W.reference_decode(f) for f in files]
    heads = synth.head_v5(len(files), seed=91)
    pipe = RaggedPipeline("cuda")
    for given in (files, [torch.from_numpy(f).cuda() for f in frames]):
        out = pipe.run(given, torch.from_numpy(heads).cuda())
        torch.cuda.synchronize()
        offs = out.roi_offsets.cpu().numpy()
        for i, (frame, head) in enumerate(zip(frames, heads)):
            tensor, dets, keep, _, crops = O.run_image(frame, head, 0.25, 0.45)
            k = int(out.count[i])
            assert int(out.status[i]) == 0 and out.keep[i, :k].cpu().tolist() == keep
            assert np.array_equal(out.tensor[i : i + 1].cpu().numpy(), tensor), shapes[i]
            assert np.array_equal(out.dets[i, :k].cpu().numpy(), dets)
            assert np.array_equal(out.crops[offs[i] : offs[i + 1]].cpu().numpy(), crops), shapes[i]


def test_pipelines_report_the_decode_status_of_each_file():
    """A damaged upload must not go unnoticed inside a batch: both pipelines hand back the JPEG
    decode status per request (0 ok, bit 0 damaged and concealed like libjpeg, bit 1 the data
    simply ends — the request the reference answers with "Failed to decode image from bytes")."""
    pytest.importorskip("cv2")
    from shared.processing.pipeline import HostPipeline, RaggedPipeline
    from workloads import jpeg_cases as W

    h, w = 240, 320
    files = [W.encode(W.photo(h, w, seed=40 + i), 85, "420") for i in range(6)]
    files[1] = files[1][: len(files[1]) // 2]                # cut off: no marker behind the data
    files[4] = files[4][: len(files[4]) // 2] + b"\xff\xd9"  # cut off and closed
    want = [0, 3, 0, 0, 1, 0]
    heads = torch.from_numpy(synth.head_v5(len(files), seed=41))
    out = RaggedPipeline("cuda").run(files, heads.cuda())
    torch.cuda.synchronize()
    assert out.jpeg_status.cpu().tolist() == want
    assert RaggedPipeline("cuda").run([torch.zeros(8, 8, 3, dtype=torch.uint8, device="cuda")],
                                      heads[:1].cuda()).jpeg_status is None
    host = HostPipeline(8, h, w, chunk=4)
    res = host.run(files, heads.pin_memory())
    assert res["jpeg_status"].tolist() == want
    frame = W.reference_decode(files[4])  # the concealed frame is libjpeg's, so is everything after it
    tensor, dets, keep, _, crops = O.run_image(frame, heads[4].numpy(), 0.25, 0.45)
    assert np.array_equal(res["tensor"][4:5].numpy(), tensor)
    raw = host.run(torch.zeros(2, h, w, 3, dtype=torch.uint8).pin_memory(), heads[:2].pin_memory())
    assert raw["jpeg_status"] is None


def _check_shards(shards, bounds, want, chunk):
    for res, (lo, hi) in zip(shards, bounds):
        if hi == lo:
            assert res is None
            continue
        for li, i in enumerate(range(lo, hi)):
            tensor, dets, keep, _, crops = want[i]
            k = int(res["count"][li])
            assert int(res["status"][li]) == 0 and k == len(keep)
            assert res["keep"][li, :k].tolist() == keep
            assert np.array_equal(res["tensor"][li : li + 1].numpy(), tensor)
            assert np.array_equal(res["dets"][li, :k].numpy(), dets)
            s, c = int(res["crop_start"][li]), int(res["crop_count"][li])
            assert c == k and np.array_equal(res["crops"][s : s + c].numpy(), crops)


@pytest.mark.parametrize("n_devices", [1, 2])
def test_sharded_host_pipeline(n_devices):
    """One host batch over several GPUs from one process: contiguous shards, one
    HostPipeline and one launching thread per GPU, no collective.  Every image of every shard
    equals the oracle; the two-GPU case needs a box with two GPUs."""
    from shared.processing.pipeline import ShardedHostPipeline
    if torch.cuda.device_count() < n_devices:
        pytest.skip(f"needs {n_devices} GPUs")
    h, w, n = 480, 640, 9
    frames, heads = synth.frames(n, h, w, seed=66), synth.head_v5(n, seed=67)
    want = _oracle(frames, heads)
    sp = ShardedHostPipeline(n, h, w, devices=list(range(n_devices)), chunk=2, slots=2)
    hf, hh = torch.from_numpy(frames).pin_memory(), torch.from_numpy(heads).pin_memory()
    ticket = sp.submit(hf, hh)
    assert ticket.bounds[0][0] == 0 and ticket.bounds[-1][1] == n
    _check_shards(ticket.wait(), ticket.bounds, want, 2)
    # a smaller batch right after (ragged shards, one GPU may get fewer images), then asynchronous reuse
    t2 = sp.submit(hf[:5], hh[:5])
    t3 = sp.submit(hf, hh)
    _check_shards(t2.wait(), t2.bounds, want[:5], 2)
    _check_shards(t3.wait(), t3.bounds, want, 2)
    with pytest.raises(ValueError, match="max_batch"):
        sp.submit(torch.cat([hf, hf]), torch.cat([hh, hh]))
    sp.close()


def test_host_pipeline_crop_index_and_batch_limit():
    from shared.processing.pipeline import HostPipeline
    h, w, n = 480, 640, 5
    frames, heads = synth.frames(n, h, w, seed=68), synth.head_v5(n, seed=69)
    want = _oracle(frames, heads)
    hp = HostPipeline(n, h, w, "cuda", chunk=2, slots=2)
    ticket = hp.submit(torch.from_numpy(frames).pin_memory(), torch.from_numpy(heads).pin_memory())
    res = ticket.wait()
    for i in range(n):
        assert np.array_equal(ticket.crops_of(i).numpy(), want[i][4])
        assert int(res["crop_count"][i]) == len(want[i][2])
    with pytest.raises(ValueError, match="max_batch"):
        hp.submit(torch.zeros((n + 1, h, w, 3), dtype=torch.uint8), torch.zeros((n + 1, 25200, 85)))
