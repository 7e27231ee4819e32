"""The whole hot path as one object: preprocess -> decode+NMS -> fan-out crops.

``DevicePipeline.run`` chains the four kernels on device-resident inputs (what the reference's
``monolithic/app/inference.py:161-193`` does per request, here per batch and with no host
round trip).  ``HostPipeline.run`` is the reference-facing form: HOST buffers in, HOST
buffers out; it cuts the batch into chunks and drives H2D copy, kernels and D2H copy on
rotating CUDA streams from pinned staging memory so transfers overlap compute.

torch provides memory, streams and copies; all arithmetic is libia_b200.so.
"""

from __future__ import annotations

from dataclasses import dataclass

import numpy as np
import torch

from . import device as D

KERNELS_PER_STEP = 5  # K1, K2 (decode+filter), K3 (NMS), K4a (ROI rects), K4b (crop-resize)


@dataclass
class PipelineOutput:
    tensor: torch.Tensor       # float32 [N,3,640,640]
    dets: torch.Tensor         # float32 [N,max_det,6] letterbox space
    keep: torch.Tensor         # int32   [N,max_det]
    count: torch.Tensor        # int32   [N]
    status: torch.Tensor       # int32   [N]
    crops: torch.Tensor        # float32 [roi_capacity,3,224,224]
    roi_offsets: torch.Tensor  # int32   [N+1]
    rois: torch.Tensor         # int32   [roi_capacity,8]
    # JPEG input only, int32 [N]: 0 = libjpeg's pixels; bit 0 = damaged stream (concealed the way
    # libjpeg does, see ia_jpeg_decode); bit 1 = the data ends before the last block without a
    # marker — the reference raises "Failed to decode image from bytes" for that request
    jpeg_status: torch.Tensor | None = None


class DevicePipeline:
    """Pre-allocated buffers for batches of up to ``max_batch`` equally-sized frames."""

    def __init__(self, max_batch: int, src_h: int, src_w: int, device: torch.device | str = "cuda",
                 anchors: int = 25200, classes: int = 80, conf: float = 0.25, iou: float = 0.45,
                 target: int = 640, crop_size: int = 224, max_det: int = 32,
                 max_candidates: int = 4096, rois_per_image: int = 8) -> None:
        self.device = torch.device(device)
        self.max_batch, self.src_h, self.src_w = max_batch, src_h, src_w
        self.anchors, self.classes, self.conf, self.iou = anchors, classes, conf, iou
        self.target, self.crop_size, self.max_det = target, crop_size, max_det
        self.max_candidates = min(max_candidates, anchors)
        self.roi_capacity = max_batch * rois_per_image
        self.geom = D.letterbox_geometry(src_h, src_w, target)
        dev = self.device
        self.tensor = torch.empty((max_batch, 3, target, target), dtype=torch.float32, device=dev)
        self.crops = torch.empty((self.roi_capacity, 3, crop_size, crop_size), dtype=torch.float32, device=dev)
        need = D.lib().ia_decode_nms_workspace_bytes(max_batch, anchors, self.max_candidates)
        self.workspace = torch.empty(need, dtype=torch.uint8, device=dev)
        self._side = None  # side stream for run(overlap=True)

    def run(self, frames: torch.Tensor, head: torch.Tensor, events=None, overlap: bool = False) -> PipelineOutput:
        """``events``: optional list of 4 CUDA events recorded before K1, before K2, between K3
        and K4, and after K4 (bench.py's per-kernel timing).  ``overlap=True`` runs K1 on a
        side stream next to K2 -> K3 -> K4 (K1 shares no data with them): the latency-bound
        NMS / ROI kernels then hide behind the bandwidth-bound ones."""
        n = frames.shape[0]
        if overlap and not events:
            cur = torch.cuda.current_stream(self.device)
            if self._side is None:
                self._side = torch.cuda.Stream(device=self.device)
            self._side.wait_stream(cur)
            with torch.cuda.stream(self._side):
                tensor = D.preprocess_batch_cuda(frames, self.target, out=self.tensor[:n])
            dets, keep, count, status = D.decode_nms_cuda(
                head, self.conf, self.iou, max_candidates=self.max_candidates, max_det=self.max_det,
                workspace=self.workspace)
            crops, offsets, rois = D.crop_batch_cuda(frames, dets, count, self.geom, self.crop_size,
                                                     roi_capacity=self.roi_capacity, out=self.crops)
            cur.wait_stream(self._side)
            return PipelineOutput(tensor, dets, keep, count, status, crops, offsets, rois)
        if events:
            events[0].record()
        tensor = D.preprocess_batch_cuda(frames, self.target, out=self.tensor[:n])
        if events:
            events[1].record()
        dets, keep, count, status = D.decode_nms_cuda(
            head, self.conf, self.iou, max_candidates=self.max_candidates, max_det=self.max_det,
            workspace=self.workspace)
        if events:
            events[2].record()
        crops, offsets, rois = D.crop_batch_cuda(frames, dets, count, self.geom, self.crop_size,
                                                 roi_capacity=self.roi_capacity, out=self.crops)
        if events:
            events[3].record()
        return PipelineOutput(tensor, dets, keep, count, status, crops, offsets, rois)


class GraphedPipeline:
    """The device chain captured once into a CUDA graph and replayed: for small batches the
    six launches of a step are launch-latency-bound, a graph replays them in one submission.
    Inputs are copied into the captured static buffers; outputs are the captured tensors."""

    def __init__(self, pipe: DevicePipeline, batch: int) -> None:
        dev = pipe.device
        self.pipe, self.batch = pipe, batch
        self.frames = torch.zeros((batch, pipe.src_h, pipe.src_w, 3), dtype=torch.uint8, device=dev)
        self.head = torch.zeros((batch, pipe.anchors, 5 + pipe.classes), dtype=torch.float32, device=dev)
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):  # warm-up outside capture (lazy module loading, attributes)
            for _ in range(2):
                pipe.run(self.frames, self.head)
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.out = pipe.run(self.frames, self.head)

    def run(self, frames: torch.Tensor, head: torch.Tensor) -> PipelineOutput:
        self.frames.copy_(frames, non_blocking=True)
        self.head.copy_(head, non_blocking=True)
        self.graph.replay()
        return self.out

    def replay(self) -> PipelineOutput:
        """Re-run on whatever is in ``self.frames`` / ``self.head``."""
        self.graph.replay()
        return self.out


class RaggedPipeline:
    """The chain for a batch of requests as a service receives them: frames of different
    sizes, given as uint8 [H_i,W_i,3] CUDA tensors or as baseline JPEG files (``bytes``, decoded
    on the device).  One launch each for K1 (up to 32 shapes), K2, K3 and K4; the detections of
    image i are in its own letterbox space, exactly as the reference's per-request loop
    (monolithic/app/inference.py:153-193) produces them."""

    def __init__(self, device: torch.device | str = "cuda", anchors: int = 25200, classes: int = 80,
                 conf: float = 0.25, iou: float = 0.45, target: int = 640, crop_size: int = 224,
                 max_det: int = 32, max_candidates: int = 4096, rois_per_image: int = 8) -> None:
        self.device = torch.device(device)
        self.anchors, self.classes, self.conf, self.iou = anchors, classes, conf, iou
        self.target, self.crop_size, self.max_det = target, crop_size, max_det
        self.max_candidates = min(max_candidates, anchors)
        self.rois_per_image = rois_per_image
        self._decoder = None
        self._workspace = None

    def run(self, frames, head: torch.Tensor) -> PipelineOutput:
        if frames and isinstance(frames[0], (bytes, bytearray, memoryview)):
            if self._decoder is None:
                from .jpeg import JpegDecoder
                self._decoder = JpegDecoder(self.device)
            decoded = self._decoder.decode([bytes(f) for f in frames])
            frames, jpeg_status = decoded.frames, decoded.status
        else:
            jpeg_status = None
        n = len(frames)
        need = D.lib().ia_decode_nms_workspace_bytes(n, self.anchors, self.max_candidates)
        if self._workspace is None or self._workspace.numel() < need:
            self._workspace = torch.empty(need, dtype=torch.uint8, device=self.device)
        tensor = D.preprocess_ragged_cuda(frames, self.target)
        dets, keep, count, status = D.decode_nms_cuda(
            head, self.conf, self.iou, max_candidates=self.max_candidates, max_det=self.max_det,
            workspace=self._workspace)
        crops, offsets, rois = D.crop_ragged_cuda(frames, dets, count, self.target, self.crop_size,
                                                  roi_capacity=n * self.rois_per_image)
        return PipelineOutput(tensor, dets, keep, count, status, crops, offsets, rois, jpeg_status)


class HostTicket:
    """Handle of one submitted batch: ``wait()`` blocks until its outputs are in host memory
    and returns them (views into the pipeline's pinned result set; valid until the
    pipeline's next-but-one ``submit``)."""

    def __init__(self, owner, slots, pending, views: dict) -> None:
        self._owner, self._slots, self._pending, self._views, self.done = owner, slots, pending, views, False

    def finish_last_chunk(self) -> None:
        if self._pending is not None:
            self._owner._copy_crops(self._pending)
            self._pending = None

    def wait(self) -> dict:
        if not self.done:
            self.finish_last_chunk()
            for slot in self._slots:
                ev = torch.cuda.Event()
                ev.record(self._owner.streams[slot])
                ev.synchronize()
            self.done = True
            # Where image i's crops are: crops[crop_start[i] : crop_start[i] + crop_count[i]].  (The
            # crops of a chunk are packed, chunks start at multiples of chunk * rois_per_image:
            # callers index through these two vectors instead of re-deriving that.)
            v, o = self._views, self._owner
            n = v["count"].shape[0]
            offs = v["roi_offsets"].numpy()
            idx = np.arange(n)
            ci, li = idx // o.chunk, idx % o.chunk
            v["crop_start"] = ci * (o.chunk * o.rpi) + offs[ci, li]
            v["crop_count"] = offs[ci, li + 1] - offs[ci, li]
        return self._views

    def crops_of(self, i: int) -> torch.Tensor:
        """float32 [k_i, 3, S, S]: the model inputs of image i's detections (host tensor view)."""
        v = self.wait()
        lo = int(v["crop_start"][i])
        return v["crops"][lo : lo + int(v["crop_count"][i])]


class HostPipeline:
    """Host buffers in, host buffers out (the e2e path).  ``chunk`` images move per stream
    slot; ``slots`` chunks are in flight at once, so copy-in, kernels and copy-out overlap.
    ``submit`` is asynchronous and double-buffers the pinned result buffers: the copies of
    batch k+1 start while batch k is still draining, which keeps the PCIe link busy in both
    directions; ``run`` is the synchronous form."""

    RESULT_SETS = 2

    def __init__(self, max_batch: int, src_h: int, src_w: int, device: torch.device | str = "cuda",
                 chunk: int = 32, slots: int = 3, **kw) -> None:
        self.device = torch.device(device)
        self.chunk, self.slots, self.max_batch = chunk, slots, max_batch
        self.src_h, self.src_w = src_h, src_w
        self.pipes = [DevicePipeline(chunk, src_h, src_w, device, **kw) for _ in range(slots)]
        p = self.pipes[0]
        self.max_det, self.rpi = p.max_det, p.roi_capacity // chunk
        dev = self.device
        self.streams = [torch.cuda.Stream(device=dev) for _ in range(slots)]
        self.d_frames = [torch.empty((chunk, src_h, src_w, 3), dtype=torch.uint8, device=dev) for _ in range(slots)]
        self.d_heads = [torch.empty((chunk, p.anchors, 5 + p.classes), dtype=torch.float32, device=dev)
                        for _ in range(slots)]

        def pinned(shape, dtype):
            return torch.empty(shape, dtype=dtype, pin_memory=True)

        nchunks = (max_batch + chunk - 1) // chunk
        self.results = [{
            "tensor": pinned((max_batch, 3, p.target, p.target), torch.float32),
            "dets": pinned((max_batch, p.max_det, 6), torch.float32),
            "keep": pinned((max_batch, p.max_det), torch.int32),
            "count": pinned((max_batch,), torch.int32),
            "status": pinned((max_batch,), torch.int32),
            "jpeg_status": pinned((max_batch,), torch.int32),
            "crops": pinned((max_batch * self.rpi, 3, p.crop_size, p.crop_size), torch.float32),
            "roi_offsets": pinned((nchunks, chunk + 1), torch.int32),
        } for _ in range(self.RESULT_SETS)]
        self._tickets: list[HostTicket | None] = [None] * self.RESULT_SETS
        self._submits = 0
        self._chunks = 0
        self.last_d2h_bytes = 0
        self.last_h2d_bytes = 0
        self._decoders = None  # one JpegDecoder per stream slot, created on first JPEG submit

    @property
    def h2d_bytes_per_image(self) -> int:
        p = self.pipes[0]
        return self.src_h * self.src_w * 3 + p.anchors * (5 + p.classes) * 4

    def d2h_bytes(self, n: int) -> int:
        """Bytes copied back by the most recent batch (counted from the tensors copied: the
        fixed per-image outputs plus exactly the crops that exist)."""
        return self.last_d2h_bytes

    def _copy_crops(self, pending) -> None:
        """Second half of a chunk: once its ROI total has reached the host, copy back exactly
        that many crops (instead of the bounded slab) on the chunk's own stream."""
        slot, ev, res, ci, lo, m, crops = pending
        ev.synchronize()
        total = min(int(res["roi_offsets"][ci, m]), m * self.rpi)
        if total:
            with torch.cuda.stream(self.streams[slot]):
                base = lo * self.rpi
                res["crops"][base : base + total].copy_(crops[:total], non_blocking=True)
        self.last_d2h_bytes += total * crops[0].numel() * 4

    def submit(self, frames, heads: torch.Tensor, tensors_to_host: bool = True) -> HostTicket:
        """``frames`` uint8 [N,H,W,3] and ``heads`` float32 [N,A,85]: pinned (or pageable) HOST
        tensors, which must stay untouched until the ticket has been waited on.  ``frames`` may
        instead be a list of N baseline JPEG files (``bytes``) of the pipeline's frame size: they
        are uploaded compressed and decoded on the device.

        With both models on the GPU nothing big needs to cross PCIe except the
        frames: ``heads`` may be a CUDA tensor (the detector's output, used in place), and
        ``tensors_to_host=False`` leaves the two model inputs (``tensor``, ``crops``) on the
        device — only detections, counts and ROI offsets are copied back."""
        jpeg = isinstance(frames, (list, tuple))
        n = len(frames) if jpeg else frames.shape[0]
        if n > self.max_batch:
            raise ValueError(f"batch of {n} images exceeds this pipeline's max_batch={self.max_batch}")
        if jpeg and self._decoders is None:
            from concurrent.futures import ThreadPoolExecutor

            from .jpeg import JpegDecoder
            self._decoders = [JpegDecoder(self.device) for _ in range(self.slots)]
            # stages the next chunks (one per other stream slot) while chunk i is launched
            self._stager = ThreadPoolExecutor(max_workers=max(1, self.slots - 1))
        self.last_h2d_bytes = 0
        rs = self._submits % self.RESULT_SETS
        self._submits += 1
        if self._tickets[rs] is not None:
            self._tickets[rs].wait()  # that result set is about to be overwritten
        for t in self._tickets:  # an earlier batch's last chunk still owns a slot's crop buffer
            if t is not None:
                t.finish_last_chunk()
        res = self.results[rs]
        p = self.pipes[0]
        per_image = (3 * p.target * p.target * 4 if tensors_to_host else 0) + p.max_det * (6 * 4 + 4) + 4 + 4
        self.last_d2h_bytes = n * (per_image + (4 if jpeg else 0))  # + the decode status of each file
        cur = torch.cuda.current_stream(self.device)
        for s in self.streams:
            s.wait_stream(cur)
        used = set()
        pending = None
        starts = list(range(0, n, self.chunk))
        first_chunk = self._chunks

        def stage(ci):  # host half of the JPEG decode of chunk ci (parse + copy into pinned memory)
            lo = starts[ci]
            dec = self._decoders[(first_chunk + ci) % self.slots]
            return self._stager.submit(dec.stage, frames[lo:min(lo + self.chunk, n)])

        ahead = max(1, self.slots - 1)  # chunks i+1 .. i+ahead use the other slots' decoders
        staged = [stage(ci) for ci in range(min(ahead, len(starts)))] if jpeg else []
        for ci, lo in enumerate(starts):
            hi = min(lo + self.chunk, n)
            m = hi - lo
            slot = self._chunks % self.slots
            self._chunks += 1
            used.add(slot)
            pipe = self.pipes[slot]
            if jpeg:
                ready = staged.pop(0).result()
                if ci + ahead < len(starts):
                    staged.append(stage(ci + ahead))
            with torch.cuda.stream(self.streams[slot]):
                df, dh = self.d_frames[slot][:m], self.d_heads[slot][:m]
                if heads.is_cuda:
                    dh = heads[lo:hi]
                else:
                    dh.copy_(heads[lo:hi], non_blocking=True)  # first: the copy engine never waits for a kernel
                    self.last_h2d_bytes += dh.numel() * 4
                if jpeg:
                    decoded = self._decoders[slot].launch(ready, out=[df[i] for i in range(m)])
                    res["jpeg_status"][lo:hi].copy_(decoded.status, non_blocking=True)
                    self.last_h2d_bytes += sum(len(b) for b in frames[lo:hi])
                else:
                    df.copy_(frames[lo:hi], non_blocking=True)
                    self.last_h2d_bytes += df.numel()
                out = pipe.run(df, dh)
                res["roi_offsets"][ci, : m + 1].copy_(out.roi_offsets, non_blocking=True)
                ev = torch.cuda.Event()
                ev.record()
                if tensors_to_host:
                    res["tensor"][lo:hi].copy_(out.tensor, non_blocking=True)
                res["dets"][lo:hi].copy_(out.dets, non_blocking=True)
                res["keep"][lo:hi].copy_(out.keep, non_blocking=True)
                res["count"][lo:hi].copy_(out.count, non_blocking=True)
                res["status"][lo:hi].copy_(out.status, non_blocking=True)
                self.last_d2h_bytes += (m + 1) * 4
            if pending is not None:
                self._copy_crops(pending)  # previous chunk: its total is on the host by now
            pending = (slot, ev, res, ci, lo, m, out.crops) if tensors_to_host else None
        views = {"tensor": res["tensor"][:n], "dets": res["dets"][:n], "keep": res["keep"][:n],
                 "count": res["count"][:n], "status": res["status"][:n], "crops": res["crops"],
                 "jpeg_status": res["jpeg_status"][:n] if jpeg else None,
                 "roi_offsets": res["roi_offsets"]}
        ticket = HostTicket(self, sorted(used), pending, views)
        self._tickets[rs] = ticket
        return ticket

    def run(self, frames, heads: torch.Tensor, tensors_to_host: bool = True) -> dict:
        """Synchronous form: returns when every output of the batch is in host memory."""
        return self.submit(frames, heads, tensors_to_host).wait()

    def drain(self) -> None:
        for t in self._tickets:
            if t is not None:
                t.wait()


class ShardedTicket:
    """Handle of one batch submitted to a ShardedHostPipeline: ``wait()`` returns one result
    dict per GPU (the HostPipeline dict of that GPU's shard) — ``bounds[g]`` is the [lo, hi)
    slice of the batch GPU g processed."""

    def __init__(self, futures, bounds) -> None:
        self._futures, self.bounds = futures, bounds
        self._tickets = None

    def tickets(self) -> list:
        if self._tickets is None:
            self._tickets = [f.result() if f is not None else None for f in self._futures]
        return self._tickets

    def wait(self) -> list:
        return [t.wait() if t is not None else None for t in self.tickets()]


class ShardedHostPipeline:
    """ONE host batch spread over several GPUs of the box from ONE process: images are
    independent, so the batch is cut into contiguous shards (``sharding.shard_bounds``), each GPU
    runs its shard through its own ``HostPipeline`` — own streams, own device buffers, own pinned
    result buffers — driven by its own host thread, and nothing is exchanged between the GPUs: no
    ``torch.distributed``, no NCCL, no peer access.  The inputs are read straight from the caller's
    one (pinned) host buffer; per-GPU results stay in per-GPU pinned buffers and are returned as a
    list (``ShardedTicket``), since gathering them would only add a host copy."""

    def __init__(self, max_batch: int, src_h: int, src_w: int, devices=None, chunk: int = 32, slots: int = 3,
                 **kw) -> None:
        from concurrent.futures import ThreadPoolExecutor

        from .sharding import shard_bounds
        if devices is None:
            devices = list(range(torch.cuda.device_count()))
        self.devices = [torch.device("cuda", d) if isinstance(d, int) else torch.device(d) for d in devices]
        if not self.devices:
            raise ValueError("ShardedHostPipeline needs at least one CUDA device")
        self.max_batch = max_batch
        self._shard_bounds = shard_bounds
        per = max(hi - lo for lo, hi in shard_bounds(max_batch, len(self.devices)))
        self.pipes = []
        for d in self.devices:
            with torch.cuda.device(d):
                self.pipes.append(HostPipeline(max(per, 1), src_h, src_w, d, chunk=min(chunk, max(per, 1)),
                                               slots=slots, **kw))
        # one launching thread per GPU: copies and launches of different GPUs are issued concurrently
        self._threads = [ThreadPoolExecutor(max_workers=1) for _ in self.devices]

    def _submit_one(self, g: int, frames, heads, tensors_to_host: bool):
        with torch.cuda.device(self.devices[g]):
            return self.pipes[g].submit(frames, heads, tensors_to_host)

    def submit(self, frames, heads: torch.Tensor, tensors_to_host: bool = True) -> ShardedTicket:
        """``frames`` uint8 [N,H,W,3] (or a list of N JPEG files) and ``heads`` float32 [N,A,85] in
        HOST memory, N <= max_batch; they must stay untouched until the ticket has been waited on."""
        n = len(frames) if isinstance(frames, (list, tuple)) else frames.shape[0]
        if n > self.max_batch:
            raise ValueError(f"batch of {n} images exceeds this pipeline's max_batch={self.max_batch}")
        bounds = self._shard_bounds(n, len(self.devices))
        futures = [self._threads[g].submit(self._submit_one, g, frames[lo:hi], heads[lo:hi], tensors_to_host)
                   if hi > lo else None for g, (lo, hi) in enumerate(bounds)]
        return ShardedTicket(futures, bounds)

    def run(self, frames, heads: torch.Tensor, tensors_to_host: bool = True) -> list:
        return self.submit(frames, heads, tensors_to_host).wait()

    def drain(self) -> None:
        for g, p in enumerate(self.pipes):
            self._threads[g].submit(p.drain).result()

    def close(self) -> None:
        self.drain()
        for t in self._threads:
            t.shutdown()
