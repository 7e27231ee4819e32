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

    def run(self, frames: torch.Tensor, head: torch.Tensor, events=None) -> PipelineOutput:
        """``events``: optional list of 4 CUDA events recorded before K1, before K2, between K3
        and K4, and after K4 (bench.py's per-kernel timing)."""
        n = frames.shape[0]
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


class HostPipeline:
    """Host buffers in, host buffers out (the e2e path).  ``chunk`` images move per stream
    slot; ``slots`` chunks are in flight at once (copy-in / compute / copy-out overlap)."""

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

        self.h_tensor = pinned((max_batch, 3, p.target, p.target), torch.float32)
        self.h_dets = pinned((max_batch, p.max_det, 6), torch.float32)
        self.h_keep = pinned((max_batch, p.max_det), torch.int32)
        self.h_count = pinned((max_batch,), torch.int32)
        self.h_status = pinned((max_batch,), torch.int32)
        self.h_crops = pinned((max_batch * self.rpi, 3, p.crop_size, p.crop_size), torch.float32)
        self.h_offsets = pinned((max_batch // chunk + 1, chunk + 1), torch.int32)

    @property
    def h2d_bytes_per_image(self) -> int:
        p = self.pipes[0]
        return self.src_h * self.src_w * 3 + p.anchors * (5 + p.classes) * 4

    def d2h_bytes(self, n: int, total_rois: int) -> int:
        p = self.pipes[0]
        per_image = 3 * p.target * p.target * 4 + p.max_det * (6 * 4 + 4) + 8 + 4
        return n * per_image + total_rois * 3 * p.crop_size * p.crop_size * 4

    def run(self, frames: torch.Tensor, heads: torch.Tensor) -> dict:
        """``frames`` uint8 [N,H,W,3] and ``heads`` float32 [N,A,85]: pinned (or pageable) HOST
        tensors.  Returns host views: tensor, dets, keep, count, status, crops (per chunk,
        ``rpi`` slots per image) and per-chunk roi offsets.  Synchronises before returning."""
        n = frames.shape[0]
        assert n <= self.max_batch
        cur = torch.cuda.current_stream(self.device)
        for s in self.streams:
            s.wait_stream(cur)
        for ci, lo in enumerate(range(0, n, self.chunk)):
            hi = min(lo + self.chunk, n)
            m = hi - lo
            slot = ci % self.slots
            pipe = self.pipes[slot]
            with torch.cuda.stream(self.streams[slot]):
                df, dh = self.d_frames[slot][:m], self.d_heads[slot][:m]
                df.copy_(frames[lo:hi], non_blocking=True)
                dh.copy_(heads[lo:hi], non_blocking=True)
                out = pipe.run(df, dh)
                self.h_tensor[lo:hi].copy_(out.tensor, non_blocking=True)
                self.h_dets[lo:hi].copy_(out.dets, non_blocking=True)
                self.h_keep[lo:hi].copy_(out.keep, non_blocking=True)
                self.h_count[lo:hi].copy_(out.count, non_blocking=True)
                self.h_status[lo:hi].copy_(out.status, non_blocking=True)
                self.h_offsets[ci, : m + 1].copy_(out.roi_offsets, non_blocking=True)
                # ROI count is only known on the device: copy the chunk's bounded crop slab
                rows = m * self.rpi
                self.h_crops[lo * self.rpi : lo * self.rpi + rows].copy_(out.crops[:rows], non_blocking=True)
        for s in self.streams:
            cur.wait_stream(s)
        cur.synchronize()
        return {"tensor": self.h_tensor[:n], "dets": self.h_dets[:n], "keep": self.h_keep[:n],
                "count": self.h_count[:n], "status": self.h_status[:n], "crops": self.h_crops,
                "roi_offsets": self.h_offsets}
