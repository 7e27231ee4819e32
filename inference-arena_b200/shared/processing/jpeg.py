"""JPEG decode on the GPU: what ``load_image_from_bytes``
(reference src/shared/processing/transforms.py:77-110) gets from ``cv2.imdecode`` +
``cvtColor(BGR2RGB)``, for baseline / extended-sequential Huffman JPEGs, bit for bit.  Only the
compressed bytes cross PCIe; frames are produced in HBM where K1 and K4 read them.

Progressive, arithmetic-coded, 12-bit, CMYK and multi-scan files raise ``UnsupportedJpeg``;
they (and PNG etc.) remain the business of the host-side ``transforms.load_image_from_bytes``.
"""

from __future__ import annotations

import ctypes
from ctypes import c_int32, c_size_t, c_uint8, c_uint16, c_uint32, c_void_p

import numpy as np
import torch

from . import _native
from ._native import check, lib

IA_ERR_UNSUPPORTED = -3


class UnsupportedJpeg(ValueError):
    """A valid JPEG of a flavour the device decoder does not take."""


class JpegJob(ctypes.Structure):
    """``ia_jpeg_job`` (include/ia_b200.h), field for field."""

    _fields_ = [
        ("width", c_int32), ("height", c_int32), ("out_width", c_int32), ("out_height", c_int32),
        ("orientation", c_int32), ("ncomp", c_int32), ("h", c_int32 * 3), ("v", c_int32 * 3),
        ("hmax", c_int32), ("vmax", c_int32), ("comp_w", c_int32 * 3), ("comp_h", c_int32 * 3),
        ("comp_pitch", c_int32 * 3), ("hexp", c_int32 * 3), ("vexp", c_int32 * 3),
        ("mcux", c_int32), ("mcuy", c_int32),
        ("blocks_per_mcu", c_int32), ("total_blocks", c_int32), ("restart_interval", c_int32),
        ("num_segments", c_int32), ("scan_offset", c_uint32), ("file_bytes", c_uint32),
        ("block_comp", c_uint8 * 12), ("block_bx", c_uint8 * 12), ("block_by", c_uint8 * 12),
        ("comp_dc", c_uint8 * 4), ("comp_ac", c_uint8 * 4), ("quant", (c_uint16 * 64) * 3),
        ("dc_counts", (c_uint8 * 16) * 4), ("dc_vals", (c_uint8 * 16) * 4),
        ("ac_counts", (c_uint8 * 16) * 4), ("ac_vals", (c_uint8 * 256) * 4),
        ("file_dev", c_void_p), ("clean_dev", c_void_p), ("segment_dev", c_void_p), ("coef_dev", c_void_p),
        ("plane_dev", c_void_p * 3), ("out_dev", c_void_p), ("chain_dev", c_void_p), ("tile_dev", c_void_p),
        ("ticket_dev", c_void_p),
    ]


def _bind():
    handle = lib()
    if not hasattr(handle, "_jpeg_bound"):
        handle.ia_jpeg_parse.argtypes = [c_void_p, c_size_t, ctypes.POINTER(JpegJob)]
        handle.ia_jpeg_workspace_bytes.argtypes = [ctypes.POINTER(JpegJob), ctypes.c_int]
        handle.ia_jpeg_plan.argtypes = [ctypes.POINTER(JpegJob), ctypes.c_int, ctypes.POINTER(c_void_p),
                                        ctypes.POINTER(c_void_p), c_void_p, c_size_t]
        handle.ia_jpeg_decode.argtypes = [ctypes.POINTER(JpegJob), c_void_p, ctypes.c_int, c_void_p, c_void_p]
        if handle.ia_jpeg_job_size() != ctypes.sizeof(JpegJob):
            raise RuntimeError("ia_jpeg_job layout mismatch between libia_b200.so and shared/processing/jpeg.py")
        handle._jpeg_bound = True
    return handle


def parse_jpeg(image_bytes: bytes) -> JpegJob:
    """Headers of one file (host only, no CUDA).  ``ValueError("Failed to decode ...")`` for
    anything that is not a JPEG — the reference's message, transforms.py:100-106."""
    if not image_bytes:
        raise ValueError("Failed to decode image from bytes: empty input")
    job = JpegJob()
    rc = _bind().ia_jpeg_parse(ctypes.cast(ctypes.c_char_p(image_bytes), c_void_p), len(image_bytes),
                               ctypes.byref(job))
    if rc == IA_ERR_UNSUPPORTED:
        raise UnsupportedJpeg("Failed to decode image from bytes: JPEG flavour not supported on the device "
                              "(progressive / arithmetic / 12-bit / CMYK / multi-scan)")
    if rc != 0:
        raise ValueError("Failed to decode image from bytes")
    return job


class JpegBatch:
    """Device buffers of one decoded batch (kept for inspection by tests and tools)."""

    def __init__(self, jobs, frames, status, workspace, files_dev, jobs_dev=None):
        self.jobs, self.frames, self.status = jobs, frames, status
        self.workspace, self.files_dev, self.jobs_dev = workspace, files_dev, jobs_dev

    def chain_records(self, i: int) -> torch.Tensor:
        """int32 [chunks, 16] hand-over records of image i (word 10: repeat rounds | tasks << 16)."""
        j = self.jobs[i]
        off = j.chain_dev - self.workspace.data_ptr()
        chunks = (j.file_bytes - j.scan_offset + 65535) // 65536 + 1
        return self.workspace[off:off + chunks * 64].view(torch.int32).view(chunks, 16)

    def rerun(self) -> None:
        """Launch the four decode kernels again on the resident compressed bytes (timing)."""
        n = len(self.frames)
        check(_bind().ia_jpeg_decode(self.jobs, self.jobs_dev.data_ptr(), n, self.status.data_ptr(),
                                     torch.cuda.current_stream(self.status.device).cuda_stream), "ia_jpeg_decode")

    def coefficients(self, i: int) -> torch.Tensor:
        """int16 [total_blocks, 64] of image i (natural order, DC absolute, not dequantised)."""
        j = self.jobs[i]
        off = j.coef_dev - self.workspace.data_ptr()
        return self.workspace[off:off + j.total_blocks * 128].view(torch.int16).view(j.total_blocks, 64)


_COPY_THREADS = 4
_PARALLEL_COPY_BYTES = 8 << 20
_pool = None


def _copy_pool():
    global _pool
    if _pool is None:
        from concurrent.futures import ThreadPoolExecutor
        _pool = ThreadPoolExecutor(max_workers=_COPY_THREADS, thread_name_prefix="ia-jpeg-stage")
    return _pool


class JpegDecoder:
    """Reusable decoder for one device: keeps two pinned staging buffers (compressed bytes +
    job records travel in one host-to-device copy each) and the device workspace between
    calls, so that a call costs the header parse, one memcpy into pinned memory and the
    launches.  Calls are ordered on the current stream; use one decoder per stream.  The
    workspace-backed views of a returned batch (``coefficients``, ``chain_records``) are valid
    until the next ``decode`` on this decoder; ``frames`` are fresh tensors owned by the caller."""

    def __init__(self, device: torch.device | str | None = None) -> None:
        if not torch.cuda.is_available():
            raise RuntimeError("JpegDecoder needs a CUDA device (no CPU fallback in this path)")
        dev = torch.device(device if device is not None else "cuda")
        self.device = dev if dev.index is not None else torch.device("cuda", torch.cuda.current_device())
        self._handle = _bind()
        self._staging = [None, None]  # pinned uint8 tensors
        self._busy = [None, None]  # event after the last copy out of each
        self._turn = 0
        self._workspace = None

    def _pinned(self, need: int) -> torch.Tensor:
        self._turn ^= 1
        k = self._turn
        if self._busy[k] is not None:
            self._busy[k].synchronize()  # the copy out of this buffer two calls ago
        if self._staging[k] is None or self._staging[k].numel() < need:
            self._staging[k] = torch.empty(max(need, 1 << 20) * 5 // 4, dtype=torch.uint8, pin_memory=True)
        return self._staging[k]

    def decode(self, blobs: list[bytes], out: list[torch.Tensor] | None = None) -> JpegBatch:
        """``out``: optional destination frames (uint8 CUDA tensors [H, W, 3], contiguous, one per
        file, of exactly the decoded size) instead of fresh ones — e.g. rows of a batch buffer."""
        return self.launch(self.stage(blobs), out)

    def stage(self, blobs: list[bytes]):
        """Host half of ``decode``: parse the headers and copy the files into pinned memory.
        Touches no stream, so a worker thread may run it for the next batch while the current
        one is being launched (at most one ``stage`` ahead of ``launch`` per decoder)."""
        handle = self._handle
        n = len(blobs)
        jobs = (JpegJob * max(n, 1))()
        for i, blob in enumerate(blobs):
            if not blob:
                raise ValueError("Failed to decode image from bytes: empty input")
            rc = handle.ia_jpeg_parse(ctypes.cast(ctypes.c_char_p(blob), c_void_p), len(blob), ctypes.byref(jobs[i]))
            if rc == IA_ERR_UNSUPPORTED:
                raise UnsupportedJpeg("Failed to decode image from bytes: JPEG flavour not supported on the "
                                      "device (progressive / arithmetic / 12-bit / CMYK / multi-scan)")
            if rc != 0:
                raise ValueError("Failed to decode image from bytes")
        if n == 0:
            return (jobs, 0, [], 0, None, 0)
        job_bytes = n * ctypes.sizeof(JpegJob)
        offsets, total = [], (job_bytes + 255) & ~255  # job records first, then the files
        for blob in blobs:
            offsets.append(total)
            total += (len(blob) + 31) & ~15
        staging = self._pinned(total)
        turn = self._turn
        host = staging.numpy()

        def copy(lo: int, hi: int) -> None:
            for off, blob in zip(offsets[lo:hi], blobs[lo:hi]):
                host[off:off + len(blob)] = np.frombuffer(blob, dtype=np.uint8)

        if total >= _PARALLEL_COPY_BYTES and n >= 2 * _COPY_THREADS:
            # tens of megabytes: one core copies ~10 GB/s, less than PCIe moves; numpy releases the GIL
            step = -(-n // _COPY_THREADS)
            list(_copy_pool().map(lambda lo: copy(lo, min(lo + step, n)), range(0, n, step)))
        else:
            copy(0, n)
        return (jobs, n, offsets, total, staging, turn)

    def launch(self, staged, out: list[torch.Tensor] | None = None) -> JpegBatch:
        """Device half of ``decode``: one host-to-device copy and the kernels, on the current stream."""
        handle, dev = self._handle, self.device
        jobs, n, offsets, total, staging, turn = staged
        if n == 0:
            return JpegBatch(jobs, [], torch.empty(0, dtype=torch.int32, device=dev), None, None)
        job_bytes = n * ctypes.sizeof(JpegJob)
        with torch.cuda.device(dev):
            stream = torch.cuda.current_stream(dev)
            upload = torch.empty(total, dtype=torch.uint8, device=dev)
            if out is None:
                frames = [torch.empty((jobs[i].out_height, jobs[i].out_width, 3), dtype=torch.uint8, device=dev)
                          for i in range(n)]
            else:
                frames = list(out)
                if len(frames) != n:
                    raise ValueError(f"out holds {len(frames)} frames for {n} files")
                for i, f in enumerate(frames):
                    want = (jobs[i].out_height, jobs[i].out_width, 3)
                    if tuple(f.shape) != want or f.dtype != torch.uint8 or not f.is_cuda or not f.is_contiguous():
                        raise ValueError(f"out[{i}] must be a contiguous CUDA uint8 tensor of shape {want}")
            need = handle.ia_jpeg_workspace_bytes(jobs, n)
            if self._workspace is None or self._workspace.numel() < need:
                self._workspace = torch.empty(need * 5 // 4, dtype=torch.uint8, device=dev)
            workspace = self._workspace
            file_ptrs = (c_void_p * n)(*[upload.data_ptr() + off for off in offsets])
            out_ptrs = (c_void_p * n)(*[f.data_ptr() for f in frames])
            check(handle.ia_jpeg_plan(jobs, n, file_ptrs, out_ptrs, workspace.data_ptr(), workspace.numel()),
                  "ia_jpeg_plan")
            ctypes.memmove(staging.data_ptr(), ctypes.addressof(jobs), job_bytes)  # planned records
            upload.copy_(staging[:total], non_blocking=True)
            busy = torch.cuda.Event()
            busy.record(stream)
            self._busy[turn] = busy
            status = torch.empty(n, dtype=torch.int32, device=dev)
            check(handle.ia_jpeg_decode(jobs, upload.data_ptr(), n, status.data_ptr(), stream.cuda_stream),
                  "ia_jpeg_decode")
        return JpegBatch(jobs, frames, status, workspace, upload, upload[:job_bytes])


_decoders: dict = {}


def decode_jpeg_batch_cuda(blobs: list[bytes], device: torch.device | str | None = None) -> JpegBatch:
    """Decode ``blobs`` (whole JPEG files) on ``device``; ``.frames[i]`` is uint8 [H, W, 3] RGB.
    Uses one cached :class:`JpegDecoder` per (device, current stream)."""
    if not torch.cuda.is_available():
        raise RuntimeError("decode_jpeg_batch_cuda needs a CUDA device (no CPU fallback in this path)")
    dev = torch.device(device if device is not None else "cuda")
    if dev.index is None:
        dev = torch.device("cuda", torch.cuda.current_device())
    key = (dev.index, torch.cuda.current_stream(dev).cuda_stream)
    if key not in _decoders:
        _decoders[key] = JpegDecoder(dev)
    return _decoders[key].decode(blobs)


def load_jpeg_from_bytes(image_bytes: bytes) -> np.ndarray:
    """``load_image_from_bytes`` for a baseline JPEG, decoded on the GPU: RGB uint8 [H, W, 3]."""
    batch = decode_jpeg_batch_cuda([image_bytes])
    frame = batch.frames[0].cpu().numpy()  # synchronises
    if int(batch.status[0]) & 2:
        # the data ends before the last MCU without a marker behind it: libjpeg runs out of
        # input, cv2.imdecode returns None and the reference raises (transforms.py:104-105)
        raise ValueError("Failed to decode image from bytes")
    return frame
