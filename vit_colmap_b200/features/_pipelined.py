"""Host <-> device plumbing shared by the GPU extractors: double-buffered pinned staging, uploads
and downloads on their own streams, and a software pipeline over batches that keeps two batches
queued on the device, so the staging memcpy + host->device copy of batch k+2, the read-back of batch k
and the caller's database writes for batch k-1 all run under the kernels of batch k+1.

The reference processes one image at a time, synchronously (vit_extractor.py:727-762); the
per-image results are the same, only the schedule differs.
"""
from __future__ import annotations

import time
from concurrent.futures import ThreadPoolExecutor
from typing import Iterable, Iterator, List, Optional, Tuple

import numpy as np

Batch = Tuple[List[np.ndarray], Tuple[int, int], Tuple[int, int], object]  # images, orig_wh, resized_wh, extra


def decode_ahead(files, prepare, workers: int = 8, lookahead: int = 0):
    """Ordered iterator of ``(index, path, prepare(path))`` with the ``prepare`` calls (cv2.imread, the
    resize to a multiple of 14, a host-side detector: all release the GIL) running on a thread pool up
    to ``lookahead`` files ahead of the consumer (default 2 x workers).  A 1600 x 1200 PNG takes 30-40 ms to
    decode, an order of magnitude more than its share of the GPU batch, so a serial ``for f in files:
    cv2.imread(f)`` loop -- the reference's, vit_extractor.py:729-739 -- would leave the device idle.  An
    exception inside ``prepare`` is returned in place of the result; the order, and with it the image ids the
    single database writer hands out, is the sorted file order."""
    files = list(files)
    lookahead = lookahead or 2 * workers

    def run(path):
        try:
            return prepare(path)
        except Exception as e:  # noqa: BLE001
            return e

    with ThreadPoolExecutor(max_workers=workers, thread_name_prefix="vc-decode") as pool:
        jobs = []
        nxt = 0
        for k, path in enumerate(files):
            while nxt < len(files) and nxt < k + lookahead:
                jobs.append(pool.submit(run, files[nxt]))
                nxt += 1
            yield k, path, jobs[k].result()
            jobs[k] = None


class BatchPipeline:
    """Mixin.  The extractor provides ``self.device`` and
    ``_device_stage(batch_u8, orig_wh, resized_wh, extra) -> (keypoints, desc_u8, counts, keep)``
    (device tensors [B,cap,k], [B,cap,d], [B]; ``keep`` is returned to the caller untouched)."""

    _SLOTS = 2
    _COPY_THREADS = 4   # numpy memcpy releases the GIL: pageable -> pinned staging in parallel

    def _pool(self) -> ThreadPoolExecutor:
        if getattr(self, "_copy_pool", None) is None:
            self._copy_pool = ThreadPoolExecutor(self._COPY_THREADS, thread_name_prefix="vc-stage")
            self.pipe_stats = {"stage_s": 0.0, "launch_s": 0.0, "finish_s": 0.0, "batches": 0}
        return self._copy_pool

    def _plumbing(self):
        import torch
        if getattr(self, "_copy_stream", None) is None:
            self._copy_stream = torch.cuda.Stream(device=self.device)
            self._pin_in = [None] * self._SLOTS
            self._dev_in = [None] * self._SLOTS
        return torch

    def _stage_upload(self, images_bgr: List[np.ndarray], slot: int = 0):
        """Host images -> device batch [B, H, W, 3] uint8 in input slot ``slot``.  Each image is
        copied once into the slot's pinned staging buffer and its DMA is enqueued immediately on the
        copy stream (the memcpy of image i+1 overlaps the DMA of image i).  Returns (batch, event)."""
        torch = self._plumbing()
        B = len(images_bgr)
        shape = (B,) + tuple(images_bgr[0].shape)
        n = int(np.prod(shape))
        if self._pin_in[slot] is None or self._pin_in[slot].numel() < n:
            torch.cuda.current_stream(self.device).synchronize()  # nobody may still read the old buffers
            self._copy_stream.synchronize()
            self._pin_in[slot] = torch.empty(n, dtype=torch.uint8).pin_memory()
            self._dev_in[slot] = torch.empty(n, dtype=torch.uint8, device=self.device)
        stage = self._pin_in[slot][:n].view(shape)
        dev = self._dev_in[slot][:n].view(shape)
        stage_np = stage.numpy()
        for img in images_bgr:
            if img.shape != shape[1:]:
                raise ValueError("images of one batch must have equal sizes")
        t0 = time.perf_counter()
        jobs = [self._pool().submit(np.copyto, stage_np[i], img) for i, img in enumerate(images_bgr)]
        with torch.cuda.stream(self._copy_stream):
            for i, job in enumerate(jobs):   # DMA of image i goes out as soon as its memcpy is done
                job.result()
                dev[i].copy_(stage[i], non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self._copy_stream)
        self.pipe_stats["stage_s"] += time.perf_counter() - t0
        return dev, ev

    def _upload(self, images_bgr: List[np.ndarray]):
        """Synchronous-order convenience: the batch is ready for the current stream."""
        torch = self._plumbing()
        dev, ev = self._stage_upload(images_bgr, 0)
        torch.cuda.current_stream(self.device).wait_event(ev)
        return dev

    def _enqueue_download(self, kps, du8, counts, keep, n_images: int, slot: int = 0):
        """Queues the device->host copies of one batch's outputs (whole [B, cap, .] buffers + counts)
        on the download stream, behind the batch's kernels, into the slot's pinned buffers.  Nothing
        blocks here; _finish waits on the returned event only, so later batches keep the GPU busy."""
        torch = self._plumbing()
        if getattr(self, "_d2h_stream", None) is None:
            self._d2h_stream = torch.cuda.Stream(device=self.device)
            self._pin_out = [dict() for _ in range(self._SLOTS)]
        done = torch.cuda.Event()
        done.record(torch.cuda.current_stream(self.device))
        hosts = []
        with torch.cuda.stream(self._d2h_stream):
            self._d2h_stream.wait_event(done)
            for name, t in (("kps", kps), ("desc", du8), ("counts", counts)):
                if t.device.type == "cpu":
                    hosts.append(t)
                    continue
                t = t.contiguous()
                buf = self._pin_out[slot].get(name)
                if buf is None or buf.numel() < t.numel() or buf.dtype != t.dtype:
                    buf = torch.empty(max(t.numel(), 1), dtype=t.dtype).pin_memory()
                    self._pin_out[slot][name] = buf
                host = buf[: t.numel()].view(t.shape)
                host.copy_(t, non_blocking=True)
                hosts.append(host)
            ev = torch.cuda.Event()
            ev.record(self._d2h_stream)
        self.last_d2h_bytes = sum(h.numel() * h.element_size() for h in hosts if h.is_pinned())
        # the device tensors stay referenced until the event has been waited on
        return (hosts, ev, (kps, du8, counts), keep, n_images)

    def _finish(self, pending):
        hosts, ev, _alive, keep, B = pending
        ev.synchronize()
        t0 = time.perf_counter()
        kps_h, du8_h, n = hosts[0].numpy(), hosts[1].numpy(), hosts[2].numpy()
        if int(n.sum()) * (kps_h.shape[2] * 4 + du8_h.shape[2]) > (1 << 20):
            rows = list(self._pool().map(lambda i: (kps_h[i, : n[i]].copy(), du8_h[i, : n[i]].copy()), range(B)))
        else:
            rows = [(kps_h[i, : n[i]].copy(), du8_h[i, : n[i]].copy()) for i in range(B)]
        self.pipe_stats["finish_s"] += time.perf_counter() - t0
        self.pipe_stats["batches"] += 1
        return rows, keep

    def _run_batch(self, images_bgr: List[np.ndarray], orig_wh, resized_wh, extra=None):
        """One batch, start to finish (the serial form of extract_batches)."""
        batch = self._upload(images_bgr)
        return self._finish(self._enqueue_download(*self._device_stage(batch, orig_wh, resized_wh, extra),
                                                   len(images_bgr), 0))

    def extract_batches(self, batches: Iterable[Batch], on_error=None) -> Iterator[Tuple[int, Optional[list], object]]:
        """Software pipeline over batches.  Yields (batch_index, per-image [(keypoints, descriptors)]
        or None when the batch failed, keep) in order.  Two batches are always queued on the device:
        while the host reads back batch k, stages batch k+2 and the caller handles batch k-1, the
        GPU runs batch k+1."""
        torch = self._plumbing()
        stream = torch.cuda.current_stream(self.device)
        it = iter(batches)

        def submit(slot):
            """stage + upload + launch + queue the download of the next batch; None at the end."""
            b = next(it, None)
            if b is None:
                return None
            try:
                dev, ev = self._stage_upload(b[0], slot)
                t0 = time.perf_counter()
                stream.wait_event(ev)
                out = self._device_stage(dev, b[1], b[2], b[3] if len(b) > 3 else None)
                pending = self._enqueue_download(*out, len(b[0]), slot)
                self.pipe_stats["launch_s"] += time.perf_counter() - t0
                return pending
            except Exception as e:  # noqa: BLE001
                return e

        queue = [submit(0)]
        if queue[0] is None:
            return
        queue.append(submit(1))
        k = 0
        while queue and queue[0] is not None:
            pending = queue.pop(0)
            result = None
            if isinstance(pending, Exception):
                if on_error is None:
                    raise pending
                on_error(k, pending)
            else:
                try:
                    result = self._finish(pending)
                except Exception as e:  # noqa: BLE001
                    if on_error is None:
                        raise
                    on_error(k, e)
            if queue and queue[-1] is not None:
                queue.append(submit(k % self._SLOTS))  # slot of batch k: kernels and copies are done
            yield (k, result[0] if result else None, result[1] if result else None)
            k += 1

    def extract_stream(self, items, batch_size: int, write, on_error, first_alone=None) -> None:
        """Groups ``items`` = iterator of (meta, image, orig_wh, resized_wh, extra) into equal-size
        batches (at most ``batch_size``; the very first item alone while ``first_alone()`` holds),
        runs them through extract_batches and calls ``write(meta, keypoints, descriptors)`` per image
        in input order.  ``extra`` (per image) reaches _device_stage as a list."""
        metas: List[list] = []

        def batches():
            pending: list = []

            def emit():
                metas.append([p[0] for p in pending])
                extra = [p[4] for p in pending]
                out = ([p[1] for p in pending], pending[0][2], pending[0][3],
                       None if all(e is None for e in extra) else extra)
                pending.clear()
                return out

            for item in items:
                alone = first_alone is not None and not metas and not pending and first_alone()
                if pending and (pending[0][2] != item[2] or pending[0][1].shape != item[1].shape
                                or len(pending) >= batch_size):
                    yield emit()
                pending.append(item)
                if alone:
                    yield emit()
            if pending:
                yield emit()

        for k, out, _ in self.extract_batches(batches(), on_error=on_error):
            if out is None:
                continue
            for meta, (kps, desc) in zip(metas[k], out):
                write(meta, kps, desc)
