"""On-disk depth formats written right after the forward path, with batched device->host copies and
background writers.

Formats (byte-compatible with the reference):
* ``.raw`` float32 -- ``Benchmark/image_io.py:128-168`` (save) / ``:102-124`` (load): header ``int32 h, int32 w,
  int32 cv_type (= CV_32F + ((d-1) << 3)), uint64 pixel_size (= 4*d)``, then the pixels as float32 in row-major
  order (the reference transposes and then iterates in Fortran order, which is the original C order).
* normalised 8-bit PNG -- ``Benchmark/benchmark.py:760-771`` + ``image_io.py:171-178``: ``(x - min) / (max - min +
  tiny)`` over the WHOLE batch, ``* 255`` truncated to uint8.
* 16-bit millimetre PNG -- ``ReSIDE/test_nyu.py:86-92`` (the uint16 map comes from ede_depth_epilogue).
"""
import os
import queue
import struct
import threading

import numpy as np
import torch

CV_32F, CV_CN_SHIFT, CV_CN_MAX = 5, 3, 512


def raw_float32_bytes(image: np.ndarray) -> bytes:
    """The exact file content of image_io.save_raw_float32_image for an [h,w] or [h,w,d] array."""
    if image.ndim == 2:
        (h, w), d = image.shape, 1
    elif image.ndim == 3:
        h, w, d = image.shape
    else:
        raise ValueError(f"expected an [h,w] or [h,w,d] image, got shape {image.shape}")
    if d > CV_CN_MAX:
        raise Exception("Cannot save image with more than 512 channels")
    head = struct.pack("i", h) + struct.pack("i", w) + struct.pack("i", CV_32F + ((d - 1) << CV_CN_SHIFT)) + struct.pack("Q", d * 4)
    return head + np.ascontiguousarray(image, dtype=np.float32).tobytes("C")


def save_raw_float32_image(file_name, image):
    with open(file_name, "wb") as f:
        f.write(raw_float32_bytes(np.asarray(image)))


def load_raw_float32_image(file_name):
    with open(file_name, "rb") as f:
        h, w, cv_type = struct.unpack("iii", f.read(12))
        pixel_size = struct.unpack("Q", f.read(8))[0]
        d = ((cv_type - CV_32F) >> CV_CN_SHIFT) + 1
        if d < 1 or d != pixel_size // 4:
            raise Exception("Incompatible pixel_size(%d) and cv_type(%d)" % (pixel_size, cv_type))
        if d > CV_CN_MAX:
            raise Exception("Cannot save image with more than 512 channels")
        data = np.frombuffer(f.read(), dtype=np.float32)
        return data.reshape(h, w) if d == 1 else data.reshape(h, w, d)


def normalised_u8(batch: np.ndarray) -> np.ndarray:
    """benchmark.py:760-762 + image_io.py:175-176 on a float32 batch: min/max over the whole batch."""
    b = np.asarray(batch, dtype=np.float32)
    lo, hi = b.min(), b.max()
    n = (b - lo) / (hi - lo + np.finfo(np.float32).tiny)
    return (np.float32(255.0) * n).astype("uint8")


class AsyncDepthWriter:
    """Batched D2H into pinned staging buffers on a side stream + a pool of writer threads, so that file I/O never
    stalls the forward stream (the reference writes synchronously per image, benchmark.py:764-773)."""

    def __init__(self, num_threads: int = 4):
        self._q = queue.Queue()
        self._threads = [threading.Thread(target=self._work, daemon=True) for _ in range(num_threads)]
        for t in self._threads:
            t.start()
        self._stream = None
        self._errors = []

    def _work(self):
        while True:
            job = self._q.get()
            if job is None:
                return
            try:
                job()
            except Exception as e:  # noqa: BLE001
                self._errors.append(e)
            finally:
                self._q.task_done()

    def write_batch(self, maps: torch.Tensor, paths, kind: str = "raw"):
        """maps: CUDA [N,H,W] / [N,1,H,W] float32 (kind 'raw' / 'png8') or uint16 (kind 'png16'); one path per map."""
        if maps.device.type != "cuda":
            raise RuntimeError("AsyncDepthWriter.write_batch: CUDA tensor required")
        if maps.dim() == 4:
            maps = maps[:, 0]
        if len(paths) != maps.shape[0]:
            raise ValueError(f"{len(paths)} paths for {maps.shape[0]} maps")
        if self._stream is None:
            self._stream = torch.cuda.Stream(maps.device)
        host = torch.empty(tuple(maps.shape), dtype=maps.dtype).pin_memory()
        done = torch.cuda.Event()
        self._stream.wait_stream(torch.cuda.current_stream(maps.device))
        with torch.cuda.stream(self._stream):
            host.copy_(maps, non_blocking=True)
            done.record(self._stream)
        maps.record_stream(self._stream)

        def job():
            done.synchronize()
            arr = host.numpy()
            if kind == "png8":
                arr = normalised_u8(arr)
            for a, p in zip(arr, paths):
                if kind == "raw":
                    save_raw_float32_image(p, a)
                else:
                    from PIL import Image
                    Image.fromarray(a).save(p if os.path.splitext(p)[1] else p + ".png")
        self._q.put(job)

    def cleanup(self):
        self._q.join()
        for _ in self._threads:
            self._q.put(None)
        for t in self._threads:
            t.join()
        if self._errors:
            raise self._errors[0]
