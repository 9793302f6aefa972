"""The sampling / preview path of the training loops, kept off the critical path.

Reference (train/trainunits.py:52-56, called every 10-30 iterations, e.g. wgangp.py:87-88):
    gen_imgs = self.generator(self.fixed_noise).cpu()                       # device sync + 3 MB D2H, fp32
    save_image(make_grid(gen_imgs, padding=2, normalize=True).squeeze(), path)   # min/max, normalise, grid, quantise, PNG

Here the generator forward is the native no-grad forward (train-mode BatchNorm, running statistics updated — SURVEY B4),
and everything after it runs on the device in two launches (ganb200_minmax, ganb200_make_grid_u8: the uint8 HWC bytes
torchvision would hand to the PNG encoder, bit for bit).  The 4x smaller uint8 grid is copied to pinned host memory on a
side stream; a worker thread waits for that copy's event, encodes the PNG and writes the file.  `save_grid_async` returns
as soon as the launches are queued: the training loop never synchronises with the device for a preview.
`flush()` (also registered with atexit) blocks until every queued file is on disk.
"""
import atexit
import queue
import threading
import zlib
import struct

import torch

from . import kernels as KN

_q = None
_worker = None
_copy_stream = {}
_lock = threading.Lock()


def _png_bytes(arr):
    """uint8 [H, W, 3] numpy array -> PNG file bytes (PIL when present, else a minimal zlib encoder)."""
    try:
        import io
        from PIL import Image
        buf = io.BytesIO()
        Image.fromarray(arr).save(buf, format="PNG")
        return buf.getvalue()
    except ImportError:
        h, w, _ = arr.shape
        raw = b"".join(b"\x00" + arr[y].tobytes() for y in range(h))

        def chunk(tag, data):
            c = struct.pack(">I", len(data)) + tag + data
            return c + struct.pack(">I", zlib.crc32(tag + data) & 0xffffffff)
        return (b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 2, 0, 0, 0)) +
                chunk(b"IDAT", zlib.compress(raw, 6)) + chunk(b"IEND", b""))


def _run():
    while True:
        item = _q.get()
        try:
            if item is None:
                return
            host, event, path = item
            if event is not None:
                event.synchronize()          # only this thread waits for the device
            with open(path, "wb") as f:
                f.write(_png_bytes(host.numpy()))
        finally:
            _q.task_done()


def _ensure_worker():
    global _q, _worker
    with _lock:
        if _worker is None or not _worker.is_alive():
            _q = queue.Queue()
            _worker = threading.Thread(target=_run, name="ganb200-preview", daemon=True)
            _worker.start()
            atexit.register(flush)


def grid_u8(images, nrow=8, padding=2):
    """uint8 [GH, GW, 3] tensor on images' device: torchvision make_grid(padding, normalize=True) + save_image's
    quantisation."""
    return KN.backend().make_grid_u8(images.detach().contiguous(), nrow, padding)


def save_grid_async(images, path, nrow=8, padding=2):
    """Queue `path` (a .png) to be written from `images` [N, 1|3, H, W]; returns without waiting for the device."""
    _ensure_worker()
    grid = grid_u8(images, nrow, padding)
    if grid.is_cuda:
        dev = grid.device
        cur = torch.cuda.current_stream(dev)
        side = _copy_stream.get(dev)
        if side is None:
            side = _copy_stream[dev] = torch.cuda.Stream(device=dev)
        side.wait_stream(cur)
        host = torch.empty(grid.shape, dtype=torch.uint8, pin_memory=True)
        with torch.cuda.stream(side):
            host.copy_(grid, non_blocking=True)
            grid.record_stream(side)
            event = torch.cuda.Event()
            event.record(side)
        _q.put((host, event, path))
    else:
        _q.put((grid, None, path))


def flush():
    """Block until every queued preview has been written."""
    if _q is not None:
        _q.join()
