"""Cached, data-parallel-aware input pipeline — an addition beside the drop-in
units/dataloader.py, for runs where the reference's loader starves the step.

The reference decodes, flips, bicubic-resizes and normalises every image again in every epoch, in the training process
(units/dataloader.py:5-32: DataLoader without workers; main.py:32-38).  At the 64x64 resolution the whole data set fits in
host memory after the deterministic part of that pipeline (48 KB per image as fp32): `CachedImageLoader` runs the
dataset's own transform ONCE per image with the random flip taken out, keeps the result in one (pinned) host tensor, and
per batch only gathers rows, applies the flip (a column reversal: it commutes with the symmetric bicubic resize and with
the per-channel normalisation) and copies to the device, the next batch being staged while the current one trains.

Same iteration contract as the DataLoader the loops consume (`len()`, iteration yielding `(images, labels)`,
shuffle=True, drop_last=True).  Data parallel: every rank draws the same permutation (a generator seeded with
`seed + epoch`, no communication) and takes its quarter-major slice of each GLOBAL batch (gan_b200.dist
.shard_quarter_major), so that the ranks' MiniBatchStdDev groups are those of the single-process global batch.
"""
import torch

from .dist import shard_quarter_major


def _split_flip(transform):
    """(transform without RandomHorizontalFlip, flip probability) of a torchvision Compose; (transform, 0.0) if none."""
    steps = getattr(transform, "transforms", None)
    if steps is None:
        return transform, 0.0
    keep, p = [], 0.0
    for t in steps:
        if type(t).__name__ == "RandomHorizontalFlip":
            p = float(t.p)
        else:
            keep.append(t)
    if p == 0.0:
        return transform, 0.0
    return type(transform)(keep), p


class CachedImageLoader:
    def __init__(self, dataset, batch_size, device=None, rank=0, world=1, seed=0, flip_p=None, device_cache=False):
        """`dataset`: a torchvision-style dataset with a `.transform` (units.dataloader.get_dataset / get_MNIST).
        `batch_size` is per rank; the global batch is batch_size * world.
        `device_cache`: keep the cached data set in HBM (48 KB per 64x64 RGB image: 200 000 images = 9.6 GB) and build
        every batch with ONE kernel launch (ganb200_gather_flip: row gather + horizontal flip), no host work and no
        host-to-device copy per batch; the batches equal the host path's bit for bit."""
        self.batch_size, self.rank, self.world, self.seed = int(batch_size), int(rank), int(world), int(seed)
        self.device = torch.device(device) if device is not None else torch.device("cpu")
        self.epoch = 0
        det, p = _split_flip(getattr(dataset, "transform", None))
        self.flip_p = p if flip_p is None else float(flip_p)
        original = getattr(dataset, "transform", None)
        try:  # one pass through the data set with the deterministic part of its own pipeline
            if original is not None:
                dataset.transform = det
            first, _ = dataset[0]
            n = len(dataset)
            pin = self.device.type == "cuda"
            self.images = torch.empty((n,) + tuple(first.shape), dtype=torch.float32, pin_memory=pin)
            self.labels = torch.empty(n, dtype=torch.int64)
            for i in range(n):
                x, y = dataset[i]
                self.images[i] = x
                self.labels[i] = int(y)
        finally:
            if original is not None:
                dataset.transform = original
        if self.world > 1 and self.batch_size % 4 != 0:
            raise ValueError("data parallel: the per-rank batch must be a multiple of 4 (MiniBatchStdDev groups)")
        self._stage = None
        self.device_cache = bool(device_cache) and self.device.type == "cuda"
        if self.device_cache:
            self.images_dev = self.images.to(self.device)
            self.labels_dev = self.labels.to(self.device)

    def __len__(self):
        return self.images.shape[0] // (self.batch_size * self.world)

    def _host_batch(self, idx, flips, out):
        torch.index_select(self.images, 0, idx, out=out)
        if bool(flips.any()):
            sel = flips.nonzero().flatten()
            out[sel] = out[sel].flip(-1)
        return out

    def __iter__(self):
        gen = torch.Generator().manual_seed(self.seed + self.epoch)
        self.epoch += 1
        n, bg = self.images.shape[0], self.batch_size * self.world
        perm = torch.randperm(n, generator=gen)
        flips = torch.rand(n, generator=gen) < self.flip_p   # one draw per sample of the epoch, same on every rank
        cuda = self.device.type == "cuda"
        if cuda and self._stage is None:  # two pinned staging buffers: batch k+1 is gathered while batch k is in flight
            shape = (self.batch_size,) + tuple(self.images.shape[1:])
            self._stage = [torch.empty(shape, dtype=torch.float32, pin_memory=True) for _ in range(2)]
            self._done = [None, None]
        for k in range(len(self)):
            glob = perm[k * bg:(k + 1) * bg]
            idx = shard_quarter_major(glob, self.rank, self.world) if self.world > 1 else glob
            if self.device_cache:
                from . import kernels as KN
                idx_dev = idx.to(self.device, non_blocking=True)
                yield (KN.backend().gather_flip(self.images_dev, idx_dev, flips[idx].to(self.device, non_blocking=True)),
                       self.labels_dev.index_select(0, idx_dev))
                continue
            if not cuda:
                out = torch.empty((self.batch_size,) + tuple(self.images.shape[1:]), dtype=torch.float32)
                yield self._host_batch(idx, flips[idx], out), self.labels[idx]
                continue
            slot = k % 2
            if self._done[slot] is not None:
                self._done[slot].synchronize()  # the copy that last read this staging buffer has finished
            host = self._host_batch(idx, flips[idx], self._stage[slot])
            dev = host.to(self.device, non_blocking=True)
            self._done[slot] = torch.cuda.Event()
            self._done[slot].record()
            yield dev, self.labels[idx].to(self.device, non_blocking=True)
