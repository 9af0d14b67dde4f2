"""PNG files of pileup tensors, encoded on the device (dp_png_encode, csrc/png.cu).

The reference writes one PNG per window from Python (Pileup.savePileupRGB, /root/reference/src/utils/Pileup.py:336-361) and
its training / inference drivers read them back through modules.dataset.PileupDataset (dataset.py:47-52).  Here the whole
batch of encoder output tensors becomes PNG byte streams in HBM in one launch and only finished files are copied to the
host; `write_files` stores them under the reference's file names.  No CPU fallback.
"""
from __future__ import annotations

import ctypes as C
import os
from concurrent.futures import ThreadPoolExecutor
from typing import Optional, Sequence

from . import _lib


def encode_batch(img, stream=None):
    """img: uint8 CUDA tensor [N, H, W, 4] (dp_encode_windows' out_img) -> pinned host uint8 tensor [N, file_bytes]; row i is
    the complete PNG file of image i.  Synchronises the stream it ran on."""
    import torch
    _lib.require_cuda()
    assert img.is_cuda and img.dtype == torch.uint8 and img.dim() == 4 and img.shape[3] == 4
    img = img.contiguous()
    n, H, W = int(img.shape[0]), int(img.shape[1]), int(img.shape[2])
    size = int(_lib.lib().dp_png_size(H, W))
    if size < 0:
        _lib.check(size, "dp_png_size")
    stride = (size + 15) // 16 * 16
    dev = img.device
    out = torch.empty((max(n, 1), stride), dtype=torch.uint8, device=dev)
    st = stream if stream is not None else torch.cuda.current_stream(dev)
    with torch.cuda.device(dev):
        _lib.check(_lib.lib().dp_png_encode(img.data_ptr(), n, H, W, out.data_ptr(), stride, C.c_void_p(st.cuda_stream)),
                   "dp_png_encode")
    host = torch.empty((max(n, 1), stride), dtype=torch.uint8).pin_memory()
    with torch.cuda.stream(st):
        host.copy_(out, non_blocking=True)
    st.synchronize()
    return host[:n, :size]


def write_files(paths: Sequence[str], img, threads: Optional[int] = None) -> int:
    """One PNG per image under `paths` (same order); returns the bytes written.  File writes run on a thread pool."""
    files = encode_batch(img).numpy()
    assert len(paths) == files.shape[0]

    def one(k):
        with open(paths[k], "wb") as f:
            f.write(files[k].tobytes())
        return files.shape[1]
    with ThreadPoolExecutor(max_workers=threads or min(32, os.cpu_count() or 1)) as ex:
        return sum(ex.map(one, range(len(paths))))
