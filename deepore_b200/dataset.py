"""PileupDataset / TextColor: API of /root/reference/src/modules/dataset.py (:11-24, :26-57).

`PileupDataset(csv_path, transform)` reads the two-column summary CSV written by the window driver
(pileupGenerator.py:119: png path, label digits) and yields `(transform(RGBA image), int label vector)`.
`InMemoryPileupDataset` serves encoder output tensors directly (no PNG round trip) with the same item format.
"""
from __future__ import annotations

import numpy as np


class TextColor:
    """ANSI colour codes used by the reference's progress messages (dataset.py:11-24)."""
    PURPLE = '\033[95m'
    CYAN = '\033[96m'
    DARKCYAN = '\033[36m'
    BLUE = '\033[94m'
    GREEN = '\033[92m'
    YELLOW = '\033[93m'
    RED = '\033[91m'
    BOLD = '\033[1m'
    UNDERLINE = '\033[4m'
    END = '\033[0m'


class PileupDataset:
    def __init__(self, csv_path, transform=None):
        import os
        import pandas as pd
        tmp_df = pd.read_csv(csv_path, header=None, dtype=str)
        assert tmp_df[0].apply(lambda x: os.path.isfile(x)).all(), \
            "Some images referenced in the CSV file were not found"          # dataset.py:34-35
        self.transform = transform
        self.X_train = tmp_df[0]
        # one int64 vector per window; labels grow with the insert rows of their window (Pileup.py:320-333), so the
        # container is an object array (np.array() of ragged vectors, dataset.py:45, is an error on numpy >= 1.24)
        self.y_train = np.empty(len(tmp_df.index), dtype=object)
        for k, label in enumerate(tmp_df[1]):
            self.y_train[k] = np.array([int(x) for x in str(label)], dtype=np.int64)

    def __getitem__(self, index):
        import torch
        from PIL import Image
        img = Image.open(self.X_train[index])
        if self.transform is not None:
            img = self.transform(img)
        return img, torch.from_numpy(self.y_train[index])

    def __len__(self):
        return len(self.X_train.index)


class InMemoryPileupDataset:
    """Encoder output uint8 [N,H,W,4] (+ label strings) served like PileupDataset with ToTensor: float32 (4,H,W)/255."""

    def __init__(self, images, labels):
        self.images = images
        self.labels = labels

    def __getitem__(self, index):
        import torch
        img = torch.as_tensor(np.asarray(self.images[index])).permute(2, 0, 1).float() / 255.0
        label = torch.from_numpy(np.array([int(c) for c in self.labels[index]], dtype=np.int64))
        return img.contiguous(), label

    def __len__(self):
        return len(self.images)
