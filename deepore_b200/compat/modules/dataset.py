"""`modules.dataset` -- PileupDataset / TextColor with the reference's behaviour (src/modules/dataset.py:11-57)."""
from deepore_b200.dataset import PileupDataset, TextColor  # noqa: F401
