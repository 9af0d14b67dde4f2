"""deepore_b200 -- B200-native implementation of deePore's hot path (pileup encode + CNN/GRU forward).

Public surface mirrors the reference's Python API (src/utils/Pileup.py, src/modules/model.py,
src/modules/dataset.py); the work runs in hand-written sm_100a CUDA kernels behind the C-ABI declared in
include/deepore_b200.h.  There is no CPU fallback: every compute entry point raises if the CUDA library
is missing or no GPU is present.
"""
__version__ = "0.1.0"
