#!/bin/bash
mkdir -p gpurun_out
echo "== alone"; timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -3
echo "== after the others"; timeout 2400 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
