#!/bin/bash
timeout 900 python -m pytest tests/test_model_gpu.py -m gpu -q -s -k "trained or all_windows" 2>&1 | grep -v "^$" | tail -12
