#!/bin/bash
timeout 600 python -m pytest tests/test_ops_gpu.py tests/test_fp16_gpu.py -m gpu -q --timeout 300 -k "gemm" 2>&1 | tail -8 | cut -c1-400
