#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "tensor_memory or all_decode_configs" 2>&1 | tail -12
