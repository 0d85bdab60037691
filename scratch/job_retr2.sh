#!/bin/bash
set -u
timeout 300 python -m pytest tests/test_gpu_retrieval.py -x -q 2>&1 | tail -30
