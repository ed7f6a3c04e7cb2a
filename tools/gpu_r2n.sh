#!/bin/bash
set -u
timeout 900 python -m pytest tests/test_gpu_peer.py -m gpu -x -q 2>&1 | tail -8
python tools/prof_case.py cfg4_open 3840 2160 64 auto 3 1 | tail -1
