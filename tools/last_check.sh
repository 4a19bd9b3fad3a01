#!/bin/bash
# new tests + the drain fuzz with seeds of its own on the last tree
timeout 200 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "memory_back or score_pass or drain" 2>&1 | tail -3
timeout 300 python tools/drain_fuzz.py 9000 50; timeout 200 python tools/drain_fuzz.py 9500 40 tight 2>&1 | tail -3
