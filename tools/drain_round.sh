#!/bin/bash
# drain tests + timing of the batch-scale drain + its real (overlapped) kernel timeline
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "drain" 2>&1 | tail -5
timeout 200 python tools/drain_probe.py cfg3 100000 4096 0 -1 2>&1 | tail -4
timeout 200 python tools/drain_probe.py cfg4 100000 4096 0 2>&1 | tail -2
python tools/drain_timeline.py cfg3 100000 1024 0 2>&1 | tail -10
