#!/bin/bash
timeout 900 python -m pytest tests/test_framing.py -x -q -m gpu 2>&1 | tail -4
timeout 300 python scripts/frame_one.py 2>&1 | tail -1
timeout 300 python scripts/frame_one.py --mib 256 2>&1 | tail -1
