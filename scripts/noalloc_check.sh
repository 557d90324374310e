#!/bin/bash
echo "== alone, fresh process"; timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -4
echo "== after the other GPU tests"; timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -5
