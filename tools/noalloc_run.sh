#!/bin/bash
# the allocation-free entry point: alone in a fresh process, then after the whole GPU suite
set -x
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -15
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -8
