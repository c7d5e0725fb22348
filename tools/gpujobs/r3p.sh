#!/bin/bash
# r3p: the no-alloc entry point: its test alone in a fresh process, then at the end of the whole GPU suite
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -15
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -6
