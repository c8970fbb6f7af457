#!/bin/bash
cd "$(dirname "$0")/.."
timeout 600 python -m pytest tests/test_y_comm_single.py tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider 2>&1 | tail -15
