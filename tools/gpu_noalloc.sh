#!/bin/bash
echo "== alone (fresh process)"; timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x -s 2>&1 | tail -8
echo "== after the other tests"; timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | tail -4
