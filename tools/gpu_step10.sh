#!/bin/bash
mkdir -p gpurun_out
timeout 200 python tools/cat_bench.py 2>&1 | tail -4
timeout 600 python -m pytest tests -q -m gpu -k "categorical or classes or cpp_class" 2>&1 | tail -6
