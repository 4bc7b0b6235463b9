#!/bin/bash
python tools/bm_bench.py 32 5
DWLS_LIB=exp/variants/libdwls_rb64.so python tools/bm_bench.py 32 5
DWLS_LIB=exp/variants/libdwls_rb128.so python tools/bm_bench.py 32 5
