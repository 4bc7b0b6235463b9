#!/bin/bash
python tools/facade_bench.py
DWLS_LIB=exp/variants/libdwls_x6.so python tools/facade_bench.py
DWLS_LIB=exp/variants/libdwls_x8.so python tools/facade_bench.py
python tools/facade_bench.py
