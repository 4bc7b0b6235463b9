#!/bin/bash
python -m pytest tests/test_gpu_parity.py tests/test_gpu_configs.py tests/test_gpu_batch.py tests/test_golden.py tests/test_cpp_facade.py -m gpu -x -q 2>&1 | tail -3
echo NEW; python bench.py --steps 5 --no-e2e --no-cpu --no-extras 2>/dev/null | python tools/kernel_table.py
echo NOALIAS; DWLS_CONF_ALIAS=0 python bench.py --steps 5 --no-e2e --no-cpu --no-extras 2>/dev/null | python tools/kernel_table.py | grep "value\|conf_kernel<left\|hsolve"
