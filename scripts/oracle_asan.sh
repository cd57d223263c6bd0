#!/bin/bash
# The CPU oracle under AddressSanitizer + UBSan: builds oracle/_build/libtstwo_oracle_asan.so and runs the oracle's own tests on it
# (TSTWO_ORACLE_SO points oracle/__init__.py at the instrumented library; libasan must be preloaded into python).
set -e
cd "$(dirname "$0")/.."
make -s -C oracle asan
ASAN=$(/usr/bin/gcc -print-file-name=libasan.so)
TSTWO_ORACLE_SO=$PWD/oracle/_build/libtstwo_oracle_asan.so LD_PRELOAD=$ASAN ASAN_OPTIONS=detect_leaks=0 \
  python -m pytest tests/test_oracle_fields.py tests/test_oracle_poly.py tests/test_oracle_merkle.py tests/test_oracle_fri.py tests/test_oracle_decommit.py -q -x "$@"
