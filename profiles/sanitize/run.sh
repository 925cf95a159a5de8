#!/bin/bash
# AddressSanitizer + UndefinedBehaviorSanitizer over the plain-host code of the path (no GPU needed):
#   1. euromillioner_b200/csrc/host_logic.hpp through profiles/sanitize/host_logic_probe.cpp
#   2. the CPU oracle (oracle/spark_rf_oracle.cpp), built with the sanitizers and driven through its known-answer
#      tests and a small fit + predict (LD_PRELOAD of libasan because the host process is python)
# compute-sanitizer on the CUDA kernels is refused by this GPU pool; this is the sanitizer evidence there is.
# Usage: bash profiles/sanitize/run.sh > profiles/r2_sanitizers.txt 2>&1
set -e
cd "$(dirname "$0")/../.."
TMP=$(mktemp -d)
SAN="-fsanitize=address,undefined -fno-omit-frame-pointer -fno-sanitize-recover=undefined"
echo "== host_logic.hpp probe (g++ $SAN)"
g++ -O1 -g -std=c++17 $SAN -o $TMP/host_probe profiles/sanitize/host_logic_probe.cpp
ASAN_OPTIONS=detect_leaks=1 $TMP/host_probe
echo "== oracle under ASan/UBSan"
g++ -O1 -g -std=c++17 -fPIC -ffp-contract=off -fopenmp -Wall -Wno-unused-function $SAN -shared -o $TMP/liboracle.so oracle/spark_rf_oracle.cpp
cp oracle/liboracle.so $TMP/liboracle.keep
cp $TMP/liboracle.so oracle/liboracle.so
trap 'cp $TMP/liboracle.keep oracle/liboracle.so; touch oracle/liboracle.so' EXIT
LD_PRELOAD=$(gcc -print-file-name=libasan.so) ASAN_OPTIONS=detect_leaks=0 OMP_NUM_THREADS=4 \
  python -m pytest tests/test_oracle_kat.py tests/test_tools.py -x -q -p no:cacheprovider 2>&1 | tail -5
echo "== done: no sanitizer report above means none was raised (both sanitizers abort the process on a finding)"
