#!/usr/bin/env bash
# Installs the UNMODIFIED reference under baseline/_ref for bench.py --impl reference.
#  1. the prescribed pip install (the reference has no setup.py/pyproject, so this is
#     expected to fail; the outcome is logged to baseline/_ref/INSTALL.log);
#  2. fallback: byte-for-byte copy of the source tree to baseline/_ref/src and a build of
#     its .cu/.c files with the reference's own command line (README.md:11-14) plus
#     -arch=sm_100a, into baseline/_ref/bin.  No source file is edited.
set -u
HERE="$(cd "$(dirname "$0")" && pwd)"
SRC="${1:-/root/reference}"
DST="$HERE/_ref"
mkdir -p "$DST/bin"
LOG="$DST/INSTALL.log"
: > "$LOG"
if [ ! -d "$SRC" ]; then echo "reference tree $SRC not found" | tee -a "$LOG"; exit 1; fi
echo "== pip install attempt ($(date -u +%FT%TZ))" >> "$LOG"
TMP="$(mktemp -d)"; cp -r "$SRC" "$TMP/reference"; chmod -R u+w "$TMP/reference"
python -m pip install --no-index --no-build-isolation --find-links /opt/wheelhouse \
    --target "$DST/pip" "$TMP/reference" >> "$LOG" 2>&1
echo "pip exit code: $?" >> "$LOG"
rm -rf "$TMP"
echo "== source copy + nvcc build" >> "$LOG"
rm -rf "$DST/src"; cp -r "$SRC" "$DST/src"; chmod -R u+w "$DST/src"; rm -rf "$DST/src/.git"
NV="nvcc --std=c++11 -Werror cross-execution-space-call -lm -arch=sm_100a"
rc=0
for lab in lab1 lab2 lab3; do
  $NV -o "$DST/bin/${lab}_cuda_exe" "$DST/src/$lab/src/main.cu" >> "$LOG" 2>&1 || rc=1
  $NV -o "$DST/bin/${lab}_to_plot"  "$DST/src/$lab/src/to_plot.cu" >> "$LOG" 2>&1 || rc=1
done
$NV -o "$DST/bin/gpu_info" "$DST/src/gpu_info/src/main.cu" >> "$LOG" 2>&1 || rc=1
gcc -o "$DST/bin/lab1_cpu_exe" "$DST/src/lab1/src/main.c" -lm >> "$LOG" 2>&1 || rc=1
gcc -o "$DST/bin/lab2_cpu_exe" "$DST/src/lab2/src/main.c" -lm >> "$LOG" 2>&1 || rc=1
gcc -o "$DST/bin/hw1" "$DST/src/hw1/src/main.c" -lm >> "$LOG" 2>&1 || rc=1
gcc -o "$DST/bin/hw2" "$DST/src/hw2/src/main.c" >> "$LOG" 2>&1 || rc=1
( cd "$DST/src" && find . -type f \( -name '*.cu' -o -name '*.c' -o -name '*.py' \) -exec sha256sum {} + ) > "$DST/SOURCES.sha256"
echo "build exit code: $rc" >> "$LOG"
ls -la "$DST/bin" >> "$LOG"
exit $rc
