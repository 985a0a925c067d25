#!/usr/bin/env bash
# Builds baseline/_ref/bin/lab1_driver: the reference's lab1 kernel (cut, unmodified, out of
# its to_plot.cu at build time) behind a host driver that generates inputs on the device.
set -eu
HERE="$(cd "$(dirname "$0")" && pwd)"
SRC="${1:-/root/reference}/lab1/src/to_plot.cu"
[ -f "$SRC" ] || SRC="$HERE/_ref/src/lab1/src/to_plot.cu"
mkdir -p "$HERE/_ref/bin"
# from the line that starts the kernel to its closing brace at column 0
awk '/^__global__ void kernel\(/{p=1} p{print} p&&/^}/{exit}' "$SRC" > "$HERE/_ref/lab1_kernel.inc"
grep -q "arr_3\[idx\] = arr_1\[idx\] - arr_2\[idx\]" "$HERE/_ref/lab1_kernel.inc"
nvcc -O3 -std=c++17 -arch=sm_100a -I"$HERE" -o "$HERE/_ref/bin/lab1_driver" "$HERE/lab1_driver.cu" -ldl
echo "built $HERE/_ref/bin/lab1_driver"
