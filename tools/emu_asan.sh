#!/bin/bash
# Out-of-bounds check of the loss-side kernels without a GPU: their SOURCE runs on the host under
# tests/native/cuda_emu.h, compiled with AddressSanitizer (global buffers and the static "shared" arrays get redzones).
set -euo pipefail
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
OUT="$(mktemp -d)"
FLAGS="-O1 -g -fsanitize=address -fno-omit-frame-pointer -ffp-contract=off -pthread -shared -fPIC \
  -I$ROOT/geometry-grounded-gaussian-splatting_b200/csrc -I$ROOT/tests/native"
g++ $FLAGS -o "$OUT/ssim_emu_asan.so" "$ROOT/tests/native/ssim_emulated.cpp"
g++ $FLAGS -o "$OUT/ncc_emu_asan.so" "$ROOT/tests/native/patch_ncc_emulated.cpp"
ASAN_OPTIONS=detect_leaks=0 LD_PRELOAD="$(gcc -print-file-name=libasan.so)" python "$ROOT/tools/emu_asan_drive.py" "$OUT"
echo "emulated kernels: no out-of-bounds access"
