#!/usr/bin/env bash
# oracle/build_ref.sh — builds oracle/_ref/libkiss_ref.so from the UNMODIFIED reference
# headers where they lie under /root/reference (never copied into the repo).
#
# g++ 13 rejects include/biovoltron/container/xbit_vector.hpp (45 out-of-class member
# definitions omit the class's requires-clause).  The recipe writes a
# patched copy of that ONE header into a scratch directory that is placed first on the
# include path, compiles, and deletes the scratch copy again — only the .so stays.
#
# Flags follow the reference's own (CMakeLists.txt:31,35: -Ofast -mavx -mavx2 -mlzcnt -mbmi
# -fopenmp) minus -ffast-math (no floating point on the path) and without -march=native so
# that the .so runs on the GPU box's host CPU as well.
set -euo pipefail
here="$(cd "$(dirname "$0")" && pwd)"
ref="${KISS_REFERENCE:-/root/reference}"
out="$here/_ref"
if [ ! -d "$ref/include/biovoltron" ]; then
  echo "build_ref: $ref not present — keeping prebuilt $out/libkiss_ref.so (if any)" >&2
  exit 0
fi
mkdir -p "$out"
scratch="$out/.patched"
rm -rf "$scratch"; mkdir -p "$scratch/biovoltron/container"
python3 - "$ref" "$scratch" <<'PY'
import sys
ref, dst = sys.argv[1], sys.argv[2]
src = open(f'{ref}/include/biovoltron/container/xbit_vector.hpp').read()
head = 'template<std::size_t N, std::unsigned_integral Block,\n         std::copy_constructible Allocator>\n'
cut = src.index('class XbitVector : private')
fixed = src[:cut] + src[cut:].replace(head, head + '  requires(!std::same_as<Block, bool>)\n')
open(f'{dst}/biovoltron/container/xbit_vector.hpp', 'w').write(fixed)
PY
g++ -std=c++20 -O3 -mavx -mavx2 -mlzcnt -mbmi -mbmi2 -fopenmp -fPIC -shared -w \
    -I"$scratch" -I"$ref/include" -I"$ref/submodules/spdlog/include" \
    "$here/ref_harness.cpp" -o "$out/libkiss_ref.so"
rm -rf "$scratch"
echo "build_ref: built $out/libkiss_ref.so"
