#!/bin/bash
# Builds kiss_b200/option_a_demo: the reference's own fm_index.hpp + sorter call sites compiled against the B200 facade
# headers (INTEGRATION.md Option A).  Needs /root/reference (keeps a prebuilt binary otherwise).  The one reference header
# g++ 13 rejects (container/xbit_vector.hpp) is patched into a scratch directory, as oracle/build_ref.sh does.
set -e
here="$(cd "$(dirname "$0")/.." && pwd)"
ref=${KISS_REFERENCE:-/root/reference}
out="$here/kiss_b200/option_a_demo"
if [ ! -d "$ref/include" ]; then
  echo "build_option_a: $ref absent, keeping $( [ -x "$out" ] && echo the prebuilt binary || echo nothing )"; exit 0
fi
scratch="$(mktemp -d)"; mkdir -p "$scratch/biovoltron/container"
python3 - "$ref" "$scratch" <<'PY'
import sys
ref, dst = sys.argv[1], sys.argv[2]
src = open(f'{ref}/include/biovoltron/container/xbit_vector.hpp').read()
head = 'template<std::size_t N, std::unsigned_integral Block,\n         std::copy_constructible Allocator>\n'
cut = src.index('class XbitVector : private')
open(f'{dst}/biovoltron/container/xbit_vector.hpp', 'w').write(src[:cut] + src[cut:].replace(head, head + '  requires(!std::same_as<Block, bool>)\n'))
PY
g++ -std=c++20 -O2 -fopenmp -mavx2 -mlzcnt -mbmi -w \
    -I"$here/kiss_b200/host/include" -I"$here/include" -I"$scratch" -I"$ref/include" -I"$ref/submodules/spdlog/include" \
    "$here/kiss_b200/host/option_a/option_a_tu.cpp" -o "$out" -L"$here/kiss_b200" -lkiss_b200 -Wl,-rpath,'$ORIGIN'
rm -rf "$scratch"
echo "build_option_a: built $out"
