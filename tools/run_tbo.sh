#!/bin/bash
# run the reference's own test-backend-ops (built unmodified into oracle/_ref) against the B200 plugin
# usage: tools/run_tbo.sh test|perf|support -o MUL_MAT [...]
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
REF="$ROOT/oracle/_ref"
if grep -q avx512vbmi /proc/cpuinfo && grep -q avx512_vnni /proc/cpuinfo && [ -x "$REF/avx512/test-backend-ops" ]; then REF="$REF/avx512"; fi
export GGML_BACKEND_PATH="$ROOT/llama.cpp-gfx906_b200/lib/libggml-b200.so"
exec "$REF/test-backend-ops" "$@"
