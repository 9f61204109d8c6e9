// oracle/llama_bench_shim.cpp — TEST INFRASTRUCTURE ONLY.  The reference's llama-bench (tools/llama-bench/llama-bench.cpp,
// compiled unmodified from /root/reference by `make -C oracle llama`) uses four helpers of the reference's `common` library
// and the four build-info strings its cmake generates.  Instead of building all of `common` (argument parser, downloads,
// chat templates ...), this file — our code — provides those few symbols with the declarations of common/common.h.
#include "common.h"

#include <thread>

int          LLAMA_BUILD_NUMBER = 0;
const char * LLAMA_COMMIT       = "reference-tree";
const char * LLAMA_COMPILER     = "g++";
const char * LLAMA_BUILD_TARGET = "x86_64-linux-gnu";

// physical-core heuristics live in common.cpp; the bench runs the model on the GPU backend, so the host thread count only
// feeds the CPU backend's pool
int32_t cpu_get_num_math() {
    const unsigned n = std::thread::hardware_concurrency();
    return n > 0 ? (int32_t) (n <= 4 ? n : n / 2) : 4;
}

bool set_process_priority(enum ggml_sched_priority prio) { return prio == GGML_SCHED_PRIO_NORMAL; }

// "0x..." hexadecimal thread mask -> bool per thread (common.cpp:parse_cpu_mask has the same contract)
bool parse_cpu_mask(const std::string & mask, bool (&boolmask)[GGML_MAX_N_THREADS]) {
    size_t start = mask.size() >= 2 && mask[0] == '0' && (mask[1] == 'x' || mask[1] == 'X') ? 2 : 0;
    const size_t n = mask.size() - start;
    if (n == 0 || n * 4 > GGML_MAX_N_THREADS) return false;
    for (size_t i = 0; i < n; ++i) {
        const char c = mask[start + i];
        int v;
        if (c >= '0' && c <= '9') v = c - '0';
        else if (c >= 'a' && c <= 'f') v = c - 'a' + 10;
        else if (c >= 'A' && c <= 'F') v = c - 'A' + 10;
        else return false;
        const size_t base = (n - 1 - i) * 4;
        for (int b = 0; b < 4; ++b) boolmask[base + b] = boolmask[base + b] || ((v >> b) & 1);
    }
    return true;
}
