// latency micro-benchmarks of the synchronisation primitives the MMQ pipeline is built from (one CTA, sm_100a):
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I llama.cpp-gfx906_b200/csrc tools/ubench/sync_lat.cu -o llama.cpp-gfx906_b200/lib/sync_lat
#include "sm100.cuh"
#include <cstdio>
#include <cuda_runtime.h>
using namespace b200q::sm100;

__global__ void __launch_bounds__(128, 1) k(long long * out, const uint8_t * src) {
    extern __shared__ __align__(1024) uint8_t sm[];
    const uint32_t base = smem_u32(sm);
    const uint32_t bar0 = base, bar1 = base + 8, bar2 = base + 16, tslot = base + 32, buf = base + 1024;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        mbar_init(bar0, 1);
        mbar_init(bar1, 1);
        mbar_init(bar2, 1);
        mbar_init(bar0 + 64, 1);
        mbar_init(bar0 + 72, 1);
        mbar_init(bar0 + 80, 1);
        mbar_fence_init();
    }
    if (warp == 0) tmem_alloc(tslot, 256);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (threadIdx.x == 0) {
        long long t0, t1;
        const int N = 64;
        // (1) try_wait on an already completed phase
        mbar_arrive(bar0);
        mbar_wait(bar0, 0);
        t0 = clock64();
        for (int i = 0; i < N; ++i) mbar_wait(bar0, 0);
        t1 = clock64();
        out[0] = (t1 - t0) / N;
        // (2) arrive + wait by the same thread
        uint32_t ph = 1;
        t0 = clock64();
        for (int i = 0; i < N; ++i) {
            mbar_arrive(bar0);
            mbar_wait(bar0, ph);
            ph ^= 1;
        }
        t1 = clock64();
        out[1] = (t1 - t0) / N;
        // (3) tcgen05.commit (nothing outstanding) + wait
        ph = 0;
        t0 = clock64();
        for (int i = 0; i < N; ++i) {
            umma_commit(bar1);
            mbar_wait(bar1, ph);
            ph ^= 1;
        }
        t1 = clock64();
        out[2] = (t1 - t0) / N;
        // (4) issue cost of commit alone (N commits need N phases: use wait only at the end of pairs)
        t0 = clock64();
        for (int i = 0; i < N; ++i) {
            umma_commit(bar1);
            if (i & 1) {  // two arrivals would overflow a count-1 barrier: wait every time but measure separately below
            }
            mbar_wait(bar1, ph);
            ph ^= 1;
            umma_commit(bar1);
            mbar_wait(bar1, ph);
            ph ^= 1;
        }
        t1 = clock64();
        out[3] = (t1 - t0) / (2 * N);
        // (5) bulk copy 32 KB from L2-resident global + wait
        uint32_t p2 = 0;
        for (int rep = 0; rep < 3; ++rep) {
            t0 = clock64();
            mbar_expect_tx(bar2, 32768);
            bulk_g2s(buf, src, 32768, bar2);
            mbar_wait(bar2, p2);
            p2 ^= 1;
            t1 = clock64();
            out[4 + rep] = t1 - t0;
        }
        // (6) 4 bulk copies of 32 KB in flight
        t0 = clock64();
        mbar_expect_tx(bar2, 4 * 32768);
        for (int i = 0; i < 4; ++i) bulk_g2s(buf + i * 32768, src + i * 32768, 32768, bar2);
        mbar_wait(bar2, p2);
        p2 ^= 1;
        t1 = clock64();
        out[7] = t1 - t0;
        // (7) 4 MMAs 128x256x16 + commit + wait (operands: whatever is in smem)
        const uint64_t ad = umma_desc_k_sw128(buf), bd = umma_desc_k_sw128(buf + 16384);
        const uint32_t idesc = umma_idesc_bf16_f32(128, 256);
        uint32_t tm;
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tm) : "r"(tslot));
        uint32_t p3 = 0;
        const int counts[4] = { 4, 8, 16, 32 };
        for (int r = 0; r < 4; ++r) {
            t0 = clock64();
            for (int i = 0; i < counts[r]; ++i) umma_f16(tm, ad + 2 * (i & 3), bd + 2 * (i & 3), idesc, i != 0);
            umma_commit(bar0 + 80);
            t1 = clock64();
            mbar_wait(bar0 + 80, p3);
            p3 ^= 1;
            out[10 + 2 * r] = t1 - t0;              // issue time
            out[11 + 2 * r] = clock64() - t0;       // issue -> complete
        }
    }
    __syncthreads();
    // (9) three commits back to back, nobody waiting (one thread), then the whole-warp elect pattern of the MMQ issuer
    if (threadIdx.x == 0) {
        long long t0 = clock64();
        for (int i = 0; i < 64; ++i) {
            umma_commit(bar1);
            umma_commit(bar2);
            umma_commit(bar0 + 80);
        }
        out[20] = (clock64() - t0) / 64;
    }
    __syncthreads();
    if (warp == 0) {
        long long t0 = clock64();
        for (int i = 0; i < 64; ++i) {
            for (int sn = 0; sn < 2; ++sn) {
                if (elect_one()) {
                    umma_commit(bar1);
                    if (sn == 1) umma_commit(bar2);
                }
                __syncwarp();
            }
        }
        if (lane == 0) out[21] = (clock64() - t0) / 64;
    }
    __syncthreads();
    // (8) cross-warp ping-pong: warp 1 lane 0 <-> warp 2 lane 0 through two barriers (round trip)
    if (threadIdx.x == 32) {
        uint32_t ph = 0;
        long long t0 = clock64();
        for (int i = 0; i < 64; ++i) {
            mbar_arrive(bar0 + 64);
            mbar_wait(bar0 + 72, ph);
            ph ^= 1;
        }
        out[8] = (clock64() - t0) / 64;
    } else if (threadIdx.x == 64) {
        uint32_t ph = 0;
        for (int i = 0; i < 64; ++i) {
            mbar_wait(bar0 + 64, ph);
            ph ^= 1;
            mbar_arrive(bar0 + 72);
        }
    }
    __syncthreads();
    if (warp == 0) {
        uint32_t tm;
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tm) : "r"(tslot));
        tmem_dealloc(tm, 256);
    }
}

__global__ void init_bars(int) {}

int main() {
    long long * out;
    uint8_t *   src;
    cudaMallocManaged(&out, 64 * sizeof(long long));
    cudaMalloc(&src, 1 << 20);
    cudaMemset(src, 1, 1 << 20);
    for (int i = 0; i < 64; ++i) out[i] = -1;
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    for (int r = 0; r < 2; ++r) {
        k<<<1, 128, 200 * 1024>>>(out, src);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) {
            printf("error: %s\n", cudaGetErrorString(e));
            return 1;
        }
    }
    printf("try_wait (ready)            %lld cycles\n", out[0]);
    printf("arrive + wait (same thread) %lld cycles\n", out[1]);
    printf("commit + wait (same thread) %lld cycles\n", out[2]);
    printf("commit + wait (repeat)      %lld cycles\n", out[3]);
    printf("bulk copy 32 KB + wait      %lld %lld %lld cycles\n", out[4], out[5], out[6]);
    printf("4 x bulk copy 32 KB + wait  %lld cycles\n", out[7]);
    printf("cross-warp round trip       %lld cycles\n", out[8]);
    printf("3 commits, no waiter        %lld cycles per iteration (one thread), %lld (warp + elect)\n", out[20], out[21]);
    for (int r = 0; r < 4; ++r) printf("%2d MMAs 128x256x16 + commit: issue %lld, complete %lld cycles\n", 4 << r, out[10 + 2 * r], out[11 + 2 * r]);
    return 0;
}
