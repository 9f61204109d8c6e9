// mxf_probe.cu — the experiment SURVEY H3 / the north star ask for before MXFP4 goes onto the native block-scaled tensor-core
// path: ONE CTA issues tcgen05.mma kind::mxf8f6f4.block_scale on a 128 x N x 128 problem whose operands the host has laid out
// exactly as the instruction reads them, so that the layouts can be validated against a CPU emulation and the issue rate
// measured before a production kernel is built around them (tools/mxf_probe.py drives it; reference behaviour being
// replaced: load_tiles_mxfp4 + the int8 MMA of ggml-cuda/mmq.cuh:699-762).
//
//   A  (weights, e2m1):  128 rows x 128 k.  Shared-memory image = the K-major 128-byte-swizzle tile of an 8-bit operand
//       (row r at r * 128 bytes, 16-byte chunk c at chunk c ^ (r & 7)); a chunk covers 16 k-elements: 8 bytes of packed
//       nibbles (element 2i in the low nibble of byte i) followed by 8 bytes of padding — the "16U4_ALIGN16B" form the
//       f8f6f4 kinds read 4-bit operands in.
//   B  (activations, e4m3): N rows x 128 k, one byte per element, same tile layout.
//   SFA / SFB (UE8M0, one per 32 k-elements): 32 x 16-byte blocks — byte [l][4 c + j] = scale of row 32 c + l, k-block j —
//       copied to TMEM with tcgen05.cp 32x128b.warpx4 (4 columns per 128 rows); the MMA of k-block j selects byte j through
//       the a_sf_id / b_sf_id fields of its instruction descriptor.
//   D = sum_k  A[m][k] 2^(sfa[m][k/32] - 127)  *  B[n][k] 2^(sfb[n][k/32] - 127),  fp32 in TMEM.
#include "../../include/b200q.h"
#include "sm100.cuh"

#include <cuda_runtime.h>

namespace b200q {
int ops_fail(int code, const char * msg);
}
using b200q::ops_fail;

namespace {
using namespace b200q::sm100;

constexpr int TILE_BYTES = 128 * 128;  // 128 rows x 128 bytes
constexpr int SF_BYTES   = 512;        // 32 x 16

// no-swizzle K-major descriptor of a block of 8-row x 16-byte core matrices stacked along M at `sbo` bytes
__device__ __forceinline__ uint64_t desc_k_noswizzle(uint32_t smem_addr, uint32_t lbo, uint32_t sbo) {
    return (uint64_t) ((smem_addr >> 4) & 0x3FFFu) | ((uint64_t) ((lbo >> 4) & 0x3FFFu) << 16) | ((uint64_t) ((sbo >> 4) & 0x3FFFu) << 32) |
           ((uint64_t) 1 << 46);
}
// cute::UMMA::InstrDescriptorBlockScaled: E2M1 x E4M3, UE8M0 scales, both operands K-major
__device__ __forceinline__ uint32_t idesc_mxf8f6f4(int M, int N, int a_sf_id, int b_sf_id) {
    return ((uint32_t) b_sf_id << 4) | (5u << 7) | (0u << 10) | ((uint32_t) (N >> 3) << 17) | (1u << 23) | ((uint32_t) (M >> 4) << 24) |
           ((uint32_t) a_sf_id << 29);
}
__device__ __forceinline__ void tmem_cp_32x128b_warpx4(uint32_t taddr, uint64_t sdesc) {
    asm volatile("tcgen05.cp.cta_group::1.32x128b.warpx4 [%0], %1;" ::"r"(taddr), "l"(sdesc) : "memory");
}
__device__ __forceinline__ void umma_mxf8f6f4(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t sfa, uint32_t sfb,
                                              uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::mxf8f6f4.block_scale [%0], %1, %2, %3, [%5], [%6], p;\n"
        "}\n" ::"r"(d_tmem),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate), "r"(sfa), "r"(sfb)
        : "memory");
}

extern __shared__ __align__(1024) uint8_t probe_smem[];

// smem: A tile | B tiles (N / 128) | SFA | SFB blocks (N / 128) | barriers
__global__ void __launch_bounds__(128, 1) mxf_probe_kernel(const uint8_t * a_img, const uint8_t * b_img, const uint8_t * sfa_img,
                                                           const uint8_t * sfb_img, float * d, int N, int iters, long long * cycles) {
    const uint32_t base = (smem_u32(probe_smem) + 1023u) & ~1023u;
    const int      nb   = N / 128;
    const uint32_t sA = base, sB = base + TILE_BYTES, sSFA = sB + nb * TILE_BYTES, sSFB = sSFA + SF_BYTES, sBar = sSFB + nb * SF_BYTES;
    const uint32_t bar_load = sBar, bar_mma = sBar + 8, tmem_slot = sBar + 16;
    const int      warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        mbar_init(bar_load, 1);
        mbar_init(bar_mma, 1);
        mbar_fence_init();
    }
    if (warp == 0) tmem_alloc(tmem_slot, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    uint32_t tmem_base;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));
    const uint32_t t_acc = tmem_base, t_sfa = tmem_base + 256, t_sfb = tmem_base + 260;  // 4 + 4 * nb columns of scale factors

    if (threadIdx.x == 0) {
        const uint32_t bytes = (uint32_t) (TILE_BYTES + nb * TILE_BYTES + SF_BYTES + nb * SF_BYTES);
        mbar_expect_tx(bar_load, bytes);
        bulk_g2s(sA, a_img, TILE_BYTES, bar_load);
        bulk_g2s(sB, b_img, (uint32_t) (nb * TILE_BYTES), bar_load);
        bulk_g2s(sSFA, sfa_img, SF_BYTES, bar_load);
        bulk_g2s(sSFB, sfb_img, (uint32_t) (nb * SF_BYTES), bar_load);
        mbar_wait(bar_load, 0);
        tc_fence_after();
        tmem_cp_32x128b_warpx4(t_sfa, desc_k_noswizzle(sSFA, 16, 128));
        for (int i = 0; i < nb; ++i) tmem_cp_32x128b_warpx4(t_sfb + 4 * i, desc_k_noswizzle(sSFB + i * SF_BYTES, 16, 128));
        const long long t0 = clock64();
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {  // 4 x K = 32: the descriptors advance by 32 bytes inside the swizzle atom
                const uint64_t ad = umma_desc_k_sw128(sA) + (uint64_t) (kk * 2);
                const uint64_t bd = umma_desc_k_sw128(sB) + (uint64_t) (kk * 2);
                umma_mxf8f6f4(t_acc, ad, bd, idesc_mxf8f6f4(128, N, kk, kk), t_sfa, t_sfb, (it | kk) ? 1u : 0u);
            }
        }
        umma_commit(bar_mma);
        mbar_wait(bar_mma, 0);
        if (cycles) *cycles = clock64() - t0;
    }
    __syncthreads();
    mbar_wait(bar_mma, 0);
    tc_fence_after();
    for (int c0 = 0; c0 < N; c0 += 32) {
        float v[32];
        tmem_ld32(t_acc + ((uint32_t) (warp * 32) << 16) + (uint32_t) c0, v);
#pragma unroll
        for (int j = 0; j < 32; ++j) d[(size_t) (warp * 32 + lane) * N + c0 + j] = v[j];
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem_base, 512);
}

}  // namespace

extern "C" {

// experiment entry point (declared in b200q.h): N in {128, 256}; images as described at the top of this file; d [128][N]
int b200q_mxf_probe(const void * a_img, const void * b_img, const void * sfa_img, const void * sfb_img, float * d, int n, int iters,
                    long long * cycles_dev, void * stream) {
    if (!a_img || !b_img || !sfa_img || !sfb_img || !d || (n != 128 && n != 256) || iters < 1) return ops_fail(B200Q_EINVAL, "mxf_probe: arguments");
    const int smem = 1024 + TILE_BYTES * (1 + n / 128) + SF_BYTES * (1 + n / 128) + 64;
    cudaError_t e = cudaFuncSetAttribute(mxf_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return ops_fail(B200Q_ECUDA, cudaGetErrorString(e));
    mxf_probe_kernel<<<1, 128, smem, (cudaStream_t) stream>>>((const uint8_t *) a_img, (const uint8_t *) b_img, (const uint8_t *) sfa_img,
                                                              (const uint8_t *) sfb_img, d, n, iters, cycles_dev);
    e = cudaGetLastError();
    return e == cudaSuccess ? 0 : ops_fail(B200Q_ECUDA, cudaGetErrorString(e));
}

}  // extern "C"
