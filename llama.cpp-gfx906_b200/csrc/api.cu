// api.cu — the C-ABI of libb200q.so (include/b200q.h): argument checking, kernel selection, launches.
// No compute happens on the host and nothing here falls back to a CPU path.
#include "../../include/b200q.h"

#include "blocks.cuh"
#include "dequant.cuh"
#include "gemv.cuh"
#include "gemv_kq.cuh"
#include "gemv_list.cuh"
#include "gemv_mma.cuh"
#include "../../include/b200q_debug.h"
#include "dense_mm.cuh"
#include "mmq.cuh"
#include "mmq_mxf.cuh"
#include "pdl.cuh"

#include <cuda_runtime.h>

#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <set>
#include <string>
#include <vector>

using namespace b200q;

namespace {

thread_local std::string g_err;

int fail(int code, const char * fmt, ...) {
    char    buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    g_err = buf;
    return code;
}

#define CUDA_TRY(expr)                                                                                     \
    do {                                                                                                   \
        cudaError_t e_ = (expr);                                                                           \
        if (e_ != cudaSuccess) {                                                                           \
            return fail(B200Q_ECUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e_), __FILE__, __LINE__); \
        }                                                                                                  \
    } while (0)

struct DevInfo {
    bool   ok = false;
    int    sms = 0;
    int    max_smem = 0;
    int    cc = 0;
};

DevInfo & dev_info() {
    static DevInfo    info[64];
    static std::mutex mu;
    int               dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) dev = 0;
    std::lock_guard<std::mutex> lk(mu);
    DevInfo & d = info[dev];
    if (!d.ok) {
        int major = 0, minor = 0;
        cudaDeviceGetAttribute(&d.sms, cudaDevAttrMultiProcessorCount, dev);
        cudaDeviceGetAttribute(&d.max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
        cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
        cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev);
        d.cc = major * 10 + minor;
        d.ok = d.sms > 0;
    }
    return d;
}

// opt every kernel in to the full dynamic shared memory carve-out once per (device, function)
int ensure_smem_attr(const void * fn, int bytes) {
    static std::mutex                            mu;
    static std::set<std::pair<int, const void *>> done;
    int                                          dev = 0;
    cudaGetDevice(&dev);
    std::lock_guard<std::mutex> lk(mu);
    if (done.count({ dev, fn })) return 0;
    CUDA_TRY(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
    done.insert({ dev, fn });
    return 0;
}

unsigned long long * g_ts = nullptr;  // debug timestamps (b200q_debug_timestamps)

inline int round_up(int x, int a) { return (x + a - 1) / a * a; }

// ---- GEMV configuration ------------------------------------------------------------------------------
struct GemvPlan {
    GemvParams p;
    int        grid = 0;
    int        smem = 0;
};

// fills the tiling fields of plan.p for `ncols` activation columns; false if it cannot fit in shared memory.
// One CTA per SM and (when it fits) <= ~100 KB of shared memory so that the next GEMV's CTAs can become
// co-resident under programmatic dependent launch and prefetch their weights.
bool plan_gemv(GemvPlan & plan, int type, int ncols, const DevInfo & di) {
    GemvParams & p = plan.p;
    p.row_bytes    = (int) ((long long) p.k / blck_elems(type) * blck_bytes(type));
    p.contiguous   = p.w_nb1 == p.row_bytes;
    p.smem_pitch   = p.contiguous ? (int) p.w_nb1 : round_up(p.row_bytes + 31, 16);
    const int pitch = p.smem_pitch;
    const int nu    = p.k / gemv_unit_elems(type);
    int       lpr   = 1;
    while (lpr < 32 && lpr < nu) lpr *= 2;
    p.lpr        = lpr;
    const int RW = 32 / lpr, RB = 2, NW = gemv_consumer_warps();

    // activations: int8 columns (padded to 128 B so the chunk swizzle stays inside a column) + (d, s) pairs
    p.act_col_bytes        = round_up(p.k, 128);
    p.ds_off               = gemv_act_off() + ncols * p.act_col_bytes;
    const long long act_end = (long long) p.ds_off + (long long) ncols * (p.k / 32) * 8;
    if (act_end > di.max_smem - 4096) return false;
    p.stage_off = round_up((int) act_end, 128);

    // pick (wpr, groups): stage = one pass of the owning group, aim for ~16 KB
    const double target = 16.0 * 1024;
    double       best   = 1e30;
    int          bw = 1, bg = 1;
    for (int wpr = 1; wpr <= NW; wpr *= 2) {
        if (wpr > 1 && (lpr < 32 || nu < 32 * wpr)) break;
        for (int g = 1; g * wpr <= NW && g <= 8; g *= 2) {
            if (g > 1 && p.nch > 1) break;
            const long long bytes = (long long) (NW / g / wpr) * RW * RB * pitch;
            double          cost  = bytes > target ? bytes / target : target / bytes;
            cost *= 1.0 + 0.08 * (wpr > 1 ? wpr : 0) + 0.02 * g;  // prefer no cross-warp reduction, fewer groups
            if (cost < best) {
                best = cost;
                bw   = wpr;
                bg   = g;
            }
        }
    }
    p.wpr    = bw;
    p.groups = bg;
    const long long rows_total = (long long) p.nch * p.m;
    for (;;) {
        const int       pass_rows  = (NW / p.groups / p.wpr) * RW * RB;
        const long long pass_bytes = (long long) pass_rows * pitch;
        p.passes                   = (int) std::max<long long>(1, (long long) (20 * 1024) / pass_bytes);
        // keep a few stages of work per CTA so the ring actually pipelines
        const long long rows_per_cta = std::max<long long>(1, rows_total / std::max(1, di.sms));
        while (p.passes > 1 && (long long) p.passes * pass_rows * 3 > rows_per_cta) --p.passes;
        p.stage_rows                = pass_rows * p.passes;
        const long long stage_bytes = 16 + (long long) p.stage_rows * pitch + 48;
        p.stage_bytes               = round_up((int) std::min<long long>(stage_bytes, 1 << 28), 128);
        const int soft              = 100 * 1024;  // two kernels co-resident
        int       S                 = (int) ((soft - p.stage_off) / p.stage_bytes);
        if (S < std::max(3, p.groups + 1)) S = (int) ((di.max_smem - p.stage_off) / p.stage_bytes);
        if (S >= std::max(2, p.groups + 1)) {
            p.nstages = std::min(S, 8);
            // a slot must always be consumed by the SAME warp group, else a fast group can test a slot's full-barrier one
            // phase early (parity ABA) when bulk copies complete out of order
            p.nstages -= p.nstages % p.groups;
            break;
        }
        if (p.groups > 1) { p.groups /= 2; continue; }   // fewer, larger stages cannot help; fewer groups can
        if (p.wpr < NW && lpr == 32 && nu >= 64 * p.wpr) { p.wpr *= 2; continue; }
        return false;
    }
    plan.smem = p.stage_off + p.nstages * p.stage_bytes;
    const int       pass_rows = p.stage_rows / p.passes;
    const long long want      = (rows_total + pass_rows - 1) / pass_rows;
    plan.grid                 = (int) std::max<long long>(1, std::min<long long>((long long) di.sms, want));
    return true;
}

int launch_gemv(GemvPlan & plan, int type, int ncols, uint32_t flags, cudaStream_t st) {
    gemv_kernel_t fn = gemv_kernel_for(type, ncols);
    if (!fn) return fail(B200Q_ENOTSUP, "gemv: unsupported type %d / ncols %d", type, ncols);
    const DevInfo & di = dev_info();
    if (int rc = ensure_smem_attr((const void *) fn, di.max_smem)) return rc;
    const bool pdl        = !(flags & B200Q_FLAG_NO_PDL);
    plan.p.early_prefetch = pdl && (flags & B200Q_FLAG_WEIGHTS_IMMUTABLE) ? 1 : 0;
    plan.p.ts             = g_ts;
    static const int early = getenv("B200Q_EARLY") ? atoi(getenv("B200Q_EARLY")) : 8;
    plan.p.early_stages   = early;

    static const bool debug = getenv("B200Q_DEBUG") != nullptr;
    if (debug) {
        fprintf(stderr, "[b200q] gemv type=%d n=%d k=%d m=%d nch=%d grid=%d smem=%d lpr=%d wpr=%d groups=%d passes=%d stage_rows=%d stage_bytes=%d S=%d contiguous=%d pdl=%d early=%d\n",
                type, ncols, plan.p.k, plan.p.m, plan.p.nch, plan.grid, plan.smem, plan.p.lpr, plan.p.wpr, plan.p.groups,
                plan.p.passes, plan.p.stage_rows, plan.p.stage_bytes, plan.p.nstages, plan.p.contiguous, (int) pdl,
                plan.p.early_prefetch);
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim          = dim3(plan.grid);
    cfg.blockDim         = dim3(gemv_threads());
    cfg.dynamicSmemBytes = plan.smem;
    cfg.stream           = st;
    cudaLaunchAttribute attr[1];
    attr[0].id                                         = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs                                          = attr;
    cfg.numAttrs                                       = pdl ? 1 : 0;
    CUDA_TRY(cudaLaunchKernelEx(&cfg, fn, plan.p));
    return 0;
}

// ---- K-quant decode GEMV (gemv_kq.cu): n = 1, q4_K / q6_K, up to KQ_MAX_SEG matrices sharing the activations ------
bool plan_kq(KqParams & p, int & grid, int & smem, int type, int64_t k, int nseg, const void * const * w, float * const * dst,
             const int64_t * m, const float * act, const DevInfo & di, const int32_t * ids = nullptr, long long expert_stride = 0,
             long long act_seg_stride = 0) {
    if (!kq_kernel_for(type)) return false;
    const bool kquant = type == T_Q4_K || type == T_Q5_K || type == T_Q6_K;
    const int  unit   = kquant ? 256 : 32;  // elements per unit of lane work (super-block | block)
    if (nseg < 1 || nseg > kq_max_segments(ids != nullptr) || k <= 0 || k % unit) return false;
    static const bool off   = getenv("B200Q_NO_KQ") != nullptr;
    static const bool off32 = getenv("B200Q_NO_KQ32") != nullptr;
    if (off || (!kquant && off32)) return false;
    p.nsb       = (int) (k / unit);
    p.k         = (int) k;
    p.row_bytes = p.nsb * blck_bytes(type);
    if ((kquant && p.row_bytes % 16) || ((uintptr_t) act & 15)) return false;
    static const int lpr_cap   = getenv("B200Q_KQ_LPR") ? atoi(getenv("B200Q_KQ_LPR")) : 8;
    static const int budget_kb = getenv("B200Q_KQ_SMEM_KB") ? atoi(getenv("B200Q_KQ_SMEM_KB")) : 110;
    const int NWc = kq_consumer_warps();
    p.off_q = 384;
    if (kquant) {
        p.off_dx    = p.off_q + p.nsb * 256;
        p.off_bs16  = round_up(p.off_dx + p.nsb * 4, 16);
        p.off_bs32  = p.off_bs16 + p.nsb * 32;
        p.off_stage = round_up(p.off_bs32 + p.nsb * 16, 128);
    } else {
        p.off_dx    = p.off_q + p.nsb * 32;  // int8 values, then (d, d * sum) per block
        p.off_bs16  = p.off_bs32 = 0;
        p.off_stage = round_up(p.off_dx + p.nsb * 8, 128);
    }
    if (act_seg_stride) {  // MoE down projection: one quantized activation image per segment
        if ((act_seg_stride * 4) % 16) return false;
        p.act_smem_stride = round_up(p.off_stage - p.off_q, 128);
        p.off_stage += (nseg - 1) * p.act_smem_stride;
    }
    if (ids && kquant && expert_stride % 16) return false;
    p.ids = ids; p.expert_stride = expert_stride; p.act_seg_stride = act_seg_stride;
    const int work = kquant ? p.nsb : std::max(1, p.nsb / 8);  // lane work per row in super-block equivalents
    // lanes per row: few lanes = many rows per stage = less per-stage overhead, but two launches must stay co-resident
    // per SM (the next one prefetches under this one's tail) with at least ~1.25 slots per warp in flight
    int lpr = 1, S = 0;
    while (lpr < lpr_cap && lpr < work) lpr *= 2;
    // a launch that streams for tens of microseconds (the output head) gains more from a deep ring of large stages than
    // from sharing the SM with its successor: give it the whole shared memory
    long long rows_total = 0;
    for (int g = 0; g < nseg; ++g) rows_total += m[g];
    const bool big    = rows_total * p.row_bytes >= (64ll << 20);
    const int  budget = big ? di.max_smem : budget_kb * 1024;
    for (;; lpr *= 2) {
        p.lpr            = lpr;
        p.rows_per_stage = 32 / lpr;
        p.stage_bytes    = round_up(p.rows_per_stage * p.row_bytes + 48, 128);  // + alignment lead, tail, funnel over-read
        S                = (budget - p.off_stage) / p.stage_bytes;
        if (S >= NWc + 2 || lpr == 32) break;
    }
    if (S < NWc / 2) S = (di.max_smem - p.off_stage) / p.stage_bytes;  // very long rows: one CTA per SM
    if (S < 2) return false;
    p.nstages = std::min(S, 32);  // each warp owns nstages / NW (+1) slots (gemv_kq.cu)
    long long total = 0;
    for (int g = 0; g < nseg; ++g) {
        if (m[g] <= 0 || m[g] > (1 << 30) || ((uintptr_t) w[g] & 15)) return false;
        p.seg[g].w      = (const uint8_t *) w[g];
        p.seg[g].dst    = dst[g];
        p.seg[g].m      = (int) m[g];
        p.seg[g].stage0 = (int) total;
        total += (m[g] + p.rows_per_stage - 1) / p.rows_per_stage;
    }
    if (total > (1ll << 30)) return false;
    p.nseg         = nseg;
    p.act          = act;
    p.total_stages = (int) total;
    grid           = (int) std::min<long long>(total, di.sms);
    smem           = p.off_stage + p.nstages * p.stage_bytes;
    return true;
}

int launch_kq(KqParams & p, int grid, int smem, int type, uint32_t flags, cudaStream_t st) {
    kq_kernel_t fn = kq_kernel_for(type, p.ids != nullptr);
    if (!fn) return fail(B200Q_ENOTSUP, "gemv_kq: unsupported type %d", type);
    const DevInfo & di = dev_info();
    if (int rc = ensure_smem_attr((const void *) fn, di.max_smem)) return rc;
    const bool pdl   = !(flags & B200Q_FLAG_NO_PDL);
    p.early_prefetch = pdl && (flags & B200Q_FLAG_WEIGHTS_IMMUTABLE) && !p.ids ? 1 : 0;  // ids belong to an earlier kernel
    static const bool debug = getenv("B200Q_DEBUG") != nullptr;
    if (debug) {
        fprintf(stderr, "[b200q] gemv_kq type=%d k=%d nseg=%d m0=%d grid=%d smem=%d lpr=%d rps=%d stage_bytes=%d S=%d stages=%d pdl=%d early=%d\n",
                type, p.k, p.nseg, p.seg[0].m, grid, smem, p.lpr, p.rows_per_stage, p.stage_bytes, p.nstages, p.total_stages,
                (int) pdl, p.early_prefetch);
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim          = dim3(grid);
    cfg.blockDim         = dim3(kq_threads());
    cfg.dynamicSmemBytes = smem;
    cfg.stream           = st;
    cudaLaunchAttribute attr[1];
    attr[0].id                                         = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs                                          = attr;
    cfg.numAttrs                                       = pdl ? 1 : 0;
    CUDA_TRY(cudaLaunchKernelEx(&cfg, fn, p));
    return 0;
}

// ---- decode GEMV for 2..8 columns on the MMA tile kernel (gemv_mma.cu): K-quants, dense rows ---------------------
bool plan_mma(MmaParams & p, int & grid, int & smem, int type, int64_t m, int64_t k, int n, const void * w, const float * act,
              long long a_nb1, float * dst, long long d_nb1, const DevInfo & di) {
    if (!mma_kernel_for(type)) return false;
    static const bool off = getenv("B200Q_NO_MMA_GEMV") != nullptr;
    const bool kquant = type == T_Q4_K || type == T_Q5_K || type == T_Q6_K;
    const int  unit   = kquant ? 256 : 32;  // elements per unit: super-block | block
    if (off || n < 2 || n > 8 || k <= 0 || k % unit || m <= 0 || m > (1 << 30)) return false;
    const int bb = blck_bytes(type);
    p.nsb        = (int) (k / unit);
    p.row_bytes  = p.nsb * bb;
    if (p.row_bytes % 16 || ((uintptr_t) w & 15) || ((uintptr_t) act & 15) || (a_nb1 & 15) || ((uintptr_t) dst & 3) || (d_nb1 & 3)) return false;
    // quantized activations: int8 images (K-quants: permuted per super-block, pitch k + 64; blocks: natural, k + 16), scales
    // (K-quants: one float, blocks: (d, d * sum)), K-quant partial sums — each column one unit longer than needed, so the 4
    // columns a warp reads at once sit in different banks — then the teams' partial tiles and the ring
    p.col_pitch = (int) k + (kquant ? 64 : 16);
    p.dx_pitch  = (p.nsb + 1) * (kquant ? 4 : 8);
    p.bs_pitch  = (p.nsb + 1) * mma_bs_bytes(type);
    p.off_q     = 384;
    p.off_dx    = p.off_q + n * p.col_pitch;
    p.off_bs    = round_up(p.off_dx + n * p.dx_pitch, 16);
    p.off_part  = round_up(p.off_bs + n * p.bs_pitch, 16);
    p.off_stage = round_up(p.off_part + mma_max_warps() * 2 * 512, 128);
    // units per stage: the largest chunk (<= the cap) that divides the row and leaves every warp a ring slot; 16 warps when
    // the ring has 16 slots, else 8.  Every row of a stage is copied as the 16-byte aligned window around its chunk
    // (lead <= 14 bytes); the pitch between rows makes the A-fragment loads conflict-free: LDS.64 of 4 rows x 4 lanes per
    // half-warp (q4_K, q5_K) -> pitch == 32 or 96 mod 128; LDS.32 of 8 rows x 4 lanes per warp (q6_K, blocks) -> pitch an
    // odd multiple of 16
    static const int kc_env = getenv("B200Q_MMA_KC") ? atoi(getenv("B200Q_MMA_KC")) : 0;
    static const int nw_env = getenv("B200Q_MMA_NW") ? atoi(getenv("B200Q_MMA_NW")) : 16;
    const int        kc_cap = kc_env ? kc_env : kquant ? 4 : 32;
    const bool       lds64  = type == T_Q4_K || type == T_Q5_K;
    int              S      = 0;
    p.nwarps                = 0;
    for (int nw = nw_env >= 16 ? 16 : 8; nw >= 8 && !p.nwarps; nw -= 8) {
        for (int c = 64; c >= 1; c >>= 1) {
            if (c > kc_cap || p.nsb % c) continue;
            p.kc          = c;
            p.chunk_bytes = c * bb;
            int pitch     = round_up(p.chunk_bytes + 14, 16);
            while (lds64 ? (pitch % 128 != 32 && pitch % 128 != 96) : (pitch / 16) % 2 == 0) pitch += 16;
            p.pitch       = pitch;
            p.stage_bytes = 16 * pitch;
            S             = (di.max_smem - p.off_stage - 64) / p.stage_bytes;  // + slack for the over-read behind the last row
            if (S >= nw && p.chunk_bytes >= 256) { p.nwarps = nw; break; }
            if (S >= nw && nw == 8) { p.nwarps = nw; break; }
        }
    }
    if (!p.nwarps) return false;
    const int NWc = p.nwarps;
    p.chunks  = p.nsb / p.kc;
    p.nstages = std::min(S, 40);
    p.groups  = (int) ((m + 15) / 16);
    grid      = std::min(p.groups, di.sms);
    // warps per 16-row group along k: the fewest sequential stages per warp
    static const int kp_env = getenv("B200Q_MMA_KP") ? atoi(getenv("B200Q_MMA_KP")) : 0;
    const int gpc = (p.groups + grid - 1) / grid;
    auto best_kp = [&](int chunks, int & kp_out) -> long long {
        long long best = -1;
        for (int kp = 1; kp <= 8; kp *= 2) {
            if (chunks % kp) continue;
            if (kp_env && kp != kp_env) continue;
            const int teams = NWc / kp;
            if (kp > 1 && teams > 8) continue;  // named barriers 1..8
            const long long cost = (long long) ((gpc + teams - 1) / teams) * (chunks / kp);
            if (best < 0 || cost < best) { best = cost; kp_out = kp; }
        }
        return best;
    };
    long long best = best_kp(p.chunks, p.kp);
    if (best < 0) return false;
    // long rows of a short matrix (ffn_down at n = 2): when a warp would walk >= 4 stages in sequence and half-size
    // stages let twice as many warps share a group, the smaller copies pay off (4096x14336, n = 2: 16.1 -> 13.0 us)
    if (!kc_env && best >= 4 && p.kc >= 2 && p.kc % 2 == 0) {
        const int kc2 = p.kc / 2, cb2 = kc2 * bb;
        int pitch2 = round_up(cb2 + 14, 16);
        while (lds64 ? (pitch2 % 128 != 32 && pitch2 % 128 != 96) : (pitch2 / 16) % 2 == 0) pitch2 += 16;
        const int S2 = (di.max_smem - p.off_stage - 64) / (16 * pitch2);
        int       kp2 = 0;
        const long long cost2 = S2 >= NWc && cb2 >= 256 ? best_kp(p.nsb / kc2, kp2) : -1;
        if (cost2 > 0 && cost2 * kc2 * 10 <= best * p.kc * 6) {
            p.kc = kc2; p.chunk_bytes = cb2; p.pitch = pitch2; p.stage_bytes = 16 * pitch2;
            p.chunks = p.nsb / kc2; p.nstages = std::min(S2, 40); p.kp = kp2;
        }
    }
    p.cpm = p.chunks / p.kp;
    // CTA pairs (148 = 2 * 74 packs perfectly) share the quantization when it is long enough to matter
    static const int cl_env = getenv("B200Q_MMA_CLUSTER") ? atoi(getenv("B200Q_MMA_CLUSTER")) : 2;
    const int        q_units = n * (kquant ? p.nsb : (p.nsb + 7) / 8);  // warp-steps of the quantizer per CTA
    p.cluster = cl_env >= 2 && grid >= 2 && q_units > 2 * p.nwarps ? 2 : 1;
    if (p.cluster == 2) grid &= ~1;
    p.w = (const uint8_t *) w; p.act = act; p.dst = dst; p.a_nb1 = a_nb1; p.d_nb1 = d_nb1;
    p.m = (int) m; p.k = (int) k; p.n = n;
    smem = p.off_stage + p.nstages * p.stage_bytes + 64;
    return true;
}

int launch_mma(MmaParams & p, int grid, int smem, int type, uint32_t flags, cudaStream_t st) {
    mma_kernel_t fn = mma_kernel_for(type);
    if (!fn) return fail(B200Q_ENOTSUP, "gemv_mma: unsupported type %d", type);
    const DevInfo & di = dev_info();
    if (int rc = ensure_smem_attr((const void *) fn, di.max_smem)) return rc;
    const bool pdl = !(flags & B200Q_FLAG_NO_PDL);
    static const bool early_on = !getenv("B200Q_MMA_EARLY") || atoi(getenv("B200Q_MMA_EARLY")) != 0;
    p.early_prefetch = pdl && early_on && (flags & B200Q_FLAG_WEIGHTS_IMMUTABLE) ? 1 : 0;
    static const bool debug = getenv("B200Q_DEBUG") != nullptr;
    if (debug) {
        fprintf(stderr, "[b200q] gemv_mma type=%d m=%d k=%d n=%d grid=%d smem=%d nw=%d cluster=%d kc=%d chunks=%d kp=%d pitch=%d stage_bytes=%d S=%d groups=%d\n",
                type, p.m, p.k, p.n, grid, smem, p.nwarps, p.cluster, p.kc, p.chunks, p.kp, p.pitch, p.stage_bytes, p.nstages, p.groups);
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim          = dim3(grid);
    cfg.blockDim         = dim3(32 * p.nwarps);
    cfg.dynamicSmemBytes = smem;
    cfg.stream           = st;
    cudaLaunchAttribute attr[2];
    int                 na = 0;
    if (pdl) {
        attr[na].id                                         = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[na].val.programmaticStreamSerializationAllowed = 1;
        ++na;
    }
    if (p.cluster > 1) {
        attr[na].id               = cudaLaunchAttributeClusterDimension;
        attr[na].val.clusterDim.x = (unsigned) p.cluster;
        attr[na].val.clusterDim.y = 1;
        attr[na].val.clusterDim.z = 1;
        ++na;
    }
    cfg.attrs    = attr;
    cfg.numAttrs = na;
    CUDA_TRY(cudaLaunchKernelEx(&cfg, fn, p));
    return 0;
}


// ---- persistent decode list (gemv_list.cu) ---------------------------------------------------------------------------
struct ListGeo {
    int unit, nsb, bb, rps_log2, nch, csb, qmode;
    long long image_bytes;  // quantized activation image in shared memory
};
// stage geometry of one op: a function of (type, k) only (gemv_list.cuh)
bool list_geometry(int type, int64_t k, ListGeo & g) {
    if (!list_type_ok(type)) return false;
    const bool kquant = type == T_Q4_K || type == T_Q5_K || type == T_Q6_K;
    g.unit = kquant ? 256 : 32;
    if (k <= 0 || k % g.unit || k / g.unit > 65535) return false;
    g.nsb   = (int) (k / g.unit);
    g.bb    = blck_bytes(type);
    g.qmode = kquant ? 0 : type == T_MXFP4 ? 2 : 1;
    const long long rb   = (long long) g.nsb * g.bb;
    int             rps;
    if (kquant && type != T_Q5_K && g.nsb <= LIST_REG_ACT_NSB) {
        // column-owning lanes (gemv_list.cuh): 32 / lpr rows side by side, up to 4 rows per lane in sequence
        int lpr = 1;
        while (lpr < 32 && lpr < g.nsb) lpr *= 2;
        const int sets = 32 / lpr;
        rps = sets;
        while (rps > 1 && rps * rb > LIST_STAGE_PAYLOAD) rps /= 2;
        if (rps == sets) {
            for (int R = 4; R > 1; R /= 2) {
                if ((long long) sets * R * rb <= LIST_STAGE_PAYLOAD) { rps = sets * R; break; }
            }
        }
    } else {
        const int work = kquant ? g.nsb : std::max(1, g.nsb / 8);
        int       lpr  = 1;
        while (lpr < 8 && lpr < work) lpr *= 2;
        rps = 32 / lpr;
        while (rps > 1 && rps * rb > LIST_STAGE_PAYLOAD) rps /= 2;
    }
    g.nch = 1;
    g.csb = g.nsb;
    if (rb > LIST_STAGE_PAYLOAD) {
        rps               = 1;
        const int max_csb = LIST_STAGE_PAYLOAD / g.bb;
        g.nch             = (g.nsb + max_csb - 1) / max_csb;
        g.csb             = (g.nsb + g.nch - 1) / g.nch;
        g.nch             = (g.nsb + g.csb - 1) / g.csb;
        if (g.nch > 16) return false;
    }
    g.rps_log2 = 0;
    while ((1 << g.rps_log2) < rps) ++g.rps_log2;
    g.image_bytes = kquant ? (long long) (((g.nsb * 256 + g.nsb * 4 + 15) & ~15) + g.nsb * 48) : (long long) g.nsb * 40;
    return true;
}

struct ListSegment {
    ListParams           p{};
    int                  grid = 0, smem = 0;
    std::vector<ListOp>  ops;
    std::vector<ListGen> gens;
    long long            total_units = 0;
};

// shared memory of a segment with these table sizes; nslots out
int list_smem_plan(int n_ops, int n_gens, long long act_stride, int max_smem, ListParams & p) {
    p.off_ops    = 8 * LIST_MAX_SLOTS + 16;  // behind the ring barriers and the barrier-point counters (gemv_list.cu)
    p.off_gens   = p.off_ops + n_ops * (int) sizeof(ListOp);
    p.off_act    = round_up(p.off_gens + n_gens * (int) sizeof(ListGen), 128);
    p.act_stride = (int) act_stride;
    const long long ring0 = round_up(p.off_act + 2 * (int) act_stride, 128);
    p.off_ring   = (int) ring0;
    const long long left = (long long) max_smem - ring0;
    return left <= 0 ? 0 : (int) std::min<long long>(left / LIST_SLOT_BYTES, LIST_MAX_SLOTS);
}

bool aligned16(const void * p, long long stride) { return ((uintptr_t) p & 15) == 0 && (stride & 15) == 0; }

}  // namespace

// ======================================================================================================
namespace b200q {
int ops_fail(int code, const char * msg) { return fail(code, "%s", msg); }  // ops.cu reports through the same channel
}

extern "C" {

int          b200q_version(void) { return 100; }
// development aid: device buffer of [grid][8] u64 that GEMV CTAs fill with %globaltimer stamps (NULL = off)
B200Q_API void b200q_debug_timestamps(void * dev_buf) { g_ts = (unsigned long long *) dev_buf; }
B200Q_API void b200q_debug_hang_buffer(void * host_mapped) { kq_set_hang_buffer((unsigned long long *) host_mapped); }
B200Q_API void b200q_debug_mmq_stamps(void * dev_buf) { mmq_set_debug((unsigned long long *) dev_buf); }
static unsigned long long * g_list_ts = nullptr;
// development aid: device buffer of [2 CTAs][8 warps][n_gens][8] u64 the list kernel fills with %globaltimer stamps (NULL = off)
B200Q_API void b200q_debug_list_stamps(void * dev_buf) { g_list_ts = (unsigned long long *) dev_buf; }
const char * b200q_last_error(void) { return g_err.c_str(); }
// host-side test hook (tests/test_plan_cpu.py): the launch plan gemv_mma would get on a device with `sms` SMs and
// `max_smem` bytes of shared memory per CTA; needs no GPU.  out[16] = nwarps, cluster, kc, chunks, kp, cpm, pitch,
// stage_bytes, nstages, groups, grid, smem, col_pitch, off_stage, chunk_bytes, row_bytes.  Returns 1 if a plan exists.
B200Q_API int b200q_debug_plan_gemv_mma(int type, int64_t m, int64_t k, int n, int sms, int max_smem, int * out) {
    DevInfo di;
    di.ok = true; di.sms = sms; di.max_smem = max_smem; di.cc = 100;
    MmaParams p{};
    int       grid = 0, smem = 0;
    alignas(16) static float dummy[4];
    if (!plan_mma(p, grid, smem, type, m, k, n, dummy, dummy, (long long) k * 4, dummy, (long long) m * 4, di)) return 0;
    const int v[16] = { p.nwarps, p.cluster, p.kc, p.chunks, p.kp, p.cpm, p.pitch, p.stage_bytes, p.nstages, p.groups, grid, smem,
                        p.col_pitch, p.off_stage, p.chunk_bytes, p.row_bytes };
    for (int i = 0; i < 16; ++i) out[i] = v[i];
    return 1;
}

int b200q_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

int b200q_device_info(int device, char * name, size_t name_len, size_t * total_mem, size_t * free_mem, int * cc_major,
                      int * cc_minor, int * sm_count, size_t * l2_bytes) {
    int n = b200q_device_count();
    if (device < 0 || device >= n) return fail(B200Q_ENODEVICE, "no CUDA device %d (count %d)", device, n);
    cudaDeviceProp prop;
    CUDA_TRY(cudaGetDeviceProperties(&prop, device));
    if (name && name_len) {
        snprintf(name, name_len, "%s", prop.name);
    }
    if (cc_major) *cc_major = prop.major;
    if (cc_minor) *cc_minor = prop.minor;
    if (sm_count) *sm_count = prop.multiProcessorCount;
    if (l2_bytes) *l2_bytes = (size_t) prop.l2CacheSize;
    if (total_mem || free_mem) {
        int cur = 0;
        CUDA_TRY(cudaGetDevice(&cur));
        CUDA_TRY(cudaSetDevice(device));
        size_t f = 0, t = 0;
        CUDA_TRY(cudaMemGetInfo(&f, &t));
        CUDA_TRY(cudaSetDevice(cur));
        if (total_mem) *total_mem = t;
        if (free_mem) *free_mem = f;
    }
    return 0;
}

// ---- memory / streams ----------------------------------------------------------------------------------
int b200q_set_device(int device) { CUDA_TRY(cudaSetDevice(device)); return 0; }
int b200q_malloc(void ** ptr, size_t bytes) {
    cudaError_t e = cudaMalloc(ptr, bytes);
    if (e == cudaErrorMemoryAllocation) {
        cudaGetLastError();
        return fail(B200Q_ENOMEM, "cudaMalloc(%zu) out of memory", bytes);
    }
    CUDA_TRY(e);
    return 0;
}
int b200q_free(void * ptr) { CUDA_TRY(cudaFree(ptr)); return 0; }
int b200q_memset(void * ptr, int value, size_t bytes, void * stream) {
    CUDA_TRY(cudaMemsetAsync(ptr, value, bytes, (cudaStream_t) stream));
    return 0;
}
int b200q_memcpy(void * dst, const void * src, size_t bytes, int kind, void * stream) {
    const cudaMemcpyKind k = kind == B200Q_H2D ? cudaMemcpyHostToDevice
                           : kind == B200Q_D2H ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice;
    CUDA_TRY(cudaMemcpyAsync(dst, src, bytes, k, (cudaStream_t) stream));
    return 0;
}
int b200q_stream_create(void ** stream) {
    cudaStream_t s;
    CUDA_TRY(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    *stream = (void *) s;
    return 0;
}
int b200q_stream_destroy(void * stream) { CUDA_TRY(cudaStreamDestroy((cudaStream_t) stream)); return 0; }
int b200q_stream_synchronize(void * stream) { CUDA_TRY(cudaStreamSynchronize((cudaStream_t) stream)); return 0; }
int b200q_host_alloc(void ** ptr, size_t bytes) { CUDA_TRY(cudaMallocHost(ptr, bytes)); return 0; }
int b200q_host_free(void * ptr) { CUDA_TRY(cudaFreeHost(ptr)); return 0; }
// ---- CUDA-graph capture of a stream's work (the plugin replays a decode step's launches as one graph; replaces the
//      reference's cudaGraph path, ggml-cuda.cu:2876-3100) -----------------------------------------------------------------
int b200q_graph_begin(void * stream) {
    CUDA_TRY(cudaStreamBeginCapture((cudaStream_t) stream, cudaStreamCaptureModeRelaxed));
    return 0;
}
int b200q_graph_end(void * stream, void ** exec_out) {
    if (!exec_out) return fail(B200Q_EINVAL, "graph_end: null output");
    *exec_out = nullptr;
    cudaGraph_t g = nullptr;
    cudaError_t e = cudaStreamEndCapture((cudaStream_t) stream, &g);
    if (e != cudaSuccess || !g) {
        cudaGetLastError();
        return fail(B200Q_ECUDA, "graph_end: capture failed: %s", cudaGetErrorString(e));
    }
    cudaGraphExec_t x = nullptr;
    e = cudaGraphInstantiate(&x, g, 0);
    cudaGraphDestroy(g);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(B200Q_ECUDA, "graph_end: instantiate failed: %s", cudaGetErrorString(e));
    }
    *exec_out = (void *) x;
    return 0;
}
int b200q_graph_launch(void * exec, void * stream) {
    if (!exec) return fail(B200Q_EINVAL, "graph_launch: null graph");
    CUDA_TRY(cudaGraphLaunch((cudaGraphExec_t) exec, (cudaStream_t) stream));
    return 0;
}
int b200q_graph_destroy(void * exec) {
    if (exec) CUDA_TRY(cudaGraphExecDestroy((cudaGraphExec_t) exec));
    return 0;
}
int b200q_event_create(void ** event) {
    cudaEvent_t e = nullptr;
    CUDA_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    *event = (void *) e;
    return 0;
}
int b200q_event_destroy(void * event) { CUDA_TRY(cudaEventDestroy((cudaEvent_t) event)); return 0; }
int b200q_event_record(void * event, void * stream) { CUDA_TRY(cudaEventRecord((cudaEvent_t) event, (cudaStream_t) stream)); return 0; }
int b200q_event_synchronize(void * event) { CUDA_TRY(cudaEventSynchronize((cudaEvent_t) event)); return 0; }
int b200q_stream_wait_event(void * stream, void * event) {
    CUDA_TRY(cudaStreamWaitEvent((cudaStream_t) stream, (cudaEvent_t) event, 0));
    return 0;
}
int b200q_enable_peer_access(int device, int peer) {
    if (device == peer) return 0;
    int can = 0;
    CUDA_TRY(cudaDeviceCanAccessPeer(&can, device, peer));
    if (!can) return fail(B200Q_ENOTSUP, "device %d cannot map the memory of device %d", device, peer);
    int cur = 0;
    CUDA_TRY(cudaGetDevice(&cur));
    CUDA_TRY(cudaSetDevice(device));
    cudaError_t e = cudaDeviceEnablePeerAccess(peer, 0);
    if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); e = cudaSuccess; }
    cudaSetDevice(cur);
    if (e != cudaSuccess) return fail(B200Q_ECUDA, "cudaDeviceEnablePeerAccess(%d -> %d): %s", device, peer, cudaGetErrorString(e));
    return 0;
}
int b200q_memcpy_peer(void * dst, int dst_device, const void * src, int src_device, size_t bytes, void * stream) {
    CUDA_TRY(cudaMemcpyPeerAsync(dst, dst_device, src, src_device, bytes, (cudaStream_t) stream));
    return 0;
}

// ggml's numbers: f16 / bf16 are plain values (the 32-value pseudo blocks the tensor-core kernel reads them in are internal)
int64_t b200q_blck_size(int type) { return (type == T_F16 || type == T_BF16) ? 1 : blck_elems(type); }
size_t  b200q_type_size(int type) { return (type == T_F16 || type == T_BF16) ? 2 : (size_t) blck_bytes(type); }
size_t  b200q_row_size(int type, int64_t ne0) {
    if (type == T_F16 || type == T_BF16) return (size_t) ne0 * 2;  // plain values (the 32-value pseudo blocks are a kernel detail)
    const int be = blck_elems(type);
    return be ? (size_t) (ne0 / be) * (size_t) blck_bytes(type) : 0;
}
int b200q_type_supported(int type) { return type != T_F32 && type != T_F32W && blck_elems(type) > 0; }
// largest row length every MUL_MAT / MUL_MAT_ID shape of `type` can be served at: the last-resort kernel (gemv_stream_kernel,
// one column, rows at an arbitrary pitch) needs the int8 activation column, its scales and two ring stages in shared
// memory.  Host-only (device attributes when a device is present, B200's otherwise): the plugin's supports_op asks this so
// that an op it accepts never comes back as B200Q_ENOTSUP.
int64_t b200q_max_k(int type) {
    const int be = blck_elems(type);
    if (type == T_F32 || dense_type(type)) return 1 << 17;  // the tensor-core kernel streams k in chunks: no row-length limit
    if (!be) return 0;
    DevInfo di = dev_info();
    if (!di.ok) { di.sms = 148; di.max_smem = 232448; }
    auto fits = [&](int64_t k) {
        GemvPlan plan{};
        GemvParams & p = plan.p;
        p.k = (int) k; p.m = 4096; p.nch = 1; p.mode = GEMV_MODE_BATCH; p.ne12 = 1; p.r2 = p.r3 = 1; p.n_used = p.n_expert = 1;
        p.w_nb1 = (long long) (k / be) * blck_bytes(type) + 16;  // not dense: the padded shared-memory pitch
        return plan_gemv(plan, type, 1, di);
    };
    const int64_t step = be == 32 ? 256 : be;
    int64_t lo = 0, hi = (1 << 17) / step;  // k = step * i
    while (lo < hi) {
        const int64_t mid = (lo + hi + 1) / 2;
        if (fits(mid * step)) lo = mid; else hi = mid - 1;
    }
    return lo * step;
}

// ---- dequantize / GET_ROWS -----------------------------------------------------------------------------
static int launch_get_rows(const GetRowsParams & p, int type, cudaStream_t st) {
    get_rows_kernel_t fn = get_rows_kernel_for(type);
    if (!fn) return fail(B200Q_ENOTSUP, "get_rows: unsupported type %d", type);
    const long long total = ((p.ne00 + 3) / 4) * p.n_rows;
    if (total <= 0) return 0;
    const DevInfo & di   = dev_info();
    const long long want = (total + 255) / 256;
    const int       grid = (int) std::min<long long>(want, (long long) di.sms * 16);
    launch_pdl(fn, dim3(grid), dim3(256), 0, st, p);  // get_rows_kernel starts with pdl_grid_sync()
    CUDA_TRY(cudaGetLastError());
    return 0;
}

int b200q_dequantize(int type, const void * src, float * dst, int64_t n_elems, void * stream) {
    const int be = blck_elems(type);
    if (!be || type == T_F32) return fail(B200Q_ENOTSUP, "dequantize: unsupported type %d", type);
    if (n_elems % be) return fail(B200Q_EINVAL, "dequantize: n_elems %lld not a multiple of the block size %d", (long long) n_elems, be);
    GetRowsParams p{};
    p.src      = (const uint8_t *) src;
    p.ids      = nullptr;
    p.dst      = dst;
    p.ne00     = n_elems;
    p.n_rows   = 1;
    p.ne10     = 1;
    p.ne11     = 1;
    p.nb01     = 0;
    p.nb1      = 0;
    p.dst_vec4 = ((uintptr_t) dst & 15) == 0;
    return launch_get_rows(p, type, (cudaStream_t) stream);
}

int b200q_get_rows(const b200q_get_rows_args * a, const void * src0, const int32_t * ids, float * dst, void * stream) {
    if (!a || !src0 || !ids || !dst) return fail(B200Q_EINVAL, "get_rows: null argument");
    const int be = blck_elems(a->type);
    if (!be) return fail(B200Q_ENOTSUP, "get_rows: unsupported type %d", a->type);
    if (a->ne00 % be) return fail(B200Q_EINVAL, "get_rows: ne00 not a multiple of the block size");
    GetRowsParams p{};
    p.src      = (const uint8_t *) src0;
    p.ids      = ids;
    p.dst      = dst;
    p.ne00     = a->ne00;
    p.ne10     = a->ne10;
    p.ne11     = a->ne11;
    p.n_rows   = a->ne10 * a->ne11 * a->ne12;
    p.nb01     = (int64_t) a->nb01;
    p.nb02     = (int64_t) a->nb02;
    p.nb03     = (int64_t) a->nb03;
    p.nb10     = (int64_t) a->nb10;
    p.nb11     = (int64_t) a->nb11;
    p.nb12     = (int64_t) a->nb12;
    p.nb1      = (int64_t) a->nb1;
    p.nb2      = (int64_t) a->nb2;
    p.nb3      = (int64_t) a->nb3;
    p.dst_vec4 = ((uintptr_t) dst & 15) == 0 && (a->nb1 & 15) == 0 && (a->nb2 & 15) == 0 && (a->nb3 & 15) == 0;
    return launch_get_rows(p, a->type, (cudaStream_t) stream);
}

// ---- MUL_MAT -------------------------------------------------------------------------------------------
// weight type as the kernels see it: f32 weights are the internal 32-value "block" type
static int wtype(int type) { return type == T_F32 ? T_F32W : type; }

static int check_mm(const b200q_mul_mat_args * a) {
    if (!a) return fail(B200Q_EINVAL, "mul_mat: null args");
    const int be = blck_elems(wtype(a->type));
    if (!be || a->type == T_F32W) return fail(B200Q_ENOTSUP, "mul_mat: unsupported weight type %d", a->type);
    // unquantized weights: any k (dense_mm.cu takes what the tensor-core kernel's 64-element chunks do not)
    if (a->k <= 0 || (!dense_type(wtype(a->type)) && a->k % be)) {
        return fail(B200Q_EINVAL, "mul_mat: k=%lld not a positive multiple of %d", (long long) a->k, be);
    }
    if (a->m < 0 || a->n < 0) return fail(B200Q_EINVAL, "mul_mat: negative dimension");
    if (a->ne02 <= 0 || a->ne03 <= 0 || a->ne12 <= 0 || a->ne13 <= 0 || a->ne12 % a->ne02 || a->ne13 % a->ne03) {
        return fail(B200Q_EINVAL, "mul_mat: bad batch dims");
    }
    return 0;
}

// unquantized weights: few columns (one pass over the rows, HBM-bound) or a k the 64-element chunks do not divide go to the
// CUDA-core kernel of dense_mm.cu; everything else to the tensor-core kernel
static bool use_dense(const b200q_mul_mat_args * a) {
    if (!dense_type(wtype(a->type))) return false;
    if (a->flags & B200Q_FLAG_FORCE_MMQ) return a->k % 64 != 0;
    return a->n <= 8 || a->k % 64 != 0;
}
static bool use_mmq(const b200q_mul_mat_args * a) {
    if (dense_type(wtype(a->type))) return !use_dense(a);
    if (a->flags & B200Q_FLAG_FORCE_GEMV) return false;
    if (a->flags & B200Q_FLAG_FORCE_MMQ) return mmq_supported(a->type, a->k);
    // from this many columns on the tensor-core kernel (dequantize to bf16 + tcgen05: ~flat in n below 64 columns) beats
    // repeated passes of the 8-column integer-MMA GEMV (gemv_mma.cu): two passes win, three are a draw, four lose
    // (measured cross-over, profiles/r01_cols_crossover.txt)
    static const int64_t min_n = getenv("B200Q_MMQ_MIN_N") ? atoll(getenv("B200Q_MMQ_MIN_N")) : 17;
    if (a->n >= min_n) return mmq_supported(a->type, a->k);
    if (a->n > 8 && !(mma_kernel_for(a->type) && a->ne12 * a->ne13 == 1 && a->nb01 == b200q_row_size(a->type, a->k) &&
                      a->nb01 % 16 == 0 && a->k % (blck_elems(a->type) == 256 ? 256 : 32) == 0)) {
        return mmq_supported(a->type, a->k);  // no multi-pass GEMV for this case
    }
    return false;
}

// MXFP4 on the native block-scaled tensor-core path (mmq_mxf.cu): plain 2-D products
static long long k_mxf_dual_units(int64_t m, int64_t n) { return ((m + 127) / 128) * ((n + 511) / 512); }
static bool use_mxf(const b200q_mul_mat_args * a) {
    if (a->type != T_MXFP4 || a->ne02 * a->ne03 * a->ne12 * a->ne13 != 1 || !mxf_supported(a->type, a->k) || (a->flags & B200Q_FLAG_FORCE_GEMV)) return false;
    if (!(a->flags & B200Q_FLAG_FORCE_MMQ) && a->n < 17) return false;
    // measured (profiles/r02_mxf_vs_bf16_v1.txt): the native kernel wins wherever the bf16 kernel is dequantization-bound
    // (1.3 - 1.8 x: few row tiles, few columns, MoE); where the bf16 kernel runs its two-accumulator units on a full GPU
    // (n > 256 with >= 88 row tiles, e.g. 11520 x 2880 at n = 512: 55 vs 66 us) it stays ahead
    static const bool always = getenv("B200Q_MXF_ALWAYS") != nullptr;
    if (!always && a->n > 256 && k_mxf_dual_units(a->m, a->n) >= 88 && a->k % 64 == 0) return false;
    return true;
}

size_t b200q_mul_mat_workspace(const b200q_mul_mat_args * a) {
    if (!a || check_mm(a)) return 0;
    if (use_mxf(a)) return mxf_workspace_bytes(a->k, a->n);
    if (!use_mmq(a)) return 0;
    return mmq_workspace_bytes(a->k, a->n, a->ne12 * a->ne13);
}

int b200q_mul_mat(const b200q_mul_mat_args * a, const void * w, const float * act, float * dst, void * workspace,
                  size_t workspace_bytes, void * stream) {
    if (int rc = check_mm(a)) return rc;
    if (!w || !act || !dst) return fail(B200Q_EINVAL, "mul_mat: null pointer");
    if (a->m == 0 || a->n == 0) return 0;
    cudaStream_t    st = (cudaStream_t) stream;
    const DevInfo & di = dev_info();
    if (!di.ok) return fail(B200Q_ENODEVICE, "mul_mat: no CUDA device");

    if (use_dense(a)) {
        if (!dense_mm_supported(wtype(a->type), a->k, a->m, a->n, a->ne12 * a->ne13)) return fail(B200Q_ENOTSUP, "mul_mat: dense shape out of range");
        MmqArgs q{};
        q.type = wtype(a->type); q.k = a->k; q.m = a->m; q.n = a->n;
        q.ne02 = a->ne02; q.ne03 = a->ne03; q.ne12 = a->ne12; q.ne13 = a->ne13;
        q.w_nb1 = (long long) a->nb01; q.w_nb2 = (long long) a->nb02; q.w_nb3 = (long long) a->nb03;
        q.a_nb1 = (long long) a->nb11; q.a_nb2 = (long long) a->nb12; q.a_nb3 = (long long) a->nb13;
        q.d_nb1 = (long long) a->nb1;  q.d_nb2 = (long long) a->nb2;  q.d_nb3 = (long long) a->nb3;
        q.w = (const uint8_t *) w; q.act = act; q.dst = dst;
        if (int rc = dense_mm_launch(q, st)) return fail(rc, "mul_mat: dense launch failed: %s", cudaGetErrorString(cudaGetLastError()));
        return 0;
    }
    if (use_mxf(a)) {
        const size_t need = mxf_workspace_bytes(a->k, a->n);
        if (workspace_bytes < need || !workspace) return fail(B200Q_EWORKSPACE, "mul_mat: workspace %zu < required %zu", workspace_bytes, need);
        MxfArgs q{};
        q.k = a->k; q.m = a->m; q.n = a->n; q.w = (const uint8_t *) w; q.act = act; q.dst = dst;
        q.w_nb1 = (long long) a->nb01; q.a_nb1 = (long long) a->nb11; q.d_nb1 = (long long) a->nb1;
        std::string err;
        mmq_forget_pack();
        if (int rc = mxf_launch(q, workspace, st, di.sms, err)) return fail(rc, "%s", err.c_str());
        return 0;
    }
    if (use_mmq(a)) {
        MmqArgs q{};
        q.type = wtype(a->type); q.k = a->k; q.m = a->m; q.n = a->n;
        q.ne02 = a->ne02; q.ne03 = a->ne03; q.ne12 = a->ne12; q.ne13 = a->ne13;
        q.w_nb1 = (long long) a->nb01; q.w_nb2 = (long long) a->nb02; q.w_nb3 = (long long) a->nb03;
        q.a_nb1 = (long long) a->nb11; q.a_nb2 = (long long) a->nb12; q.a_nb3 = (long long) a->nb13;
        q.d_nb1 = (long long) a->nb1;  q.d_nb2 = (long long) a->nb2;  q.d_nb3 = (long long) a->nb3;
        q.w = (const uint8_t *) w; q.act = act; q.dst = dst;
        q.reuse_pack = (a->flags & B200Q_FLAG_ACT_UNCHANGED) != 0;
        const size_t need = mmq_workspace_bytes(a->k, a->n, a->ne12 * a->ne13);
        if (workspace_bytes < need || !workspace) {
            return fail(B200Q_EWORKSPACE, "mul_mat: workspace %zu < required %zu", workspace_bytes, need);
        }
        std::string err;
        const int   rc = mmq_launch(q, workspace, st, di.sms, err);
        if (rc) return fail(rc, "%s", err.c_str());
        return 0;
    }

    // decode fast path: one activation column, K-quant weights, dense rows
    if (a->n == 1 && a->ne12 * a->ne13 == 1 && a->nb01 == b200q_row_size(a->type, a->k)) {
        KqParams kq{};
        int      grid = 0, smem = 0;
        const void * wl[1] = { w };
        float *      dl[1] = { dst };
        const int64_t ml[1] = { a->m };
        if (plan_kq(kq, grid, smem, a->type, a->k, 1, wl, dl, ml, act, di)) {
            return launch_kq(kq, grid, smem, a->type, a->flags, st);
        }
    }

    // 2..8 columns on K-quants: the MMA tile kernel (one pass over the weights; the columns are halved only when their
    // quantized images do not fit beside the ring in shared memory)
    if (a->n >= 2 && a->ne12 * a->ne13 == 1 && a->nb01 == b200q_row_size(a->type, a->k) && mma_kernel_for(a->type)) {
        MmaParams mp{};
        int       grid = 0, smem = 0;
        int       nc   = (int) std::min<int64_t>(a->n, 8);
        auto      col_act = [&](int64_t j) { return (const float *) ((const char *) act + j * (long long) a->nb11); };
        auto      col_dst = [&](int64_t j) { return (float *) ((char *) dst + j * (long long) a->nb1); };
        while (nc >= 2 && !plan_mma(mp, grid, smem, a->type, a->m, a->k, nc, w, act, (long long) a->nb11, dst, (long long) a->nb1, di)) {
            nc = (nc + 1) / 2;
        }
        // a last group of one column goes to the n = 1 kernel
        KqParams     kq{};
        int          kgrid = 0, ksmem = 0;
        const bool   tail1 = a->n % nc == 1;
        const void * wl[1] = { w };
        float *      dl[1] = { col_dst(a->n - 1) };
        const int64_t ml[1] = { a->m };
        if (nc >= 2 && (!tail1 || plan_kq(kq, kgrid, ksmem, a->type, a->k, 1, wl, dl, ml, col_act(a->n - 1), di))) {
            for (int64_t j0 = 0; j0 < a->n; j0 += nc) {
                const int cur = (int) std::min<int64_t>(nc, a->n - j0);
                if (cur == 1) {
                    if (int rc = launch_kq(kq, kgrid, ksmem, a->type, a->flags, st)) return rc;
                    break;
                }
                if (!plan_mma(mp, grid, smem, a->type, a->m, a->k, cur, w, col_act(j0), (long long) a->nb11, col_dst(j0), (long long) a->nb1, di)) {
                    return fail(B200Q_ENOTSUP, "mul_mat: gemv_mma plan failed");
                }
                if (int rc = launch_mma(mp, grid, smem, a->type, a->flags, st)) return rc;
            }
            return 0;
        }
    }

    // GEMV path: up to 8 activation columns per launch (more columns = several passes over the weights)
    GemvPlan plan{};
    GemvParams & p = plan.p;
    p.w            = (const uint8_t *) w;
    p.ids          = nullptr;
    p.k            = (int) a->k;
    p.m            = (int) a->m;
    p.nch          = (int) (a->ne12 * a->ne13);
    p.mode         = GEMV_MODE_BATCH;
    p.ne12         = (int) a->ne12;
    p.r2           = (int) (a->ne12 / a->ne02);
    p.r3           = (int) (a->ne13 / a->ne03);
    p.n_used = 1; p.n_expert = 1; p.b_bcast = 0; p.ids_nb1 = 0;
    p.w_nb1 = (long long) a->nb01; p.w_nb2 = (long long) a->nb02; p.w_nb3 = (long long) a->nb03;
    p.a_nb1 = (long long) a->nb11; p.a_nb2 = (long long) a->nb12; p.a_nb3 = (long long) a->nb13;
    p.d_nb1 = (long long) a->nb1;  p.d_nb2 = (long long) a->nb2;  p.d_nb3 = (long long) a->nb3;
    p.act_vec4 = aligned16(act, p.a_nb1) && (p.a_nb2 & 15) == 0 && (p.a_nb3 & 15) == 0;

    int ncols = (int) std::min<int64_t>(a->n, 8);
    while (ncols >= 1 && !plan_gemv(plan, a->type, ncols, di)) {
        ncols = ncols > 1 ? (ncols + 1) / 2 : 0;
    }
    if (ncols < 1) return fail(B200Q_ENOTSUP, "mul_mat: k=%lld too large for the GEMV shared-memory plan", (long long) a->k);
    for (int64_t j0 = 0; j0 < a->n; j0 += ncols) {
        const int nc = (int) std::min<int64_t>(ncols, a->n - j0);
        if (nc != ncols && !plan_gemv(plan, a->type, nc, di)) return fail(B200Q_ENOTSUP, "mul_mat: plan failed");
        p.act = (const float *) ((const char *) act + j0 * (long long) a->nb11);
        p.dst = (float *) ((char *) dst + j0 * (long long) a->nb1);
        if (int rc = launch_gemv(plan, a->type, nc, a->flags, st)) return rc;
    }
    return 0;
}

// ---- MUL_MAT without any allocation (b200q_mul_mat_static, include/b200q.h) --------------------------------------------
// For callers that cannot allocate (CUDA-graph capture, serving loops, harnesses that forbid cudaMalloc): the tensor-core
// kernel's scratch (the packed bf16 activation image) lives in this file-scope device array instead of a caller buffer.
// Every GPU that loads the library holds B200Q_STATIC_WORKSPACE_BYTES until exit.  The array's name is only a host-side
// stand-in: host code gets the device address from cudaGetSymbolAddress, once per device, with that device current.
static __device__ __align__(256) unsigned char g_static_workspace[B200Q_STATIC_WORKSPACE_BYTES];

static int static_workspace_ptr(void ** out) {
    static std::mutex mu;
    static void *     ptr[64] = {};
    int               dev     = -1;
    CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return fail(B200Q_EINVAL, "mul_mat_static: device ordinal %d out of range", dev);
    std::lock_guard<std::mutex> lk(mu);
    if (!ptr[dev]) {
        void * p = nullptr;
        CUDA_TRY(cudaGetSymbolAddress(&p, g_static_workspace));
        if (!p) return fail(B200Q_ECUDA, "mul_mat_static: cudaGetSymbolAddress returned a null address");
        ptr[dev] = p;
    }
    *out = ptr[dev];
    return 0;
}

int b200q_mul_mat_static(const b200q_mul_mat_args * a, const void * w, const float * act, float * dst, void * stream) {
    if (int rc = check_mm(a)) return rc;
    const size_t need = b200q_mul_mat_workspace(a);
    if (need > (size_t) B200Q_STATIC_WORKSPACE_BYTES) {
        return fail(B200Q_EWORKSPACE, "mul_mat_static: this shape needs %zu bytes of scratch, the static workspace holds %zu "
                    "(B200Q_STATIC_WORKSPACE_BYTES); use b200q_mul_mat with a caller workspace", need,
                    (size_t) B200Q_STATIC_WORKSPACE_BYTES);
    }
    void * ws = nullptr;
    if (int rc = static_workspace_ptr(&ws)) return rc;
    // nothing in the scratch has to start at zero: mmq_pack_act rewrites every byte the kernel reads (padding tiles included)
    return b200q_mul_mat(a, w, act, dst, ws, (size_t) B200Q_STATIC_WORKSPACE_BYTES, stream);
}

// ---- grouped decode GEMV: matrices that share one activation vector (q/k/v, gate/up) in one launch ------------------
static int mul_mat_group_impl(const b200q_mul_mat_group_args * a, const void * const * w, const float * act, float * const * dst,
                              int n_peers, const int64_t * peer_delta, void * stream);

int b200q_mul_mat_group(const b200q_mul_mat_group_args * a, const void * const * w, const float * act, float * const * dst,
                        void * stream) {
    return mul_mat_group_impl(a, w, act, dst, 0, nullptr, stream);
}

int b200q_mul_mat_group_peers(const b200q_mul_mat_group_args * a, const void * const * w, const float * act, float * const * dst,
                              int n_peers, const int64_t * peer_delta, void * stream) {
    if (n_peers < 0 || n_peers > 7 || (n_peers && !peer_delta)) return fail(B200Q_EINVAL, "mul_mat_group_peers: 0..7 peers");
    return mul_mat_group_impl(a, w, act, dst, n_peers, peer_delta, stream);
}

// ---- inter-process device memory (row split: one process per GPU) -------------------------------------------------
int b200q_ipc_export(void * dev_ptr, void * handle64) {
    if (!dev_ptr || !handle64) return fail(B200Q_EINVAL, "ipc_export: null argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
    CUDA_TRY(cudaIpcGetMemHandle((cudaIpcMemHandle_t *) handle64, dev_ptr));
    return 0;
}
int b200q_ipc_import(const void * handle64, void ** dev_ptr) {
    if (!dev_ptr || !handle64) return fail(B200Q_EINVAL, "ipc_import: null argument");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, sizeof(h));
    CUDA_TRY(cudaIpcOpenMemHandle(dev_ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return 0;
}
int b200q_ipc_close(void * dev_ptr) {
    if (!dev_ptr) return 0;
    CUDA_TRY(cudaIpcCloseMemHandle(dev_ptr));
    return 0;
}

static int mul_mat_group_impl(const b200q_mul_mat_group_args * a, const void * const * w, const float * act, float * const * dst,
                              int n_peers, const int64_t * peer_delta, void * stream) {
    if (!a || !w || !act || !dst) return fail(B200Q_EINVAL, "mul_mat_group: null argument");
    if (a->n_mats < 1 || a->n_mats > B200Q_MAX_GROUP) return fail(B200Q_EINVAL, "mul_mat_group: n_mats must be 1..%d", B200Q_MAX_GROUP);
    const int be = blck_elems(a->type);
    if (!be || a->type == T_F32) return fail(B200Q_ENOTSUP, "mul_mat_group: unsupported weight type %d", a->type);
    if (a->k <= 0 || a->k % be) return fail(B200Q_EINVAL, "mul_mat_group: bad k");
    const DevInfo & di = dev_info();
    if (!di.ok) return fail(B200Q_ENODEVICE, "mul_mat_group: no CUDA device");
    for (int g = 0; g < a->n_mats; ++g) {
        if (!w[g] || !dst[g] || a->m[g] < 0) return fail(B200Q_EINVAL, "mul_mat_group: bad matrix %d", g);
    }
    KqParams kq{};
    int      grid = 0, smem = 0;
    if (plan_kq(kq, grid, smem, a->type, a->k, a->n_mats, w, dst, a->m, act, di)) {
        kq.n_peers = n_peers;
        for (int r = 0; r < n_peers; ++r) kq.peer_delta[r] = (long long) peer_delta[r];
        return launch_kq(kq, grid, smem, a->type, a->flags, (cudaStream_t) stream);
    }
    if (n_peers) return fail(B200Q_ENOTSUP, "mul_mat_group_peers: shape / type not handled by the ring kernel");
    // anything the grouped kernel does not take: one launch per matrix (same results)
    for (int g = 0; g < a->n_mats; ++g) {
        b200q_mul_mat_args q{};
        q.type = a->type; q.k = a->k; q.m = a->m[g]; q.n = 1;
        q.ne02 = q.ne03 = q.ne12 = q.ne13 = 1;
        q.nb01 = b200q_row_size(a->type, a->k); q.nb02 = q.nb03 = q.nb01 * (size_t) a->m[g];
        q.nb11 = q.nb12 = q.nb13 = (size_t) a->k * 4;
        q.nb1 = q.nb2 = q.nb3 = (size_t) a->m[g] * 4;
        q.flags = a->flags;
        if (int rc = b200q_mul_mat(&q, w[g], act, dst[g], nullptr, 0, stream)) return rc;
    }
    return 0;
}

// ---- MUL_MAT_ID ----------------------------------------------------------------------------------------
// (token, slot) rows from which MUL_MAT_ID groups the rows per expert and runs the tensor-core kernel; below it every
// row is one GEMV channel
static bool moe_use_mmq(const b200q_mul_mat_id_args * a) {
    static const long long min_rows = getenv("B200Q_MOE_MMQ_MIN") ? atoll(getenv("B200Q_MOE_MMQ_MIN")) : 32;
    if (dense_type(wtype(a->type))) return mmq_supported(wtype(a->type), a->k) && a->n_expert <= 1024;
    return a->n_tokens * a->n_used >= min_rows && mmq_supported(a->type, a->k) && a->n_expert <= 1024;
}

static bool moe_use_mxf(const b200q_mul_mat_id_args * a) {
    static const long long min_rows = getenv("B200Q_MOE_MMQ_MIN") ? atoll(getenv("B200Q_MOE_MMQ_MIN")) : 32;
    return a->type == T_MXFP4 && mxf_supported(a->type, a->k) && a->n_tokens * a->n_used >= min_rows && a->n_expert <= 1024;
}

size_t b200q_mul_mat_id_workspace(const b200q_mul_mat_id_args * a) {
    if (a && moe_use_mxf(a)) return mxf_moe_workspace_bytes(a->k, a->n_tokens, a->n_used, a->n_expert);
    if (!a || !moe_use_mmq(a)) return 0;
    return mmq_moe_workspace_bytes(a->k, a->n_tokens, a->n_used, a->n_expert);
}

// b200q_mul_mat_id_bias hands the bias to the decode branch of b200q_mul_mat_id through these (same thread, same call)
static thread_local const float * g_moe_bias        = nullptr;
static thread_local long long     g_moe_bias_stride = 0;

int b200q_mul_mat_id(const b200q_mul_mat_id_args * a, const void * as, const float * b, const int32_t * ids, float * dst,
                     void * workspace, size_t workspace_bytes, void * stream) {
    if (!a || !as || !b || !ids || !dst) return fail(B200Q_EINVAL, "mul_mat_id: null argument");
    const int be = blck_elems(wtype(a->type));
    if (!be || a->type == T_F32W) return fail(B200Q_ENOTSUP, "mul_mat_id: unsupported weight type %d", a->type);
    if (a->k <= 0 || a->k % be) return fail(B200Q_EINVAL, "mul_mat_id: bad k");
    if (dense_type(wtype(a->type)) && !moe_use_mmq(a)) {
        return fail(B200Q_ENOTSUP, "mul_mat_id: f16 / bf16 / f32 experts need k %% 64 == 0 and <= 1024 experts");
    }
    if (a->b_ne1 != 1 && a->b_ne1 != a->n_used) return fail(B200Q_EINVAL, "mul_mat_id: b->ne[1] must be 1 or n_used");
    if (a->m == 0 || a->n_tokens == 0 || a->n_used == 0) return 0;
    const DevInfo & di = dev_info();
    if (!di.ok) return fail(B200Q_ENODEVICE, "mul_mat_id: no CUDA device");

    if (moe_use_mxf(a)) {
        const size_t need = mxf_moe_workspace_bytes(a->k, a->n_tokens, a->n_used, a->n_expert);
        if (!workspace || workspace_bytes < need) return fail(B200Q_EINVAL, "mul_mat_id: workspace of %zu bytes needed (b200q_mul_mat_id_workspace)", need);
        MxfArgs q{};
        q.k = a->k; q.m = a->m; q.w = (const uint8_t *) as; q.act = b; q.dst = dst;
        q.w_nb1 = (long long) a->nb01; q.w_nb2 = (long long) a->nb02;
        q.a_nb1 = (long long) a->nb11; q.a_nb2 = (long long) a->nb12;
        q.d_nb1 = (long long) a->nb1;  q.d_nb2 = (long long) a->nb2;
        q.ids = ids; q.ids_nb1 = (long long) a->ids_nb1; q.n_expert = a->n_expert; q.n_used = a->n_used; q.n_tokens = a->n_tokens;
        q.b_bcast = a->b_ne1 == 1;
        std::string err;
        mmq_forget_pack();
        if (int rc = mxf_launch(q, workspace, (cudaStream_t) stream, di.sms, err)) return fail(rc, "%s", err.c_str());
        return 0;
    }
    if (moe_use_mmq(a)) {
        const size_t need = mmq_moe_workspace_bytes(a->k, a->n_tokens, a->n_used, a->n_expert);
        if (!workspace || workspace_bytes < need) {
            return fail(B200Q_EINVAL, "mul_mat_id: workspace of %zu bytes needed (b200q_mul_mat_id_workspace)", need);
        }
        MmqMoeArgs q{};
        q.type = wtype(a->type); q.k = a->k; q.m = a->m; q.n_expert = a->n_expert; q.n_used = a->n_used; q.n_tokens = a->n_tokens;
        q.b_bcast = a->b_ne1 == 1;
        q.w_nb1 = (long long) a->nb01; q.w_nb2 = (long long) a->nb02;
        q.a_nb1 = (long long) a->nb11; q.a_nb2 = (long long) a->nb12;
        q.ids_nb1 = (long long) a->ids_nb1;
        q.d_nb1 = (long long) a->nb1; q.d_nb2 = (long long) a->nb2;
        q.w = (const uint8_t *) as; q.act = b; q.ids = ids; q.dst = dst;
        std::string err;
        if (int rc = mmq_moe_launch(q, workspace, (cudaStream_t) stream, di.sms, err)) return fail(rc, "%s", err.c_str());
        return 0;
    }

    // decode (one token): the selected experts are the segments of ONE launch of the ring kernel
    static const bool no_kq_moe = getenv("B200Q_NO_KQ_MOE") != nullptr;
    if (a->n_tokens == 1 && a->n_used <= KQ_MAX_SEG && a->nb01 == b200q_row_size(a->type, a->k) && !no_kq_moe &&
        (a->b_ne1 == 1 || a->nb11 % 4 == 0) && a->nb1 % 4 == 0) {
        KqParams     kq{};
        int          grid = 0, smem = 0;
        const void * wl[KQ_MAX_SEG];
        float *      dl[KQ_MAX_SEG];
        int64_t      ml[KQ_MAX_SEG];
        for (int g = 0; g < a->n_used; ++g) {
            wl[g] = as;  // expert 0; the kernel adds ids[g] * nb02
            dl[g] = (float *) ((char *) dst + (size_t) g * a->nb1);
            ml[g] = a->m;
        }
        if (plan_kq(kq, grid, smem, a->type, a->k, (int) a->n_used, wl, dl, ml, b, di, ids, (long long) a->nb02,
                    a->b_ne1 == 1 ? 0 : (long long) (a->nb11 / 4))) {
            kq.bias = g_moe_bias; kq.bias_stride = g_moe_bias_stride;
            return launch_kq(kq, grid, smem, a->type, a->flags, (cudaStream_t) stream);
        }
    }
    if (g_moe_bias) return fail(B200Q_ENOTSUP, "mul_mat_id_bias: only the one-token ring-kernel path adds the bias");

    // one GEMV channel per (token, slot); expert ids are read on the device (no host round trip, cf. mmq.cu:161)
    GemvPlan plan{};
    GemvParams & p = plan.p;
    p.w = (const uint8_t *) as; p.act = b; p.dst = dst; p.ids = ids;
    p.k = (int) a->k; p.m = (int) a->m;
    p.nch  = (int) (a->n_tokens * a->n_used);
    p.mode = GEMV_MODE_MOE;
    p.ne12 = 1; p.r2 = 1; p.r3 = 1;
    p.n_used = (int) a->n_used; p.n_expert = (int) a->n_expert; p.b_bcast = a->b_ne1 == 1;
    p.w_nb1 = (long long) a->nb01; p.w_nb2 = (long long) a->nb02; p.w_nb3 = 0;
    p.a_nb1 = (long long) a->nb11; p.a_nb2 = (long long) a->nb12; p.a_nb3 = 0;
    p.d_nb1 = (long long) a->nb1;  p.d_nb2 = (long long) a->nb2;  p.d_nb3 = 0;
    p.ids_nb1  = (long long) a->ids_nb1;
    p.act_vec4 = aligned16(b, p.a_nb1) && (p.a_nb2 & 15) == 0;
    if (!plan_gemv(plan, a->type, 1, di)) return fail(B200Q_ENOTSUP, "mul_mat_id: k too large");
    return launch_gemv(plan, a->type, 1, a->flags, (cudaStream_t) stream);
}

// MUL_MAT_ID followed by ADD_ID (the experts' bias rows) in one launch: decode only (one token, the ring kernel's epilogue adds
// bias[ids[g]][row]); B200Q_ENOTSUP otherwise — the caller then runs the two operators
int b200q_mul_mat_id_bias(const b200q_mul_mat_id_args * a, const void * as, const float * b, const int32_t * ids, const float * bias,
                          size_t bias_nb1, float * dst, void * workspace, size_t workspace_bytes, void * stream) {
    if (!a || !bias || bias_nb1 % 4) return fail(B200Q_EINVAL, "mul_mat_id_bias: null / unaligned bias");
    if (a->n_tokens != 1 || moe_use_mxf(a) || moe_use_mmq(a)) return fail(B200Q_ENOTSUP, "mul_mat_id_bias: one token only");
    g_moe_bias = bias; g_moe_bias_stride = (long long) (bias_nb1 / 4);
    const int rc = b200q_mul_mat_id(a, as, b, ids, dst, workspace, workspace_bytes, stream);
    g_moe_bias = nullptr; g_moe_bias_stride = 0;
    return rc;
}

// ---- persistent decode list --------------------------------------------------------------------------------------------
}  // extern "C" (reopened below)

struct b200q_list {
    int                      device = 0;
    std::vector<ListSegment> segs;
    void *                   dev_tables = nullptr;
    unsigned *               own_sync   = nullptr;  // allocated here for single-GPU lists
    int                      n_bp = 0, n_gens = 0;
};

namespace {
// host-side plan of a list on a device with `sms` SMs and `max_smem` bytes of shared memory per CTA (no CUDA calls:
// tests/test_plan_cpu.py runs it without a GPU); table pointers are still unset
int list_plan(const b200q_list_op * ops, int n_ops, int world, int sms, int max_smem, std::vector<ListSegment> & segs) {
    if (!ops || n_ops < 1) return fail(B200Q_EINVAL, "list: no ops");
    segs.clear();
    ListSegment  cur;
    long long    cur_stride = 0;
    long long    units      = 0;
    const float * prev_act  = nullptr, * prev_act2 = nullptr;
    int           prev_q    = -1;
    std::vector<long long> op_units;  // units of every op of the current launch
    auto close = [&]() {
        ListParams & p = cur.p;
        p.n_ops       = (int) cur.ops.size();
        p.n_gens      = (int) cur.gens.size();
        p.nslots      = list_smem_plan(p.n_ops, p.n_gens, cur_stride, max_smem, p);
        p.final_bp    = world > 1 ? 1 : 0;
        int nbp = 0;
        for (auto & g : cur.gens) if (g.wait) g.bp = (uint16_t) ++nbp;
        p.n_bp   = nbp + p.final_bp;
        cur.smem = p.off_ring + p.nslots * LIST_SLOT_BYTES;
        // row split: every GPU counts the same number of CTAs in at a barrier point, whatever its share of the rows is
        cur.grid = world > 1 ? sms : (int) std::max<long long>(1, std::min<long long>(units, sms));
        cur.total_units = units;
        // every op's units in `grid` contiguous ranges; the surplus ranges rotate on from op to op (gemv_list.cuh)
        long long rot = 0;
        for (size_t j = 0; j < cur.ops.size(); ++j) {
            cur.ops[j].base = (int32_t) (op_units[j] / cur.grid);
            cur.ops[j].rem  = (int32_t) (op_units[j] % cur.grid);
            cur.ops[j].rot  = (int32_t) rot;
            rot = (rot + cur.grid - cur.ops[j].rem) % cur.grid;  // c' = (c + rot) % G: the next op's surplus starts where this one's ended
        }
        segs.push_back(std::move(cur));
        cur = ListSegment{};
        op_units.clear();
        cur_stride = 0; units = 0; prev_act = prev_act2 = nullptr; prev_q = -1;
    };
    for (int i = 0; i < n_ops; ++i) {
        const b200q_list_op & o = ops[i];
        ListGeo g{};
        if (!list_geometry(o.type, o.k, g)) return fail(B200Q_ENOTSUP, "list: op %d: type %d / k %lld not served", i, o.type, (long long) o.k);
        if (o.m < 0 || o.m > (1 << 30) || !o.act || ((uintptr_t) o.act & 15) || ((uintptr_t) o.w & 15) || (o.m && (!o.w || !o.dst))) {
            return fail(B200Q_EINVAL, "list: op %d: bad pointer / alignment / m", i);
        }
        const float * act2 = (o.flags & B200Q_LIST_SWIGLU) ? o.act2 : nullptr;
        if ((o.flags & B200Q_LIST_SWIGLU) && (!act2 || ((uintptr_t) act2 & 15))) return fail(B200Q_EINVAL, "list: op %d: B200Q_LIST_SWIGLU needs an aligned act2", i);
        const bool new_gen = o.act != prev_act || act2 != prev_act2 || g.qmode != prev_q || (o.flags & B200Q_LIST_DEP) || cur.gens.empty();
        // does the op still fit this launch (tables + two activation images + >= LIST_MIN_SLOTS ring slots)?
        const long long stride = std::max<long long>(cur_stride, round_up((int) g.image_bytes, 128));
        ListParams      probe{};
        int slots = list_smem_plan((int) cur.ops.size() + 1, (int) cur.gens.size() + (new_gen ? 1 : 0), stride, max_smem, probe);  // with this op added
        const bool full = cur.gens.size() + 1 >= 65535 || units + o.m >= (1ll << 30);
        (void) probe;
        if (!cur.ops.empty() && (slots < LIST_MIN_SLOTS || full)) {
            close();
            slots = list_smem_plan(1, 1, round_up((int) g.image_bytes, 128), max_smem, probe);
        }
        if (slots < LIST_MIN_SLOTS) return fail(B200Q_ENOTSUP, "list: op %d: k = %lld leaves %d ring slots", i, (long long) o.k, slots);
        const bool first_of_launch = cur.ops.empty();
        if (new_gen || first_of_launch) {
            ListGen gg{};
            gg.act   = o.act;
            gg.nsb   = (uint16_t) g.nsb;
            gg.qmode = (uint8_t) g.qmode;
            gg.xform = act2 ? 1 : 0;
            gg.act2  = act2;
            // the first generation of a launch follows a kernel boundary (and, in a row split, the previous launch's final
            // barrier point): no wait needed
            gg.wait  = (o.flags & B200Q_LIST_DEP) && !first_of_launch ? 1 : 0;
            cur.gens.push_back(gg);
        }
        cur_stride = std::max<long long>(cur_stride, round_up((int) g.image_bytes, 128));
        ListOp r{};
        r.w = (const uint8_t *) o.w; r.dst = o.dst; r.m = (int32_t) o.m;
        r.rb = (uint32_t) ((long long) g.nsb * g.bb); r.bb = (uint8_t) g.bb;
        r.nsb = (uint16_t) g.nsb; r.csb = (uint16_t) g.csb; r.gen = (uint16_t) (cur.gens.size() - 1);
        r.type = (uint8_t) o.type; r.geo = (uint8_t) (g.rps_log2 | ((g.nch - 1) << 4));
        cur.ops.push_back(r);
        op_units.push_back((o.m + (1ll << g.rps_log2) - 1) >> g.rps_log2);
        units += op_units.back();
        prev_act = o.act; prev_act2 = act2; prev_q = g.qmode;
    }
    close();
    return 0;
}
}  // namespace

extern "C" {

// host-side test hook (tests/test_plan_cpu.py): plan a list of n_ops (type, m, k) triples that all read distinct vectors
// (dep[i] != 0: B200Q_LIST_DEP) on a device with `sms` SMs / `max_smem` bytes; out[8 * n_ops] per op = launch, unit0,
// rps, nch, csb, gen, nsb, 0; seg_out[8 * launches] = n_ops, n_gens, total_units, n_bp, nslots, smem, grid, act_stride.
// Returns the number of launches (<= max_segs), or a negative status.
B200Q_API int b200q_debug_plan_list(const int * types, const int64_t * m, const int64_t * k, const int * dep, const int * same_act,
                                    int n_ops, int world, int sms, int max_smem, int * out, int * seg_out, int max_segs) {
    std::vector<b200q_list_op> ops(n_ops);
    alignas(16) static float dummy_act[1 << 12];
    alignas(16) static uint8_t dummy_w[16];
    for (int i = 0; i < n_ops; ++i) {
        ops[i].type = types[i]; ops[i].m = m[i]; ops[i].k = k[i]; ops[i].w = dummy_w;
        ops[i].act = i > 0 && same_act && same_act[i] ? ops[i - 1].act : dummy_act + 4 * (i % 1024);
        ops[i].dst = dummy_act; ops[i].flags = dep && dep[i] ? B200Q_LIST_DEP : 0;
    }
    std::vector<ListSegment> segs;
    if (int rc = list_plan(ops.data(), n_ops, world, sms, max_smem, segs)) return rc;
    if ((int) segs.size() > max_segs) return B200Q_EINVAL;
    int io = 0;
    for (size_t si = 0; si < segs.size(); ++si) {
        const ListSegment & sg = segs[si];
        for (int j = 0; j < sg.p.n_ops; ++j, ++io) {  // per op: launch, units, rps, nch, csb, gen, nsb, rot
            const ListOp & r = sg.ops[j];
            const int v[8] = { (int) si, r.base * sg.grid + r.rem, 1 << (r.geo & 15), (r.geo >> 4) + 1, r.csb, r.gen, r.nsb, r.rot };
            for (int q = 0; q < 8; ++q) out[8 * io + q] = v[q];
        }
        const int v[8] = { sg.p.n_ops, sg.p.n_gens, (int) sg.total_units, sg.p.n_bp, sg.p.nslots, sg.smem, sg.grid, sg.p.act_stride };
        for (int q = 0; q < 8; ++q) seg_out[8 * si + q] = v[q];
    }
    return (int) segs.size();
}

int b200q_list_create(const b200q_list_op * ops, int n_ops, const b200q_list_peers * peers, b200q_list ** out) {
    if (!out) return fail(B200Q_EINVAL, "list_create: null out");
    *out = nullptr;
    const DevInfo & di = dev_info();
    if (!di.ok) return fail(B200Q_ENODEVICE, "list_create: no CUDA device");
    const int world = peers ? peers->world : 1;
    if (peers && (world < 1 || world > 8 || peers->rank < 0 || peers->rank >= world || !peers->sync || (world > 1 && !peers->peer_delta))) {
        return fail(B200Q_EINVAL, "list_create: bad peers (1..8 ranks)");
    }
    auto * L = new b200q_list;
    cudaGetDevice(&L->device);
    if (int rc = list_plan(ops, n_ops, world, di.sms, di.max_smem, L->segs)) { delete L; return rc; }
    size_t bytes = 0;
    for (auto & sg : L->segs) bytes += sg.ops.size() * sizeof(ListOp) + sg.gens.size() * sizeof(ListGen);
    auto cleanup = [&](int rc) { if (L->dev_tables) cudaFree(L->dev_tables); if (L->own_sync) cudaFree(L->own_sync); delete L; return rc; };
    if (cudaMalloc(&L->dev_tables, bytes) != cudaSuccess) { cudaGetLastError(); return cleanup(fail(B200Q_ENOMEM, "list_create: cudaMalloc(%zu)", bytes)); }
    unsigned * sync = nullptr;
    if (world > 1) {
        sync = (unsigned *) peers->sync[peers->rank];
    } else {
        if (cudaMalloc((void **) &L->own_sync, 256) != cudaSuccess || cudaMemset(L->own_sync, 0, 256) != cudaSuccess) {
            cudaGetLastError();
            return cleanup(fail(B200Q_ENOMEM, "list_create: sync block"));
        }
        sync = L->own_sync;
    }
    size_t off = 0;
    for (auto & sg : L->segs) {
        ListParams & p = sg.p;
        p.ops = (const ListOp *) ((char *) L->dev_tables + off);
        if (cudaMemcpy((void *) p.ops, sg.ops.data(), sg.ops.size() * sizeof(ListOp), cudaMemcpyHostToDevice) != cudaSuccess) return cleanup(fail(B200Q_ECUDA, "list_create: table upload"));
        off += sg.ops.size() * sizeof(ListOp);
        p.gens = (const ListGen *) ((char *) L->dev_tables + off);
        if (cudaMemcpy((void *) p.gens, sg.gens.data(), sg.gens.size() * sizeof(ListGen), cudaMemcpyHostToDevice) != cudaSuccess) return cleanup(fail(B200Q_ECUDA, "list_create: table upload"));
        off += sg.gens.size() * sizeof(ListGen);
        p.sync  = sync;
        p.world = world;
        p.rank  = peers ? peers->rank : 0;
        p.n_peers = world - 1;
        for (int r = 0, q = 0; r < world; ++r) {
            if (r == p.rank) continue;
            p.peer_sync[q]  = (unsigned *) peers->sync[r];
            p.peer_delta[q] = (long long) peers->peer_delta[q];
            ++q;
        }
        L->n_bp += p.n_bp;
        L->n_gens += p.n_gens;
    }
    *out = L;
    return 0;
}

int b200q_list_launch(b200q_list * L, void * stream) {
    if (!L) return fail(B200Q_EINVAL, "list_launch: null list");
    const DevInfo & di = dev_info();
    static const bool debug = getenv("B200Q_DEBUG") != nullptr;
    for (auto & sg : L->segs) {
        if (sg.total_units == 0 && sg.p.world == 1) continue;
        bool q5k = false;
        for (const ListOp & o : sg.ops) q5k = q5k || o.type == T_Q5_K;
        list_kernel_t fn = list_kernel(q5k);
        if (int rc = ensure_smem_attr((const void *) fn, di.max_smem)) return rc;
        if (debug) {
            fprintf(stderr, "[b200q] gemv_list ops=%d gens=%d units=%d bp=%d slots=%d smem=%d grid=%d world=%d\n", sg.p.n_ops, sg.p.n_gens,
                    (int) sg.total_units, sg.p.n_bp, sg.p.nslots, sg.smem, sg.grid, sg.p.world);
        }
        cudaLaunchConfig_t cfg{};
        cfg.gridDim          = dim3(sg.grid);
        cfg.blockDim         = dim3(list_threads());
        cfg.dynamicSmemBytes = sg.smem;
        cfg.stream           = (cudaStream_t) stream;
        cudaLaunchAttribute attr[1];
        attr[0].id              = cudaLaunchAttributeCooperative;  // all CTAs co-resident: the barrier points spin on each other
        attr[0].val.cooperative = 1;
        cfg.attrs               = attr;
        cfg.numAttrs            = 1;
        sg.p.ts = g_list_ts;
        CUDA_TRY(cudaLaunchKernelEx(&cfg, fn, sg.p));
    }
    return 0;
}

int b200q_list_destroy(b200q_list * L) {
    if (!L) return 0;
    if (L->dev_tables) cudaFree(L->dev_tables);
    if (L->own_sync) cudaFree(L->own_sync);
    delete L;
    return 0;
}

int b200q_list_info(const b200q_list * L, int * out) {
    if (!L || !out) return fail(B200Q_EINVAL, "list_info: null argument");
    int launches = 0;
    for (auto & sg : L->segs) launches += sg.total_units > 0 || sg.p.world > 1;
    out[0] = launches; out[1] = L->n_bp; out[2] = L->n_gens; out[3] = L->segs.empty() ? 0 : L->segs[0].p.nslots;
    return 0;
}

// ---- host round trip -----------------------------------------------------------------------------------
int b200q_mul_mat_host(int type, const void * w_dev, const float * act_host, float * dst_host, int64_t m, int64_t k,
                       int64_t n, void * act_dev_scratch, void * dst_dev_scratch, void * workspace,
                       size_t workspace_bytes, void * stream) {
    cudaStream_t st = (cudaStream_t) stream;
    CUDA_TRY(cudaMemcpyAsync(act_dev_scratch, act_host, (size_t) (n * k) * sizeof(float), cudaMemcpyHostToDevice, st));
    b200q_mul_mat_args a{};
    a.type = type; a.k = k; a.m = m; a.n = n;
    a.ne02 = a.ne03 = a.ne12 = a.ne13 = 1;
    a.nb01 = b200q_row_size(type, k); a.nb02 = a.nb01 * (size_t) m; a.nb03 = a.nb02;
    a.nb11 = (size_t) k * 4; a.nb12 = a.nb11 * (size_t) n; a.nb13 = a.nb12;
    a.nb1  = (size_t) m * 4; a.nb2 = a.nb1 * (size_t) n; a.nb3 = a.nb2;
    a.flags = B200Q_FLAG_WEIGHTS_IMMUTABLE;
    if (int rc = b200q_mul_mat(&a, w_dev, (const float *) act_dev_scratch, (float *) dst_dev_scratch, workspace,
                               workspace_bytes, stream)) {
        return rc;
    }
    CUDA_TRY(cudaMemcpyAsync(dst_host, dst_dev_scratch, (size_t) (n * m) * sizeof(float), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    return 0;
}

}  // extern "C"
