// gemv_list.cu — persistent decode kernel: ONE launch walks a whole list of n = 1 matmuls (b200q_list_*).
//
// It replaces the reference's per-matmul launch pair (quantize_q8_1 + mul_mat_vec_q, ggml-cuda/quantize.cu:148-160,
// mmvq.cu:500-570) for runs of decode GEMVs, and removes what bounded gemv_kq.cu's chain of launches (profiles/
// r01_gemv_timeline_v1.txt: ~3.5 us per launch between "predecessor completed" and "first weight byte consumed",
// during which at most the next launch's 110 KB/SM prefetch moves):
//
//   * one CTA per SM lives for the whole list; units of work (a few rows of one matrix) are dealt round-robin over
//     (CTA, warp) along the concatenated list, so every CTA walks every matrix and the list is balanced to +-1 unit;
//   * every warp owns 2-3 ring slots and refills a slot the moment it has consumed it (slot ownership as in gemv_kq.cu)
//     with the NEXT stage of its sequence — which may belong to a later matrix: the weights are immutable, so the HBM
//     stream never stops at a matrix boundary, a dependency or an activation reload;
//   * activations are quantized per CTA into one of two shared-memory images ("generations": consecutive ops that read
//     the same vector share one) by a third warpgroup of 4 QUANTIZER warps that runs one generation ahead of the 8 compute
//     warps (image FULL / FREE named barriers; setmaxnreg gives the compute warps 208 registers, the quantizers 88 —
//     2 x 208 + 88 = the 3 x 168 a scheduler partition was allocated): the L2 round trips of the f32 vector and the compute
//     warps' skew at a switch left the critical path (they were ~1.2 us per generation, 18 % of the 8B pass: every shape
//     alone ran at 81-87 % of HBM, the pass at 69 %; now 87 %).  Quantizing a dependent vector with all 12 warps was tried:
//     inlined it spills in the stage loop, out of line the call forces the 168-register ABI (both slower);
//   * a generation whose vector is produced by earlier ops of the list (or by the other GPUs of a row split) sits behind
//     a BARRIER POINT: every CTA counts itself in when its 8 warps have stored (and fenced) everything before it
//     (one red.release per CTA), readers poll with ld.acquire; with row split the rank that completes a barrier point
//     publishes `base + bp` into its slot of every peer's sync block over NVLink (release at system scope) and readers
//     wait for all ranks' slots — per-matmul arrival flags instead of a collective or a kernel boundary;
//   * rows longer than a ring slot are cut into k-chunks that are consecutive stages of one warp, the partial sums stay
//     in registers (k = 14336 q6_K, k = 28672 of the 70B shapes).
// Fixed summation order per (type, k): bit-reproducible, independent of grid, list composition and row split.
#include "gemv_list.cuh"

#include "blocks.cuh"
#include "kq_dot.cuh"
#include "sm100.cuh"

#include <cuda_runtime.h>

namespace b200q {

namespace {

using namespace sm100;
using namespace kqdot;

constexpr int CTHREADS = LIST_WARPS * 32;                    // compute warps (warpgroups 0 and 1)
constexpr int NTHREADS = (LIST_WARPS + LIST_QWARPS) * 32;    // + the quantizer warpgroup
// named barriers: 0 __syncthreads; 2, 3 barrier point (compute warps); 4, 5 image FULL; 6, 7 image FREE; 8 barrier point done
constexpr int BAR_FULL = 4, BAR_FREE = 6, BAR_BPDONE = 8;
constexpr uint32_t OFF_FULL = 0;    // full-barriers, 8 B each (LIST_MAX_SLOTS)
constexpr uint32_t OFF_CNT  = 8 * LIST_MAX_SLOTS;  // 2 x u32: warps of this CTA that reached barrier point (bp & 1)

extern __shared__ __align__(128) uint8_t list_smem[];

struct ActLayout {
    int off_q, off_dx, off_bs16, off_bs32, nsb;
};

__device__ __forceinline__ unsigned ld_acquire_gpu(const unsigned * p) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned * p) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(unsigned * p, unsigned v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned atom_add_release_gpu(unsigned * p, unsigned v) {
    unsigned old;
    asm volatile("atom.add.release.gpu.global.u32 %0, [%1], %2;" : "=r"(old) : "l"(p), "r"(v) : "memory");
    return old;
}
__device__ __forceinline__ unsigned long long gtime() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
__device__ __forceinline__ void named_bar_arrive(int id, int nthreads) {
    asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// one op record from the shared-memory table (48 bytes = 3 x int4, gemv_list.cuh)
struct OpRec {
    const uint8_t * w;
    float *         dst;
    int             m, rb, nsb, csb, gen, type, rps_log2, nch, bb;
};
__device__ __forceinline__ OpRec op_load(uint32_t tab, int op) {
    OpRec      r;
    const int4 l = lds128(tab + 48u * (uint32_t) op);
    const int4 h = lds128(tab + 48u * (uint32_t) op + 32u);
    r.w          = (const uint8_t *) (((unsigned long long) (uint32_t) l.y << 32) | (uint32_t) l.x);
    r.dst        = (float *) (((unsigned long long) (uint32_t) l.w << 32) | (uint32_t) l.z);
    r.m          = (int) lds32(tab + 48u * (uint32_t) op + 16u);
    r.rb         = h.x;
    r.nsb        = (int) ((uint32_t) h.y & 0xffffu);
    r.csb        = (int) ((uint32_t) h.y >> 16);
    r.gen        = (int) ((uint32_t) h.z & 0xffffu);
    r.type       = (int) (((uint32_t) h.z >> 16) & 0xffu);
    r.rps_log2   = (int) (((uint32_t) h.z >> 24) & 0xfu);
    r.nch        = (int) ((uint32_t) h.z >> 28) + 1;
    r.bb         = (int) ((uint32_t) h.w & 0xffu);
    return r;
}

// a warp's position in its own stage sequence: chunk `ch` of unit `unit` of op `op`; `left` units of this op remain for
// this warp, `woff` = units this CTA has taken so far, mod LIST_WARPS (the warp that gets the CTA's next unit)
struct Walk {
    int op, left, unit, ch, woff;
};

template <int T> __device__ __forceinline__ float dot_units(const ActLayout & al, uint32_t abase, uint32_t row, int s0, int s1, int step) {
    float acc = 0.0f;
    for (int s = s0; s < s1; s += step) {
        if constexpr (T == T_Q4_K || T == T_Q5_K || T == T_Q6_K) acc += SbDot<T>::run(al, abase, row, s);
        else acc += B32Dot<T>::run(al, abase, row, s);
    }
    return acc;
}

#ifdef B200Q_LIST_TRACE
constexpr int LIST_TRACE_EVENTS = 4096;
#define EV(tag_, val_)                                                                                                \
    do {                                                                                                              \
        if (p.ts && blockIdx.x == 0 && nev < LIST_TRACE_EVENTS) {                                                     \
            if (lane == 0) {                                                                                          \
                unsigned long long * e_ = p.ts + ((size_t) warp * LIST_TRACE_EVENTS + nev) * 2;                       \
                e_[0] = ((unsigned long long) (tag_) << 32) | (unsigned) (val_);                                      \
                e_[1] = (unsigned long long) clock64();                                                               \
            }                                                                                                         \
            ++nev;                                                                                                    \
        }                                                                                                             \
    } while (0)
#else
#define EV(tag_, val_) do { } while (0)
#endif

// Q5K: the instantiation that also serves q5_K ops.  Kept apart because its extra case costs the common instantiation two
// spilled registers and 5 % of the headline pass (5.39 -> 5.10 TB/s); lists without q5_K launch the kernel without it.
template <bool Q5K> __global__ void __launch_bounds__(NTHREADS, 1) gemv_list_kernel(const ListParams p) {
    const uint32_t sb   = smem_u32(list_smem);
    const int      warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int      G    = (int) gridDim.x;
    const uint32_t tab  = sb + (uint32_t) p.off_ops;
    const uint32_t gtab = sb + (uint32_t) p.off_gens;
#ifdef B200Q_LIST_TRACE
    int nev = 0;
#endif

    // tables -> shared memory
    {
        const int4 * so = (const int4 *) p.ops;
        const int    no = p.n_ops * 3;
        for (int i = threadIdx.x; i < no; i += NTHREADS) {
            const int4 v = __ldg(so + i);
            asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(tab + 16u * i), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
        }
        const int4 * sg = (const int4 *) p.gens;
        for (int i = threadIdx.x; i < p.n_gens * 2; i += NTHREADS) {
            const int4 v = __ldg(sg + i);
            asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(gtab + 16u * i), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
        }
    }
    if (threadIdx.x == 0) {
        for (int s = 0; s < p.nslots; ++s) mbar_init(sb + OFF_FULL + 8 * s, 1);
        asm volatile("st.shared.v2.u32 [%0], {%1, %1};" ::"r"(sb + OFF_CNT), "r"(0) : "memory");
        mbar_fence_init();
    }
    __syncthreads();

    const int      S     = p.nslots;
    const int      D     = S / LIST_WARPS + (warp < S % LIST_WARPS ? 1 : 0);
    const int      slot0 = warp * (S / LIST_WARPS) + min(warp, S % LIST_WARPS);
    const int      n_ops = p.n_ops;
    const unsigned base  = p.world > 1 ? ld_acquire_gpu(p.sync + LIST_SYNC_BASE) : 0u;
    const uint64_t pol   = policy_evict_first();

    // next op in which this warp has units (every op is visited so that woff stays in step across the CTA's warps)
    auto next_op = [&](Walk & w) {
        for (;;) {
            if (++w.op >= n_ops) {
                w.left = 0;
                return;
            }
            const int4 q = lds128(tab + 48u * (uint32_t) w.op + 16u);  // m, base, rem, rot
            int        cp = (int) blockIdx.x + q.w;
            if (cp >= G) cp -= G;
            const int ncta  = q.y + (cp < q.z ? 1 : 0);
            const int first = (warp - w.woff + LIST_WARPS) % LIST_WARPS;
            w.woff          = (w.woff + ncta) % LIST_WARPS;
            if (ncta > first) {
                w.unit = cp * q.y + min(cp, q.z) + first;
                w.left = (ncta - first + LIST_WARPS - 1) / LIST_WARPS;
                w.ch   = 0;
                return;
            }
        }
    };
    auto advance = [&](Walk & w, int nch) {
        if (++w.ch < nch) return;
        w.ch = 0;
        w.unit += LIST_WARPS;
        if (--w.left == 0) next_op(w);
    };
    // bulk copy of the stage at `w` into ring slot `slot` (warp-uniform code, one elected lane issues; rows of the
    // 32-element block formats need not be 16-byte multiples: copy the aligned window around the bytes)
    auto issue = [&](const Walk & w, int slot) -> int {
        const uint32_t o  = tab + 48u * (uint32_t) w.op;
        const int4     l  = lds128(o);
        const int4     h  = lds128(o + 32u);
        const int      m  = (int) lds32(o + 16u);
        const int      rb = h.x, nsb = (int) ((uint32_t) h.y & 0xffffu), csb = (int) ((uint32_t) h.y >> 16);
        const int      rps_log2 = (int) (((uint32_t) h.z >> 24) & 0xfu), nch = (int) ((uint32_t) h.z >> 28) + 1;
        const int      bb  = (int) ((uint32_t) h.w & 0xffu);
        const int      r0  = w.unit << rps_log2;
        const int      nr  = min(1 << rps_log2, m - r0);
        const int      s0  = w.ch * csb;
        const int      len = nch == 1 ? nr * rb : (min(nsb, s0 + csb) - s0) * bb;
        const unsigned long long src = (((unsigned long long) (uint32_t) l.y << 32) | (uint32_t) l.x) +
                                       (unsigned long long) r0 * (unsigned) rb + (unsigned) (s0 * bb);
        const uint32_t lead  = (uint32_t) src & 15u;
        const uint32_t bytes = (lead + (uint32_t) len + 15u) & ~15u;
        if (elect_one()) {
            const uint32_t full = sb + OFF_FULL + 8 * slot;
            mbar_expect_tx(full, bytes);
            bulk_g2s_hint(sb + (uint32_t) p.off_ring + (uint32_t) slot * LIST_SLOT_BYTES, (const void *) (src - lead), bytes, full, pol);
        }
        return nch;
    };

    // ---- barrier point: everything before it is complete on every CTA (and every GPU) -------------------------------
    // All warps of the CTA meet (bar.sync orders their result stores before warp 0's release at CTA scope; the release is
    // cumulative), ONE thread counts the CTA in and polls — 148 pollers on the counter's L2 line instead of 1184, which sat
    // in front of the arrivals (4.4 -> 2.4 us per barrier point) — and a second CTA barrier lets everybody continue behind
    // its acquire.
    auto barrier_point = [&](int bp) {
        named_bar_sync(2, CTHREADS);
        if (warp == 0) {
            if (lane == 0) {
                if (p.world > 1) {
                    // row split: this CTA's peer stores must have landed before it is counted in (one system-scope fence per
                    // CTA), then it counts itself in ON EVERY GPU — local and remote reductions, no hop through a "last CTA" —
                    // and polls its own GPU's counter: all CTAs of all GPUs are in when it reads base + bp * G * world
                    asm volatile("fence.acq_rel.sys;" ::: "memory");
                    asm volatile("red.relaxed.sys.global.add.u32 [%0], 1;" ::"l"(p.sync + LIST_SYNC_ARRIVE) : "memory");
                    for (int r = 0; r < p.n_peers; ++r) {
                        asm volatile("red.relaxed.sys.global.add.u32 [%0], 1;" ::"l"(p.peer_sync[r] + LIST_SYNC_ARRIVE) : "memory");
                    }
                    const unsigned want = base + (unsigned) bp * (unsigned) G * (unsigned) p.world;
                    while ((int) (ld_acquire_sys(p.sync + LIST_SYNC_ARRIVE) - want) < 0) {}
                } else {
                    const unsigned want = (unsigned) bp * (unsigned) G;
                    atom_add_release_gpu(p.sync + LIST_SYNC_PROGRESS, 1u);
                    while ((int) (ld_acquire_gpu(p.sync + LIST_SYNC_PROGRESS) - want) < 0) {}
                }
            }
            __syncwarp();
        }
        named_bar_sync(3, CTHREADS);
    };

    // ---- generations ------------------------------------------------------------------------------------------------
    // decode of one generation record + the layout of its image
    ActLayout al{};
    uint32_t  abase = 0;
    auto gen_layout = [&](int g, const float *& act, int & qmode, int & wait, int & bp) {
        const int4 gi  = lds128(gtab + 32u * (uint32_t) g);
        act            = (const float *) (((unsigned long long) (uint32_t) gi.y << 32) | (uint32_t) gi.x);
        const int nsb  = (int) ((uint32_t) gi.z & 0xffffu);
        qmode          = (int) (((uint32_t) gi.z >> 16) & 0xffu);
        wait           = (int) ((uint32_t) gi.z >> 24);
        bp             = (int) ((uint32_t) gi.w & 0xffffu);
        al.nsb   = nsb;
        al.off_q = 0;
        if (qmode == 0) {
            al.off_dx   = nsb * 256;
            al.off_bs16 = (al.off_dx + nsb * 4 + 15) & ~15;
            al.off_bs32 = al.off_bs16 + nsb * 32;
        } else {
            al.off_dx   = nsb * 32;
            al.off_bs16 = al.off_bs32 = 0;
        }
        abase = sb + (uint32_t) p.off_act + (uint32_t) (g & 1) * (uint32_t) p.act_stride;
    };

    // ---- quantizer warpgroup: image g & 1 <- generation g, one generation ahead of the compute warps -------------------
    if (warp >= LIST_WARPS) {
        asm volatile("setmaxnreg.dec.sync.aligned.u32 88;");
        const int qw = warp - LIST_WARPS;
        for (int g = 0; g < p.n_gens; ++g) {
            const float * act;
            int           qmode, wait, bp;
            gen_layout(g, act, qmode, wait, bp);
            if (g >= 2) named_bar_sync(BAR_FREE + (g & 1), NTHREADS);  // the compute warps are done with generation g - 2
            if (wait) named_bar_sync(BAR_BPDONE, NTHREADS);            // ... and have passed the barrier point in front of g
            const int4 gx = lds128(gtab + 32u * (uint32_t) g + 16u);  // xform (upper half of the record's last word) and its second vector
            if (((uint32_t) lds32(gtab + 32u * (uint32_t) g + 12u) >> 16) == 1u) {
                const ActSwiglu src{ act, (const float *) (((unsigned long long) (uint32_t) gx.y << 32) | (uint32_t) gx.x) };
                if (qmode == 0) quantize_act_from<LIST_QWARPS>(al, src, abase, qw, lane);
                else if (qmode == 1) quantize_act32_from<false, LIST_QWARPS>(al, src, abase, qw, lane);
                else quantize_act32_from<true, LIST_QWARPS>(al, src, abase, qw, lane);
            } else if (qmode == 0) quantize_act<LIST_QWARPS>(al, act, abase, qw, lane);
            else if (qmode == 1) quantize_act32<false, LIST_QWARPS>(al, act, abase, qw, lane);
            else quantize_act32<true, LIST_QWARPS>(al, act, abase, qw, lane);
            named_bar_arrive(BAR_FULL + (g & 1), NTHREADS);
        }
    } else {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 208;");

    // compute warps: switch to generation g (every warp takes part in every switch, with or without work in it)
    RegAct RA;
    auto gen_switch = [&](int g) {
        const float * act;
        int           qmode, wait, bp;
        EV(1, g);
        // this warp no longer reads the previous image; generation g + 1 will overwrite it
        if (g >= 1 && g + 1 < p.n_gens) named_bar_arrive(BAR_FREE + ((g - 1) & 1), NTHREADS);
        gen_layout(g, act, qmode, wait, bp);
        if (wait) {
            barrier_point(bp);
            named_bar_arrive(BAR_BPDONE, NTHREADS);
        }
        EV(2, g);
        named_bar_sync(BAR_FULL + (g & 1), NTHREADS);
        EV(3, g);
        // column-owning lanes: this generation's quantized activations of "my" super-block columns -> registers
        if (qmode == 0 && al.nsb <= LIST_REG_ACT_NSB) {
            const int lpr = al.nsb == 1 ? 1 : 1 << (32 - __clz(al.nsb - 1));
            const int col = lane & (lpr - 1);
            if (col < al.nsb) RA.load(al, abase, col);
        }
        EV(4, g);
    };

    // ---- main loop ---------------------------------------------------------------------------------------------------
    Walk wi{ -1, 0, 0, 0, 0 };
    next_op(wi);
    Walk wc = wi;
    for (int i = 0; i < D && wi.op < n_ops; ++i) {
        const int nch = issue(wi, slot0 + i);
        advance(wi, nch);
    }
    int      g = -1, slot = 0, cur_op = -1;
    uint32_t parity = 0;
    float    acc = 0.0f;
    OpRec    r{};
    // per-op lane roles
    bool regs = false;
    int  lpr_log2 = 0, sub = 0, lrow = 0, R = 1;
    while (wc.op < n_ops) {
        if (wc.op != cur_op) {
            cur_op = wc.op;
            r      = op_load(tab, cur_op);
            while (g < r.gen) gen_switch(++g);
            const bool kq = r.type == T_Q4_K || r.type == T_Q6_K;  // q5_K (Q5_K_M files) walks its rows with activations in shared memory
            regs          = kq && r.nsb <= LIST_REG_ACT_NSB;
            // lanes per row: column-owning lanes cover the columns; row-owning lanes split a row 32 / rps ways
            lpr_log2 = regs ? (r.nsb == 1 ? 0 : 32 - __clz(r.nsb - 1)) : 5 - r.rps_log2;
            sub      = lane & ((1 << lpr_log2) - 1);
            lrow     = lane >> lpr_log2;
            R        = regs ? max(1, (1 << r.rps_log2) >> (5 - lpr_log2)) : 1;
        }
        EV(5, wc.op);
        mbar_wait(sb + OFF_FULL + 8 * (slot0 + slot), parity);
        EV(6, wc.unit);
        const int  rb   = r.rb, bb = r.bb;
        const int  r0   = wc.unit << r.rps_log2;
        const int  nr   = min(1 << r.rps_log2, r.m - r0);
        const int  s0   = wc.ch * r.csb;
        const int  s1   = min(r.nsb, s0 + r.csb);
        const int  lpr  = 1 << lpr_log2;
        const uint32_t lead = ((uint32_t) (uintptr_t) r.w + (uint32_t) r0 * (uint32_t) rb + (uint32_t) (s0 * bb)) & 15u;
        // address of the (virtual) start of the stage's first row inside the slot: a chunk holds bytes [s0 * bb, s1 * bb) of it
        const uint32_t row0 = sb + (uint32_t) p.off_ring + (uint32_t) (slot0 + slot) * LIST_SLOT_BYTES + lead - (uint32_t) (s0 * bb);
        const int  sets_log2 = 5 - lpr_log2;
        const bool last_ch   = wc.ch == r.nch - 1;
        if (wc.ch == 0) acc = 0.0f;
        auto finish_row = [&](float v, int rr) {  // fixed-order reduction over the lanes of a row, result (and peer) stores
            for (int o = lpr >> 1; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            if (sub == 0 && rr < nr) {
                float * d = r.dst + r0 + rr;
                *d = v;
                for (int q = 0; q < p.n_peers; ++q) *(float *) ((char *) d + p.peer_delta[q]) = v;  // NVLink peer stores
            }
        };
        if (regs) {
#pragma unroll 1
            for (int i = 0; i < R; i += 2) {  // the rows this lane walks, two at a time (independent instruction streams)
                const int  rr0 = lrow + (i << sets_log2), rr1 = lrow + ((i + 1) << sets_log2);
                float      a0 = r.nch > 1 ? acc : 0.0f, a1 = 0.0f;
                const bool pair = i + 1 < R;  // warp-uniform; a lane whose second row is past a ragged end computes on stale slot bytes and drops it
                if (rr0 < nr && sub >= s0 && sub < s1) {
                    const uint32_t rowa = row0 + (uint32_t) rr0 * (uint32_t) rb, rowb = row0 + (uint32_t) rr1 * (uint32_t) rb;
                    if (r.type == T_Q4_K) {
                        if (pair) {
                            const float x0 = SbDot<T_Q4_K>::core(rowa + (uint32_t) sub * 144u, RA);
                            const float x1 = SbDot<T_Q4_K>::core(rowb + (uint32_t) sub * 144u, RA);
                            a0 += x0;
                            a1 += x1;
                        } else {
                            a0 += SbDot<T_Q4_K>::core(rowa + (uint32_t) sub * 144u, RA);
                        }
                    } else {
                        if (pair) {
                            const float x0 = SbDot<T_Q6_K>::core(rowa + (uint32_t) sub * 210u, RA);
                            const float x1 = SbDot<T_Q6_K>::core(rowb + (uint32_t) sub * 210u, RA);
                            a0 += x0;
                            a1 += x1;
                        } else {
                            a0 += SbDot<T_Q6_K>::core(rowa + (uint32_t) sub * 210u, RA);
                        }
                    }
                }
                if (!last_ch) {
                    acc = a0;
                } else if (!pair) {
                    finish_row(a0, rr0);
                } else if (lpr == 1) {
                    finish_row(a0, rr0);
                    finish_row(a1, rr1);
                } else {
                    // two rows at once: the upper half of a row's lanes takes over row 1, the lower half row 0 (one exchange),
                    // then both reduce side by side — log2(lpr) shuffles for the pair instead of 2 log2(lpr)
                    const int   h    = lpr >> 1;
                    const bool  up   = (sub & h) != 0;
                    const float keep = up ? a1 : a0, send = up ? a0 : a1;
                    float       v    = keep + __shfl_xor_sync(0xffffffffu, send, h);
                    for (int o = h >> 1; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                    const int rr = up ? rr1 : rr0;
                    if ((sub & (h - 1)) == 0 && rr < nr) {
                        float * d = r.dst + r0 + rr;
                        *d = v;
                        for (int q = 0; q < p.n_peers; ++q) *(float *) ((char *) d + p.peer_delta[q]) = v;  // NVLink peer stores
                    }
                }
            }
        } else {
            if (lrow < nr) {
                const uint32_t row = row0 + (uint32_t) lrow * (uint32_t) rb;
                switch (r.type) {
                    case T_Q4_K:  acc += dot_units<T_Q4_K>(al, abase, row, s0 + sub, s1, lpr); break;
                    case T_Q5_K:  if constexpr (Q5K) acc += dot_units<T_Q5_K>(al, abase, row, s0 + sub, s1, lpr); break;
                    case T_Q6_K:  acc += dot_units<T_Q6_K>(al, abase, row, s0 + sub, s1, lpr); break;
                    case T_Q8_0:  acc += dot_units<T_Q8_0>(al, abase, row, s0 + sub, s1, lpr); break;
                    case T_Q4_0:  acc += dot_units<T_Q4_0>(al, abase, row, s0 + sub, s1, lpr); break;
                    case T_Q4_1:  acc += dot_units<T_Q4_1>(al, abase, row, s0 + sub, s1, lpr); break;
                    case T_Q5_0:  acc += dot_units<T_Q5_0>(al, abase, row, s0 + sub, s1, lpr); break;
                    case T_Q5_1:  acc += dot_units<T_Q5_1>(al, abase, row, s0 + sub, s1, lpr); break;
                    case T_MXFP4: acc += dot_units<T_MXFP4>(al, abase, row, s0 + sub, s1, lpr); break;
                    default: break;
                }
            }
            if (last_ch) finish_row(acc, lrow);
        }
        __syncwarp();  // every lane is done reading the slot (their loads have been consumed by the arithmetic above)
        EV(7, 0);
        if (wi.op < n_ops) {
            const int nch = issue(wi, slot0 + slot);
            advance(wi, nch);
        }
        EV(9, 0);
        advance(wc, r.nch);
        if (++slot == D) {
            slot = 0;
            parity ^= 1u;
        }
    }
    while (g < p.n_gens - 1) gen_switch(++g);  // warps without work still take part in every switch
    if (p.final_bp) barrier_point(p.n_bp);
    }  // compute warps

    // ---- exit: the last CTA re-arms the sync block for the next launch ------------------------------------------------
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        const unsigned prev = atomicAdd(p.sync + LIST_SYNC_EXIT, 1u);
        if (prev == (unsigned) G - 1u) {
            p.sync[LIST_SYNC_PROGRESS] = 0u;
            p.sync[LIST_SYNC_EXIT]     = 0u;
            if (p.world > 1) p.sync[LIST_SYNC_BASE] = base + (unsigned) p.n_bp * (unsigned) G * (unsigned) p.world;
            __threadfence();
        }
    }
}

}  // namespace

list_kernel_t list_kernel(bool with_q5k) { return with_q5k ? gemv_list_kernel<true> : gemv_list_kernel<false>; }
int           list_threads() { return NTHREADS; }
bool          list_type_ok(int type) {
    switch (type) {
        case T_Q4_K: case T_Q5_K: case T_Q6_K: case T_Q8_0: case T_Q4_0: case T_Q4_1: case T_Q5_0: case T_Q5_1: case T_MXFP4: return true;
        default: return false;
    }
}

}  // namespace b200q
