// selftest_mma.cu — HOST-ONLY check of the tile arithmetic of gemv_mma.cu (gemv_mma.cuh).
// The same __host__ __device__ code the kernel runs (fragment loads from the padded stage layout, the permuted
// activation image, scale / min handling) is executed here for all 32 lanes of a warp with an EMULATED mma.sync
// (fragment layouts of the PTX ISA for m16n8k32 / m16n8k16, .u8 x .s8), and compared with
//   sum_e dequantize(w)[row][e] * dx[col][e / 256] * q[col][e]     (weights through oracle/oracle.c)
// so layout mistakes are caught before any GPU time is spent.  Exit code 0 = pass.
#include "gemv_mma.cuh"

#include "oracle.h"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <vector>

using namespace b200q;
using namespace b200q::mma;

namespace b200q { namespace mma { uint8_t * host_smem = nullptr; } }

static std::mt19937 rng(4321);

struct HostExec {
    int lane = 0, pass = 0;
    static uint32_t A[32][16][4], B[32][16][2];
    static int      C[32][16][4];
    static int      kdim[16];
    static bool     a_signed[16];
    __host__ __device__ void mma32s(int id, int (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
        mma32(id, c, a, b0, b1);
#ifndef __CUDA_ARCH__
        a_signed[id] = true;
#endif
    }
    __host__ __device__ void mma32(int id, int (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
#ifndef __CUDA_ARCH__
        if (pass == 0) {
            for (int i = 0; i < 4; ++i) A[lane][id][i] = a[i];
            B[lane][id][0] = b0; B[lane][id][1] = b1;
            kdim[id] = 32;
            a_signed[id] = false;
            for (int i = 0; i < 4; ++i) c[i] = 0;
        } else {
            for (int i = 0; i < 4; ++i) c[i] = C[lane][id][i];
        }
#endif
    }
    __host__ __device__ void mma16(int id, int (&c)[4], const uint32_t (&a)[2], uint32_t b0) {
#ifndef __CUDA_ARCH__
        if (pass == 0) {
            A[lane][id][0] = a[0]; A[lane][id][1] = a[1];
            B[lane][id][0] = b0;
            kdim[id] = 16;
            a_signed[id] = false;
            for (int i = 0; i < 4; ++i) c[i] = 0;
        } else {
            for (int i = 0; i < 4; ++i) c[i] = C[lane][id][i];
        }
#endif
    }
};
uint32_t HostExec::A[32][16][4], HostExec::B[32][16][2];
int      HostExec::C[32][16][4];
int      HostExec::kdim[16];
bool     HostExec::a_signed[16];

// mma.sync.aligned.m16n8k{32,16}.row.col.s32.u8.s8.s32 — fragment layouts of the PTX ISA ("Matrix Fragments for
// mma.m16n8k32" / "m16n8k16" with 8-bit types): groupID = lane >> 2, tig = lane & 3
//   k32: a0 = A[g][4 tig ..], a1 = A[g + 8][4 tig ..], a2 = A[g][16 + 4 tig ..], a3 = A[g + 8][16 + 4 tig ..]
//        b0 = B[4 tig ..][g], b1 = B[16 + 4 tig ..][g]
//   k16: a0 = A[g][4 tig ..], a1 = A[g + 8][4 tig ..];  b0 = B[4 tig ..][g]
//   c0 = C[g][2 tig], c1 = C[g][2 tig + 1], c2 = C[g + 8][2 tig], c3 = C[g + 8][2 tig + 1]
static void emulate(int id) {
    const int K = HostExec::kdim[id];
    int Am[16][32] = {}, Bm[32][8] = {};
    for (int lane = 0; lane < 32; ++lane) {
        const int g = lane >> 2, tig = lane & 3;
        for (int i = 0; i < (K == 32 ? 4 : 2); ++i) {
            const int row = g + 8 * (i & 1), k0 = 16 * (i >> 1) + 4 * tig;
            for (int b = 0; b < 4; ++b) {
                const int v = (int) ((HostExec::A[lane][id][i] >> (8 * b)) & 0xff);
                Am[row][k0 + b] = HostExec::a_signed[id] ? (int) (int8_t) v : v;
            }
        }
        for (int i = 0; i < (K == 32 ? 2 : 1); ++i) {
            const int k0 = 16 * i + 4 * tig;
            for (int b = 0; b < 4; ++b) Bm[k0 + b][g] = (int) (int8_t) ((HostExec::B[lane][id][i] >> (8 * b)) & 0xff);
        }
    }
    for (int lane = 0; lane < 32; ++lane) {
        const int g = lane >> 2, tig = lane & 3;
        for (int i = 0; i < 4; ++i) {
            const int row = g + 8 * (i >> 1), col = 2 * tig + (i & 1);
            int       s   = 0;
            for (int k = 0; k < K; ++k) s += Am[row][k] * Bm[k][col];
            HostExec::C[lane][id][i] = s;
        }
    }
}

static void fill_row(int type, uint8_t * row, int nblocks) {
    const int bb = blck_bytes(type);
    for (int i = 0; i < nblocks * bb; ++i) row[i] = (uint8_t) rng();
    auto put_half = [&](int off, float lo, float hi) {
        std::uniform_real_distribution<float> u(lo, hi);
        const uint16_t h = orc_fp32_to_fp16(u(rng));
        row[off] = (uint8_t) h; row[off + 1] = (uint8_t) (h >> 8);
    };
    for (int b = 0; b < nblocks; ++b) {
        const int o = b * bb;
        if (type == T_Q6_K) put_half(o + 208, -0.05f, 0.05f);
        else if (type == T_Q4_0 || type == T_Q8_0 || type == T_Q5_0) put_half(o, -2.f, 2.f);
        else if (type == T_Q4_1 || type == T_Q5_1) { put_half(o, 0.f, 2.f); put_half(o + 2, -3.f, 3.f); }
        else { put_half(o, 0.f, 0.1f); put_half(o + 2, 0.f, 0.1f); }
    }
}

template <int T> static int check_type(const char * name) {
    constexpr int  BB = BlockInfo<T>::BYTES, UE = BlockInfo<T>::ELEMS;  // unit = super-block | block
    constexpr bool KQ = MmaIsKQuant<T>::value;
    const int      k = 1024, nu = k / UE, rows = 16;
    int            errors = 0;
    for (int trial = 0; trial < 16; ++trial) {
        const int ncols = 1 + trial % 8;
        const int lead  = T == T_Q4_K || T == T_Q5_K ? 0 : 2 * (trial % 8);  // the aligned copy window starts `lead` bytes before the chunk
        const int pitch = ((lead + nu * BB + 15) / 16 * 16 + 127) / 128 * 128 + (T == T_Q4_K || T == T_Q5_K ? 32 : 48);
        // weights
        std::vector<uint8_t> w((size_t) rows * nu * BB);
        std::vector<float>   wf((size_t) rows * k);
        for (int r = 0; r < rows; ++r) {
            fill_row(T, w.data() + (size_t) r * nu * BB, nu);
            orc_dequantize_row(T, w.data() + (size_t) r * nu * BB, wf.data() + (size_t) r * k, k);
        }
        // activations: random int8 + one scale per unit
        std::vector<int8_t> q((size_t) ncols * k);
        std::vector<float>  dx((size_t) ncols * nu);
        std::uniform_real_distribution<float> ud(0.001f, 0.02f);
        for (auto & v : q) v = (int8_t) ((int) (rng() % 255) - 127);
        for (auto & v : dx) v = ud(rng);
        // shared-memory image
        const int col_pitch = k + (KQ ? 64 : 16), bsb = T == T_Q6_K ? 32 : 16, dxb = KQ ? 4 : 8;
        const int dx_pitch = (nu + 1) * dxb, bs_pitch = (nu + 1) * bsb;
        const int off_q = 128, off_dx = off_q + 8 * col_pitch, off_bs = (off_dx + ncols * dx_pitch + 15) / 16 * 16;
        const int off_stage = (off_bs + ncols * bs_pitch + 127) / 128 * 128;
        std::vector<uint8_t> smem((size_t) off_stage + rows * pitch + 64, 0xA5);
        for (int c = 0; c < ncols; ++c) {
            for (int e = 0; e < k; e += 4) {
                const int off = KQ ? (e / 256) * 256 + act_word_off<T>(e % 256) : e;
                memcpy(&smem[off_q + c * col_pitch + off], &q[(size_t) c * k + e], 4);
            }
            for (int s = 0; s < nu; ++s) {
                memcpy(&smem[off_dx + c * dx_pitch + s * dxb], &dx[c * nu + s], 4);
                if (!KQ) {
                    int sum = 0;
                    for (int i = 0; i < 32; ++i) sum += q[(size_t) c * k + s * 32 + i];
                    const float dsum = dx[c * nu + s] * (float) sum;
                    memcpy(&smem[off_dx + c * dx_pitch + s * 8 + 4], &dsum, 4);
                    continue;
                }
                const int per = T == T_Q6_K ? 16 : 32;
                for (int g = 0; g < 256 / per; ++g) {
                    int sum = 0;
                    for (int i = 0; i < per; ++i) sum += q[(size_t) c * k + s * 256 + g * per + i];
                    const int16_t s16 = (int16_t) sum;
                    memcpy(&smem[off_bs + c * bs_pitch + s * bsb + g * 2], &s16, 2);
                }
            }
        }
        for (int r = 0; r < rows; ++r) memcpy(&smem[off_stage + lead + r * pitch], w.data() + (size_t) r * nu * BB, (size_t) nu * BB);
        host_smem = smem.data();
        TileCtx t{ (uint32_t) off_q, (uint32_t) off_dx, (uint32_t) off_bs, col_pitch, dx_pitch, bs_pitch, ncols };
        float   facc[32][4] = {};
        for (int s = 0; s < nu; ++s) {
            float    scratch[32][4];
            HostExec ex;
            for (int pass = 0; pass < 2; ++pass) {
                for (int lane = 0; lane < 32; ++lane) {
                    ex.lane = lane; ex.pass = pass;
                    const uint32_t blk0 = (uint32_t) (off_stage + lead + (lane >> 2) * pitch + s * BB);
                    const uint32_t blk1 = blk0 + 8 * pitch;
                    for (int i = 0; i < 4; ++i) scratch[lane][i] = facc[lane][i];
                    TileSb<T, HostExec>::run(ex, t, lane, s, blk0, blk1, scratch[lane]);
                }
                if (pass == 0) {
                    for (int id = 0; id < (T == T_Q6_K ? 16 : KQ ? 8 : 1); ++id) emulate(id);
                }
            }
            for (int lane = 0; lane < 32; ++lane) for (int i = 0; i < 4; ++i) facc[lane][i] = scratch[lane][i];
        }
        for (int lane = 0; lane < 32; ++lane) {
            for (int i = 0; i < 4; ++i) {
                const int row = (lane >> 2) + 8 * (i >> 1), col = 2 * (lane & 3) + (i & 1);
                if (col >= ncols) continue;
                double want = 0.0, mag = 0.0;
                for (int e = 0; e < k; ++e) {
                    const double v = (double) wf[(size_t) row * k + e] * (double) dx[col * nu + e / UE] * (double) q[(size_t) col * k + e];
                    want += v; mag += fabs(v);
                }
                if (fabs((double) facc[lane][i] - want) > 2e-5 * mag + 1e-6) {
                    if (errors < 8) printf("%s: trial %d row %d col %d got %.8g want %.8g (mag %.3g)\n", name, trial, row, col, facc[lane][i], want, mag);
                    ++errors;
                }
            }
        }
    }
    printf("%-6s %s\n", name, errors ? "FAIL" : "ok");
    return errors;
}

int main() {
    int e = 0;
    e += check_type<T_Q4_K>("q4_K");
    e += check_type<T_Q5_K>("q5_K");
    e += check_type<T_Q6_K>("q6_K");
    e += check_type<T_Q4_0>("q4_0");
    e += check_type<T_Q8_0>("q8_0");
    e += check_type<T_Q4_1>("q4_1");
    e += check_type<T_Q5_0>("q5_0");
    e += check_type<T_Q5_1>("q5_1");
    printf(e ? "selftest_mma FAILED (%d)\n" : "selftest_mma passed\n", e);
    return e ? 1 : 0;
}
