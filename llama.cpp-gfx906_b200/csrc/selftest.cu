// selftest.cu — HOST-ONLY check of the device decoders (blocks.cuh) and GEMV units (units.cuh).
// The same __host__ __device__ code that the kernels run is executed here on the CPU against the
// oracle (oracle/oracle.c), so layout mistakes are caught before any GPU time is spent.
//   * decode_elem<T> must be bit-identical to orc_dequantize_row
//   * Unit<T>::decode + dp4a dot must equal  sum_i dequant(w)_i * (d_x * q_i)  to fp32 rounding
// Build: make -C llama.cpp-gfx906_b200/csrc ../lib/selftest ; exit code 0 = pass.
#include "units.cuh"

#include "oracle.h"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <vector>

using namespace b200q;

static std::mt19937 rng(1234);

// random but numerically sane blocks: random payload bytes, fp16 scales in a finite range
static void fill_row(int type, std::vector<uint8_t> & row, int nblocks) {
    const int bb = blck_bytes(type);
    row.assign((size_t) nblocks * bb + 64, 0);
    for (size_t i = 0; i < (size_t) nblocks * bb; ++i) row[i] = (uint8_t) rng();
    auto put_half = [&](size_t off, float lo, float hi) {
        std::uniform_real_distribution<float> u(lo, hi);
        uint16_t h = orc_fp32_to_fp16(u(rng));
        row[off] = (uint8_t) h; row[off + 1] = (uint8_t) (h >> 8);
    };
    for (int b = 0; b < nblocks; ++b) {
        const size_t o = (size_t) b * bb;
        switch (type) {
            case T_Q4_0: case T_Q5_0: case T_Q8_0: put_half(o, -2.f, 2.f); break;
            case T_Q4_1: case T_Q5_1: put_half(o, 0.f, 2.f); put_half(o + 2, -3.f, 3.f); break;
            case T_MXFP4: row[o] = (uint8_t) (110 + rng() % 30); break;
            case T_Q4_K: case T_Q5_K: put_half(o, 0.f, 0.1f); put_half(o + 2, 0.f, 0.1f); break;
            case T_Q6_K: put_half(o + 208, -0.05f, 0.05f); break;
            case T_Q2_K: put_half(o + 80, 0.f, 0.1f); put_half(o + 82, 0.f, 0.1f); break;
            case T_Q3_K: put_half(o + 108, -0.05f, 0.05f); break;
            case T_IQ4_NL: put_half(o, -0.05f, 0.05f); break;
            case T_IQ4_XS: put_half(o, -0.005f, 0.005f); break;
        }
    }
}

template <int T> static int check_type(const char * name) {
    using U          = Unit<T>;
    const int be     = blck_elems(T);
    const int k      = 1024;  // multiple of 256
    const int nb     = k / be;
    int       errors = 0;
    for (int trial = 0; trial < 20; ++trial) {
        std::vector<uint8_t> row;
        fill_row(T, row, nb);
        // 1. bit-exact decode, including misaligned row starts
        std::vector<float> ref(k);
        orc_dequantize_row(T, row.data(), ref.data(), k);
        for (int mis = 0; mis < 4; ++mis) {
            std::vector<uint8_t> shifted(row.size() + 16);
            memcpy(shifted.data() + mis, row.data(), row.size());
            for (int i = 0; i < k; ++i) {
                float v = decode_elem<T>(shifted.data() + mis, i);
                if (memcmp(&v, &ref[i], 4) != 0) {
                    if (errors < 5) printf("%s: decode mismatch i=%d got %g want %g\n", name, i, v, ref[i]);
                    ++errors;
                }
            }
        }
        // 2. units: int8 activations with per-32 scale
        std::vector<int8_t> q(k);
        std::vector<float2> ds(k / 32);
        std::vector<float>  xdeq(k);
        std::uniform_real_distribution<float> ud(0.001f, 0.02f);
        for (int b = 0; b < k / 32; ++b) {
            const float d = ud(rng);
            int         s = 0;
            for (int j = 0; j < 32; ++j) {
                q[b * 32 + j] = (int8_t) ((int) (rng() % 255) - 127);
                s += q[b * 32 + j];
                xdeq[b * 32 + j] = d * (float) q[b * 32 + j];
            }
            ds[b] = make_float2(d, d * (float) s);
        }
        alignas(16) static int qa_words[1024 / 4];
        memcpy(qa_words, q.data(), k);
        for (int mis = 0; mis < 4; ++mis) {
            std::vector<uint8_t> shifted(row.size() + 32);
            // keep the buffer 4-byte aligned and shift the row by `mis` bytes inside it
            uint8_t * base = (uint8_t *) (((uintptr_t) shifted.data() + 15) & ~(uintptr_t) 15);
            memcpy(base + mis, row.data(), (size_t) nb * blck_bytes(T));
            double got = 0.0, want = 0.0, mag = 0.0;
            for (int u = 0; u < k / U::ELEMS; ++u) {
                int   wq[U::NG * U::GW];
                float ws[U::NG], wm[U::NG];
                U::decode((const uint8_t *) (base + mis), u, wq, ws, wm);
                got += unit_dot<T>(wq, ws, wm, u, qa_words, ds.data());
            }
            for (int i = 0; i < k; ++i) {
                want += (double) ref[i] * (double) xdeq[i];
                mag += fabs((double) ref[i] * (double) xdeq[i]);
            }
            if (fabs(got - want) > 2e-5 * mag + 1e-6) {
                if (errors < 10) printf("%s: unit dot mismatch (mis=%d) got %.8g want %.8g (mag %.3g)\n", name, mis, got, want, mag);
                ++errors;
            }
        }
    }
    printf("%-6s %s\n", name, errors ? "FAIL" : "ok");
    return errors;
}

int main() {
    int e = 0;
    e += check_type<T_Q4_0>("q4_0");
    e += check_type<T_Q4_1>("q4_1");
    e += check_type<T_Q5_0>("q5_0");
    e += check_type<T_Q5_1>("q5_1");
    e += check_type<T_Q8_0>("q8_0");
    e += check_type<T_MXFP4>("mxfp4");
    e += check_type<T_Q4_K>("q4_K");
    e += check_type<T_Q5_K>("q5_K");
    e += check_type<T_Q6_K>("q6_K");
    e += check_type<T_Q2_K>("q2_K");
    e += check_type<T_Q3_K>("q3_K");
    e += check_type<T_IQ4_NL>("iq4_nl");
    e += check_type<T_IQ4_XS>("iq4_xs");
    printf(e ? "selftest FAILED (%d)\n" : "selftest passed\n", e);
    return e ? 1 : 0;
}
