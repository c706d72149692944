// xorwow_host.h — host restatement of cuRAND's XORWOW generator as the reference uses it.
//
// TEST INFRASTRUCTURE ONLY (part of the CPU oracle).
//
// The reference seeds one generator per pixel with curand_init(seed, /*subsequence*/pixel, /*offset*/0)
// (reference: CudaRayTracer/src/CUDAKernels.h:18-26) and draws with curand_uniform
// (src/Core/Utility.cuh:18-26). cuRAND is a third-party dependency that is not under /root/reference:
// it is the CUDA toolkit's device API (toolkit 12.9 here; upstream names v12.5 only as a Windows path,
// premake5.lua:16). The algorithm is Marsaglia's xorwow (2003) plus cuRAND's documented seeding and
// 2^67-element subsequence spacing (curand_kernel.h: _curand_init_inplace, _skipahead_sequence_inplace,
// curand(), _curand_uniform). This file restates it WITHOUT using cuRAND's precomputed tables: the
// 160x160 GF(2) one-step matrix is built from the recurrence and raised to 2^67 by squaring, so the
// toolkit's table (curand_precalc.h) can serve as an independent known answer in tests
// (tests/test_oracle_rng.py), next to the device known answers in tests/golden/rngkat.txt.
#pragma once
#include <array>
#include <cmath>
#include <cstdint>
#include <mutex>
#include <vector>

namespace orc {

struct Xorwow {
    uint32_t v[5];
    uint32_t d;
};

// curand(): one xorwow step + Weyl sequence (curand_kernel.h, curand(curandStateXORWOW_t*))
static inline uint32_t xorwow_next(Xorwow& s) {
    uint32_t t = s.v[0] ^ (s.v[0] >> 2);
    s.v[0] = s.v[1];
    s.v[1] = s.v[2];
    s.v[2] = s.v[3];
    s.v[3] = s.v[4];
    s.v[4] = (s.v[4] ^ (s.v[4] << 4)) ^ (t ^ (t << 1));
    s.d += 362437u;
    return s.v[4] + s.d;
}

// curand_uniform(): x * 2^-32 + 2^-33 evaluated as ONE fma (that is what the device code does:
// fma.rn.f32 r, x, 0f2F800000, 0f2F000000), result in (0, 1].
static inline float xorwow_uniform(Xorwow& s) {
    return fmaf((float)xorwow_next(s), 2.3283064365386963e-10f, 1.1641532182693481e-10f);
}

// ---- GF(2) 160x160 matrices, row b = image of basis vector e_b, 5 words per row -----------
struct XwMatrix {
    std::array<uint32_t, 160 * 5> w;
};

static inline void xw_apply(const XwMatrix& m, uint32_t v[5]) {
    uint32_t r[5] = {0, 0, 0, 0, 0};
    for (int i = 0; i < 5; i++)
        for (int j = 0; j < 32; j++)
            if (v[i] & (1u << j)) {
                const uint32_t* row = &m.w[(size_t)(i * 32 + j) * 5];
                for (int k = 0; k < 5; k++) r[k] ^= row[k];
            }
    for (int k = 0; k < 5; k++) v[k] = r[k];
}

static inline XwMatrix xw_one_step() {
    XwMatrix m;
    for (int b = 0; b < 160; b++) {
        Xorwow s{};
        s.v[b / 32] = 1u << (b % 32);
        s.d = 0;
        xorwow_next(s);   // linear in v; d does not feed back into v
        for (int k = 0; k < 5; k++) m.w[(size_t)b * 5 + k] = s.v[k];
    }
    return m;
}

static inline XwMatrix xw_square(const XwMatrix& a) {
    XwMatrix r;
    for (int b = 0; b < 160; b++) {
        uint32_t v[5];
        for (int k = 0; k < 5; k++) v[k] = a.w[(size_t)b * 5 + k];
        xw_apply(a, v);
        for (int k = 0; k < 5; k++) r.w[(size_t)b * 5 + k] = v[k];
    }
    return r;
}

// table[k] = T^(2^67 * 4^k), k = 0..31  (cuRAND: PRECALC_BLOCK_SIZE = 2 bits per table entry)
static inline const std::vector<XwMatrix>& xw_sequence_table() {
    static std::vector<XwMatrix> table;
    static std::once_flag once;
    std::call_once(once, [] {
        XwMatrix m = xw_one_step();
        for (int i = 0; i < 67; i++) m = xw_square(m);
        table.reserve(32);
        for (int k = 0; k < 32; k++) {
            table.push_back(m);
            m = xw_square(xw_square(m));
        }
    });
    return table;
}

// curand_init(seed, subsequence, 0, &state)
static inline Xorwow xorwow_init(uint64_t seed, uint64_t subsequence) {
    Xorwow s;
    uint32_t s0 = ((uint32_t)seed) ^ 0xaad26b49u;
    uint32_t s1 = (uint32_t)(seed >> 32) ^ 0xf7dcefddu;
    uint32_t t0 = 1099087573u * s0;
    uint32_t t1 = 2591861531u * s1;
    s.d = 6615241u + t1 + t0;
    s.v[0] = 123456789u + t0;
    s.v[1] = 362436069u ^ t0;
    s.v[2] = 521288629u + t1;
    s.v[3] = 88675123u ^ t1;
    s.v[4] = 5783321u + t0;
    const std::vector<XwMatrix>& tab = xw_sequence_table();
    uint64_t x = subsequence;
    int k = 0;
    while (x) {
        for (uint32_t t = 0; t < (uint32_t)(x & 3u); t++) xw_apply(tab[(size_t)k], s.v);
        x >>= 2;
        k++;
    }
    return s;   // d needs no update: the skipped distance is a multiple of 2^32
}

}  // namespace orc
