// MT19937 jump-ahead for the model stream ("mode S at scale").
//
// The reference's batched inference consumes ONE std::mt19937 sequentially
// (monte-carlo-model.hh:58-67).  To let hundreds of CTAs each start in the middle
// of that stream, the engine state must be advanced by D draws in O(1) kernels:
// with phi the characteristic polynomial of the MT19937 transition (degree 19937)
// and g = x^D mod phi,   X[D + j] = XOR_{i : g_i = 1} X[i + j]   for the raw state
// word sequence X.  Host side: phi by Berlekamp–Massey on the engine's own output,
// chains P_j = x^(unit * 2^j) mod phi by repeated squaring (cached per unit).
// Device side: a doubling tree — level k turns windows 0..2^k-1 into windows
// 2^k..2^(k+1)-1 with the single polynomial P_(j+k).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <vector>

constexpr int MTJ_DEG = 19937;
constexpr int MTJ_GROUPS = 90;                       /* 90 * 224 = 20160 >= 19937 extension words */
constexpr int MTJ_EXT_LEN = 624 + MTJ_GROUPS * 224;  /* words of sequence per source window */

/* x^n mod phi applied to a 624-word window on the host (any n).  false on failure. */
bool mt_jump_host(uint32_t *window, uint64_t n);

/* Set-bit exponents (ascending) of x^(unit_words * 2^j) mod phi; computed once and cached. */
const std::vector<uint16_t> &mt_jump_poly(uint64_t unit_words, int j);

struct MtJump {
    /* device copies of the position lists, keyed by (unit, j) */
    struct DevPoly {
        uint64_t unit;
        int j;
        uint16_t *pos;
        uint32_t count;
    };
    std::vector<DevPoly> polys;
    uint32_t *d_ext = nullptr;
    size_t ext_cap = 0; /* windows */

    /* Fill d_ck[c] (c = 0..count-1) with the window preceding draw c * unit_words * 2^j,
     * d_ck[0] being `base_window` itself (host pointer). */
    int run(cudaStream_t stream, const uint32_t *base_window, uint64_t count, uint64_t unit_words, int j,
            uint32_t *d_ck, uint64_t *launches, std::string *err);
    void release();
};
