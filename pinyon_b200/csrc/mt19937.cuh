// MT19937 and Pinyon prng arithmetic shared by host and device code of libpinyon_cuda.
//
// Restates libstdc++'s std::mt19937 (the engine inside Pinyon's prng,
// src/types/random.hh:8-13) as a LINEAR stream: X[0..623] is the seeded array,
// X[n] = X[n-227] ^ twist(X[n-624], X[n-623]) for n >= 624, and draw number m of the
// engine (0-based) is temper(X[624 + m]).  libstdc++'s in-place 624-word refill
// (bits/random.tcc _M_gen_rand) computes exactly this sequence, block by block.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define PCU_HD __host__ __device__ __forceinline__
#else
#define PCU_HD inline
#endif

#define MT_N 624
#define MT_M 397
#define MT_LAG (MT_N - MT_M) /* 227: X[n] depends on X[n-227] */

/* a = X[n-624], b = X[n-623], c = X[n-227] */
PCU_HD uint32_t mt_twist(uint32_t a, uint32_t b, uint32_t c) {
    const uint32_t y = (a & 0x80000000u) | (b & 0x7fffffffu);
    return c ^ (y >> 1) ^ ((b & 1u) ? 0x9908b0dfu : 0u);
}

PCU_HD uint32_t mt_temper(uint32_t y) {
    y ^= y >> 11;
    y ^= (y << 7) & 0x9d2c5680u;
    y ^= (y << 15) & 0xefc60000u;
    y ^= y >> 18;
    return y;
}

/* one step of std::mt19937::seed's initialisation chain */
PCU_HD uint32_t mt_init_step(uint32_t prev, uint32_t i) { return 1812433253u * (prev ^ (prev >> 30)) + i; }

/* x mod n for 32-bit x, n >= 1, with m = floor(2^32 / n) (m = 0xffffffff for n = 1):
 * q = mulhi(x, m) is floor(x/n) or one less, so r = x - q*n lies in [0, 2n). */
struct ModConst {
    uint32_t n, m, c32; /* c32 = 2^32 mod n */
};

inline ModConst make_mod_const(uint32_t n) {
    ModConst c;
    c.n = n;
    c.m = (n <= 1) ? 0xffffffffu : (uint32_t)(0x100000000ull / n);
    c.c32 = (n <= 1) ? 0u : (uint32_t)(0x100000000ull % n);
    return c;
}

#if defined(__CUDACC__)
__device__ __forceinline__ uint32_t mod32(uint32_t x, uint32_t n, uint32_t m) {
    const uint32_t q = __umulhi(x, m);
    const uint32_t r = x - q * n;
    return min(r, r - n);
}
/* (hi:lo) mod n == prng::random_int(n) for the u64 uniform_64() = hi<<32 | lo
 * (random.hh:26,28); valid for n <= 65536. */
__device__ __forceinline__ uint32_t mod64(uint32_t hi, uint32_t lo, const ModConst &c) {
    const uint32_t a = mod32(hi, c.n, c.m);
    const uint32_t b = mod32(lo, c.n, c.m);
    return mod32(a * c.c32 + b, c.n, c.m);
}
/* The same twist written for the B200 pipes (LOP3/SHF on the ALU pipe at 64 lanes/clk/SM,
 * IMAD on the FMA pipe at 64 more): one bit-select LOP3, one shift, (b & 1) * matrix as
 * LOP3 + IMAD, one three-way XOR -> 4 ALU + 1 FMA instructions instead of 6 + 1. */
__device__ __forceinline__ uint32_t mt_twist_dev(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t y, x;
    asm("lop3.b32 %0, %1, %2, %3, 0xE4;" : "=r"(y) : "r"(a), "r"(b), "r"(0x80000000u)); /* (a & m) | (b & ~m) */
    const uint32_t s = y >> 1;
    /* (y & 1) * matrix without touching the ALU pipe: y & 1 = y - 2 (y >> 1), so the product is
     * y * matrix - (y >> 1) * (2 * matrix) in wrapping arithmetic: two multiply-adds on the FMA pipe
     * (the twist is then 3 ALU + 2 FMA instructions; the ALU pipe is the one the generator saturates) */
    uint32_t t, mag;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(t) : "r"(s), "r"(0u - 2u * 0x9908b0dfu), "r"(0u));
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(mag) : "r"(y), "r"(0x9908b0dfu), "r"(t));
    asm("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(x) : "r"(c), "r"(s), "r"(mag));
    return x;
}

/* Per-modulus constants of the folded 64-bit reduction below (n <= 256). */
struct FoldConst {
    uint32_t n, m, c32, k1, k2; /* m = floor(2^32 / n); see mod64_fold */
    uint32_t negn;              /* -n: the remainder is one multiply-add, q * (-n) + s */
};

inline FoldConst make_fold_const(uint32_t n) {
    FoldConst f;
    const uint32_t c16 = (n <= 1) ? 0u : (uint32_t)(0x10000ull % n);
    const uint32_t c48 = (n <= 1) ? 0u : (uint32_t)(0x1000000000000ull % n);
    f.n = n;
    f.m = (n <= 1) ? 0xffffffffu : (uint32_t)(0x100000000ull / n);
    f.c32 = (n <= 1) ? 0u : (uint32_t)(0x100000000ull % n);
    f.k1 = c48 - 65536u * f.c32; /* mod 2^32 */
    f.k2 = c16 - 65536u;
    f.negn = 0u - n;
    return f;
}

/* (hi:lo) mod n == prng::random_int(n) (random.hh:26,28) for n <= 256, with ONE high multiply:
 * with ck = 2^k mod n the four 16-bit digits fold to
 *   s = (hi >> 16) c48 + (hi & 0xffff) c32 + (lo >> 16) c16 + (lo & 0xffff)  < 65535 (3 * 255 + 1) < 2^26,
 * evaluated in wrapping 32-bit arithmetic as (hi >> 16) k1 + hi c32 + (lo >> 16) k2 + lo with
 * k1 = c48 - 2^16 c32, k2 = c16 - 2^16 (two shifts on the ALU pipe, three IMADs on the FMA pipe),
 * and s is reduced once.  IMAD.HI issues at less than half the IMAD rate on sm_100 (measured:
 * 28 vs 64 lanes/clk/SM) and stalls the ALU pipe while it does, so the three of them in the
 * word-by-word reduction (mod64) were the most expensive part of a model step. */
__device__ __forceinline__ uint32_t mod64_fold(uint32_t hi, uint32_t lo, const FoldConst &f) {
    const uint32_t s = (hi >> 16) * f.k1 + hi * f.c32 + (lo >> 16) * f.k2 + lo;
    const uint32_t q = __umulhi(s, f.m); /* floor(s / n) or one less */
    uint32_t r;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(q), "r"(f.negn), "r"(s));
    return min(r, r + f.negn);
}

/* prng::uniform() (random.hh:23): generate_canonical<double,53> over a 32-bit engine;
 * first draw is the LOW word; one rounding (in the add); clamp below 1. */
__device__ __forceinline__ double prng_uniform_from_words(uint32_t first, uint32_t second) {
    const double s = __dadd_rn((double)first, __dmul_rn((double)second, 4294967296.0));
    double r = __dmul_rn(s, 5.421010862427522170037e-20); /* exact 2^-64 */
    if (r >= 1.0) r = 0.99999999999999988897769753748; /* nextafter(1, 0) */
    return r;
}
#endif

inline double prng_uniform_from_words_host(uint32_t first, uint32_t second) {
    volatile double s = (double)first + (double)second * 4294967296.0;
    double r = s / 18446744073709551616.0;
    if (r >= 1.0) r = 0.99999999999999988897769753748;
    return r;
}
