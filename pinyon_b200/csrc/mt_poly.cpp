// GF(2)[x] arithmetic modulo the characteristic polynomial of MT19937 (host side of the
// jump-ahead, see mt_jump.h).  Plain C++ compiled by the host compiler; carry-less
// multiplies use PCLMULQDQ when the CPU has it (checked at run time) and a portable
// shift-and-xor loop otherwise.
#include <algorithm>
#include <cstring>
#include <map>
#include <mutex>
#include <thread>

#include "mt19937.cuh"
#include "mt_jump.h"

#if defined(__x86_64__)
#include <immintrin.h>
#define MTP_X86 1
#endif

namespace {

using Bits = std::vector<uint64_t>;
constexpr int PW = 312;  /* 64-bit words holding degrees 0..19967 */
constexpr int TOPW = MTJ_DEG >> 6, TOPB = MTJ_DEG & 63; /* bit 19937 = word 311, bit 33 */

inline bool get_bit(const Bits &b, size_t i) { return (b[i >> 6] >> (i & 63)) & 1; }

/* dst ^= src << shift, truncated to dst.size() words (same-size operands) */
void xor_shl_fixed(Bits &dst, const Bits &src, int shift) {
    const int ws = shift >> 6, bs = shift & 63, W = (int)dst.size();
    for (int i = W - 1; i >= ws; --i) {
        uint64_t v = src[i - ws] << bs;
        if (bs && i - ws - 1 >= 0) v |= src[i - ws - 1] >> (64 - bs);
        dst[i] ^= v;
    }
}

struct Phi {
    std::vector<int> low_terms; /* exponents e < 19937 with phi_e = 1 */
    bool ok = false;
};

/* Characteristic polynomial of the MT19937 transition by Berlekamp–Massey over GF(2)
 * on 2*19937 bits of the engine's own (tempered) output; tempering and bit selection
 * are linear, and phi is irreducible, so any non-zero bit stream has minimal
 * polynomial phi. */
Phi compute_phi() {
    const int N = 2 * MTJ_DEG + 64;
    std::vector<uint8_t> s(N);
    {
        std::vector<uint32_t> x(MT_N);
        x[0] = 5489u;
        for (uint32_t i = 1; i < MT_N; ++i) x[i] = mt_init_step(x[i - 1], i);
        int produced = 0;
        while (produced < N) {
            for (uint32_t k = 0; k < MT_N; ++k) x[k] = mt_twist(x[k], x[(k + 1) % MT_N], x[(k + MT_M) % MT_N]);
            for (uint32_t k = 0; k < MT_N && produced < N; ++k) s[produced++] = mt_temper(x[k]) & 1u;
        }
    }
    const int W = PW + 3;
    Bits C(W, 0), B(W, 0), T(W, 0), R(W, 0);
    C[0] = 1;
    B[0] = 1;
    int L = 0, m = 1;
    for (int n = 0; n < N; ++n) {
        /* R bit i = s[n - i] */
        uint64_t carry = s[n];
        const int rw = std::min(W, (n >> 6) + 2);
        for (int i = 0; i < rw; ++i) {
            const uint64_t nc = R[i] >> 63;
            R[i] = (R[i] << 1) | carry;
            carry = nc;
        }
        uint64_t acc = 0;
        const int lw = (L >> 6) + 1;
        for (int i = 0; i < lw; ++i) acc ^= C[i] & R[i];
        const int d = __builtin_parityll(acc);
        if (!d) {
            ++m;
        } else if (2 * L <= n) {
            T = C;
            xor_shl_fixed(C, B, m);
            L = n + 1 - L;
            B = T;
            m = 1;
        } else {
            xor_shl_fixed(C, B, m);
            ++m;
        }
    }
    Phi phi;
    if (L != MTJ_DEG) return phi;
    /* phi_k = C_{L-k} */
    for (int k = 0; k < MTJ_DEG; ++k)
        if (get_bit(C, (size_t)(L - k))) phi.low_terms.push_back(k);
    phi.ok = get_bit(C, 0); /* leading coefficient */
    return phi;
}

const Phi &phi() {
    static const Phi p = compute_phi();
    return p;
}

/* a (up to 2*PW words) -> a mod phi in PW words.  One top-down pass: the terms of phi
 * below x^19937 start at x^19314, so folding word i only touches words at least nine
 * below it. */
void reduce(Bits &a) {
    const Phi &P = phi();
    if (a.size() < (size_t)PW + 1) a.resize(PW + 1, 0);
    for (int i = (int)a.size() - 1; i >= TOPW; --i) {
        uint64_t h = a[i];
        if (i == TOPW) h &= ~((1ull << TOPB) - 1); /* keep degrees below 19937 */
        if (!h) continue;
        a[i] ^= h;
        /* bit b of h is degree 64*i + b = 19937 + (64*i + b - 19937): add h * x^(64*i - 19937) * (phi - x^19937) */
        const int base = 64 * i - MTJ_DEG; /* >= -33 */
        for (int e : P.low_terms) {
            const int sh = base + e; /* position of h's bit 0; may be negative only when i == TOPW and then h's low bits are 0 */
            if (sh >= 0) {
                const int w = sh >> 6, b = sh & 63;
                a[w] ^= h << b;
                if (b) a[w + 1] ^= h >> (64 - b);
            } else {
                a[0] ^= h >> (-sh);
            }
        }
    }
    a.resize(PW);
}

Bits square(const Bits &a) {
    Bits out(2 * a.size() + 1, 0);
#if defined(MTP_X86) && defined(__BMI2__)
    for (size_t i = 0; i < a.size(); ++i) {
        out[2 * i] = _pdep_u64(a[i], 0x5555555555555555ull);
        out[2 * i + 1] = _pdep_u64(a[i] >> 32, 0x5555555555555555ull);
    }
#else
    static const std::vector<uint16_t> spread = [] {
        std::vector<uint16_t> t(256);
        for (int v = 0; v < 256; ++v) {
            uint16_t r = 0;
            for (int b = 0; b < 8; ++b)
                if (v & (1 << b)) r |= (uint16_t)(1u << (2 * b));
            t[v] = r;
        }
        return t;
    }();
    for (size_t i = 0; i < a.size(); ++i) {
        uint64_t lo = 0, hi = 0;
        for (int b = 0; b < 4; ++b) {
            lo |= (uint64_t)spread[(a[i] >> (8 * b)) & 0xff] << (16 * b);
            hi |= (uint64_t)spread[(a[i] >> (32 + 8 * b)) & 0xff] << (16 * b);
        }
        out[2 * i] = lo;
        out[2 * i + 1] = hi;
    }
#endif
    reduce(out);
    return out;
}

#if defined(MTP_X86)
__attribute__((target("pclmul,sse2"))) void mul_clmul(const uint64_t *a, const uint64_t *b, uint64_t *out, int n) {
    /* schoolbook over 64-bit limbs: out[i+j], out[i+j+1] ^= clmul(a[i], b[j]) */
    for (int i = 0; i < n; ++i) {
        if (!a[i]) continue;
        const __m128i va = _mm_set_epi64x(0, (long long)a[i]);
        uint64_t carry = 0;
        for (int j = 0; j < n; ++j) {
            const __m128i p = _mm_clmulepi64_si128(va, _mm_set_epi64x(0, (long long)b[j]), 0x00);
            const uint64_t lo = (uint64_t)_mm_cvtsi128_si64(p);
            const uint64_t hi = (uint64_t)_mm_cvtsi128_si64(_mm_unpackhi_epi64(p, p));
            out[i + j] ^= lo ^ carry;
            carry = hi;
        }
        out[i + n] ^= carry;
    }
}
#endif

void mul_portable(const uint64_t *a, const uint64_t *b, uint64_t *out, int n) {
    for (int i = 0; i < n; ++i) {
        uint64_t w = a[i];
        while (w) {
            const int bit = __builtin_ctzll(w);
            w &= w - 1;
            if (bit == 0) {
                for (int j = 0; j < n; ++j) out[i + j] ^= b[j];
            } else {
                uint64_t carry = 0;
                for (int j = 0; j < n; ++j) {
                    out[i + j] ^= (b[j] << bit) | carry;
                    carry = b[j] >> (64 - bit);
                }
                out[i + n] ^= carry;
            }
        }
    }
}

Bits mul(const Bits &a, const Bits &b) {
    Bits out(2 * PW + 1, 0);
#if defined(MTP_X86)
    static const bool have_clmul = __builtin_cpu_supports("pclmul");
    if (have_clmul)
        mul_clmul(a.data(), b.data(), out.data(), PW);
    else
#endif
        mul_portable(a.data(), b.data(), out.data(), PW);
    reduce(out);
    return out;
}

Bits mulx(const Bits &a) {
    Bits out(a.size() + 1, 0);
    uint64_t carry = 0;
    for (size_t i = 0; i < a.size(); ++i) {
        out[i] = (a[i] << 1) | carry;
        carry = a[i] >> 63;
    }
    out[a.size()] = carry;
    reduce(out);
    return out;
}

Bits pow_x(uint64_t n) {
    Bits r(PW, 0);
    r[0] = 1;
    if (n == 0) return r;
    for (int b = 63 - __builtin_clzll(n); b >= 0; --b) {
        r = square(r);
        if ((n >> b) & 1) r = mulx(r);
    }
    return r;
}

/* Per stride: x^(c * stride) mod phi for c = 0, 1, 2, ..., grown on demand. */
struct Family {
    std::vector<Bits> g;
    uint64_t last_use = 0;
};
std::mutex g_mu;
std::map<uint64_t, Family> g_families;
uint64_t g_tick = 0;
size_t g_family_cap_bytes = size_t(256) << 20; /* host cache of polynomial families; least recently used go first */

/* Called with g_mu held, after `keep` was grown: drop least-recently-used families until the cache fits. */
void evict_families(uint64_t keep) {
    auto bytes = [] {
        size_t b = 0;
        for (auto &kv : g_families) b += kv.second.g.size() * PW * sizeof(uint64_t);
        return b;
    };
    while (g_families.size() > 1 && bytes() > g_family_cap_bytes) {
        auto lru = g_families.end();
        for (auto it = g_families.begin(); it != g_families.end(); ++it)
            if (it->first != keep && (lru == g_families.end() || it->second.last_use < lru->second.last_use)) lru = it;
        if (lru == g_families.end()) break;
        g_families.erase(lru);
    }
}

void grow_family(Family &f, uint64_t stride, uint64_t count) {
    if (f.g.empty()) {
        Bits one(PW, 0);
        one[0] = 1;
        f.g.push_back(one);
    }
    if (f.g.size() < 2 && count > 1) f.g.push_back(pow_x(stride));
    if (f.g.size() >= count) return;
    const uint64_t have = f.g.size();
    f.g.resize(count);
    /* g[2k] = g[k]^2 (cheap), g[2k+1] = g[2k] * g[1]: everything in [2^m, 2^(m+1)) depends only
     * on entries below 2^m, so each such level is filled by independent tasks. */
    unsigned hw = std::thread::hardware_concurrency();
    const unsigned nthreads = std::max(1u, std::min(16u, hw ? hw : 1u));
    uint64_t lo = have;
    while (lo < count) {
        uint64_t level_end = 1;
        while (level_end <= lo) level_end <<= 1; /* first power of two above lo */
        const uint64_t hi = std::min<uint64_t>(count, level_end);
        /* tasks: even c in [lo, hi) with its odd successor */
        auto work = [&](unsigned tid) {
            for (uint64_t c = (lo + 1) & ~1ull; c < hi; c += 2) {
                if (((c >> 1) % nthreads) != tid) continue;
                f.g[c] = square(f.g[c >> 1]);
                if (c + 1 < hi) f.g[c + 1] = mul(f.g[c], f.g[1]);
            }
        };
        if (lo & 1) f.g[lo] = mul(f.g[lo - 1], f.g[1]); /* odd start: its even predecessor exists */
        if (hi - lo > 8 && nthreads > 1) {
            std::vector<std::thread> th;
            for (unsigned t = 0; t < nthreads; ++t) th.emplace_back(work, t);
            for (auto &t : th) t.join();
        } else {
            for (unsigned t = 0; t < nthreads; ++t) work(t);
        }
        lo = hi;
    }
}

}  // namespace

bool mt_jump_ready() { return phi().ok; }

bool mt_jump_polys(uint64_t stride, uint64_t first, uint64_t count, std::vector<uint32_t> &bitmaps) {
    if (!phi().ok) return false;
    std::lock_guard<std::mutex> lk(g_mu);
    Family &f = g_families[stride];
    grow_family(f, stride, first + count);
    f.last_use = ++g_tick;
    evict_families(stride);
    bitmaps.resize((size_t)count * MTJ_POLY_WORDS);
    for (uint64_t c = 0; c < count; ++c)
        memcpy(bitmaps.data() + (size_t)c * MTJ_POLY_WORDS, f.g[first + c].data(), sizeof(uint32_t) * MTJ_POLY_WORDS);
    return true;
}

void mt_jump_set_host_cache_bytes(size_t bytes) {
    std::lock_guard<std::mutex> lk(g_mu);
    g_family_cap_bytes = bytes;
}

static void apply_poly(const uint64_t *g, uint32_t *window);

bool mt_jump_host_family(uint32_t *window, uint64_t stride, uint64_t c) {
    if (!phi().ok) return false;
    std::lock_guard<std::mutex> lk(g_mu);
    Family &f = g_families[stride];
    grow_family(f, stride, c + 1);
    f.last_use = ++g_tick;
    evict_families(stride);
    apply_poly(f.g[c].data(), window);
    return true;
}

bool mt_jump_host(uint32_t *window, uint64_t n) {
    if (!phi().ok) return false;
    const Bits g = pow_x(n);
    apply_poly(g.data(), window);
    return true;
}

static void apply_poly(const uint64_t *gw, uint32_t *window) {
    Bits g(gw, gw + PW);
    std::vector<uint32_t> ext(MT_N + MTJ_DEG + 1);
    memcpy(ext.data(), window, sizeof(uint32_t) * MT_N);
    for (size_t i = MT_N; i < ext.size(); ++i) ext[i] = mt_twist(ext[i - MT_N], ext[i - MT_N + 1], ext[i - MT_LAG]);
    uint32_t out[MT_N];
    memset(out, 0, sizeof(out));
    for (int i = 0; i < MTJ_DEG; ++i) {
        if (!get_bit(g, (size_t)i)) continue;
        const uint32_t *e = ext.data() + i;
        for (int j = 0; j < MT_N; ++j) out[j] ^= e[j];
    }
    memcpy(window, out, sizeof(out));
}
