// MT19937 jump-ahead: host polynomial arithmetic over GF(2) and the device doubling
// tree that turns one engine window into evenly spaced checkpoints.  See mt_jump.h.
#include "mt_jump.h"

#include <algorithm>
#include <cstring>
#include <map>
#include <mutex>

#include "kernels.cuh"

namespace {

using Bits = std::vector<uint64_t>;
constexpr int PW = 313; /* 64-bit words holding degrees 0..19937(+) */

inline bool get_bit(const Bits &b, size_t i) { return (b[i >> 6] >> (i & 63)) & 1; }
inline void flip_bit(Bits &b, size_t i) { b[i >> 6] ^= 1ull << (i & 63); }

/* dst ^= src << shift, src of `sw` words, dst long enough */
void xor_shifted(Bits &dst, const uint64_t *src, size_t sw, size_t shift) {
    const size_t ws = shift >> 6, bs = shift & 63;
    if (dst.size() < ws + sw + 1) dst.resize(ws + sw + 1, 0);
    if (bs == 0) {
        for (size_t i = 0; i < sw; ++i) dst[ws + i] ^= src[i];
    } else {
        uint64_t carry = 0;
        for (size_t i = 0; i < sw; ++i) {
            dst[ws + i] ^= (src[i] << bs) | carry;
            carry = src[i] >> (64 - bs);
        }
        dst[ws + sw] ^= carry;
    }
}

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
    const int W = PW + 2;
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

/* a (any length) -> a mod phi, in PW words */
void reduce(Bits &a) {
    const Phi &P = phi();
    for (;;) {
        /* strip trailing zero words */
        size_t top = a.size();
        while (top > 0 && a[top - 1] == 0) --top;
        if (top == 0) break;
        const size_t deg = (top - 1) * 64 + (63 - __builtin_clzll(a[top - 1]));
        if (deg < (size_t)MTJ_DEG) break;
        /* H = a >> DEG */
        const size_t hbits = deg - MTJ_DEG + 1;
        const size_t hw = (hbits + 63) / 64;
        Bits H(hw + 1, 0);
        const size_t ws = MTJ_DEG >> 6, bs = MTJ_DEG & 63;
        for (size_t i = 0; i < hw; ++i) {
            uint64_t lo = (ws + i < a.size()) ? a[ws + i] >> bs : 0;
            uint64_t hi = (bs && ws + i + 1 < a.size()) ? a[ws + i + 1] << (64 - bs) : 0;
            H[i] = lo | hi;
        }
        /* clear bits >= DEG */
        a.resize(std::max(a.size(), ws + 1));
        a[ws] &= (bs ? ((1ull << bs) - 1) : 0);
        for (size_t i = ws + 1; i < a.size(); ++i) a[i] = 0;
        for (int e : P.low_terms) xor_shifted(a, H.data(), hw, (size_t)e);
    }
    a.resize(PW, 0);
}

Bits square(const Bits &a) {
    static uint16_t spread[256];
    static bool init = false;
    if (!init) {
        for (int v = 0; v < 256; ++v) {
            uint16_t r = 0;
            for (int b = 0; b < 8; ++b)
                if (v & (1 << b)) r |= (uint16_t)(1u << (2 * b));
            spread[v] = r;
        }
        init = true;
    }
    Bits out(2 * a.size(), 0);
    for (size_t i = 0; i < a.size(); ++i) {
        uint64_t lo = 0, hi = 0;
        for (int b = 0; b < 4; ++b) {
            lo |= (uint64_t)spread[(a[i] >> (8 * b)) & 0xff] << (16 * b);
            hi |= (uint64_t)spread[(a[i] >> (32 + 8 * b)) & 0xff] << (16 * b);
        }
        out[2 * i] = lo;
        out[2 * i + 1] = hi;
    }
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

std::vector<uint16_t> positions(const Bits &g) {
    std::vector<uint16_t> p;
    for (int i = 0; i < MTJ_DEG; ++i)
        if (get_bit(g, (size_t)i)) p.push_back((uint16_t)i);
    return p;
}

struct Chain {
    std::vector<Bits> polys;
    std::vector<std::vector<uint16_t>> pos;
};
std::mutex g_mu;
std::map<uint64_t, Chain> g_chains;

}  // namespace

const std::vector<uint16_t> &mt_jump_poly(uint64_t unit_words, int j) {
    std::lock_guard<std::mutex> lk(g_mu);
    Chain &c = g_chains[unit_words];
    if (c.polys.empty()) {
        c.polys.push_back(pow_x(unit_words));
        c.pos.push_back(positions(c.polys.back()));
    }
    while ((int)c.polys.size() <= j) {
        c.polys.push_back(square(c.polys.back()));
        c.pos.push_back(positions(c.polys.back()));
    }
    return c.pos[j];
}

bool mt_jump_host(uint32_t *window, uint64_t n) {
    if (!phi().ok) return false;
    const Bits g = pow_x(n);
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
    return true;
}

/* ------------------------------------------------------------------ device side */

/* Extend each source window into MTJ_EXT_LEN raw words of its sequence. */
__global__ void __launch_bounds__(GEN_THREADS) k_jump_extend(const uint32_t *ck, uint32_t *ext) {
    __shared__ uint32_t ring[RING_WORDS];
    const int t = threadIdx.x;
    const uint32_t *win = ck + (size_t)blockIdx.x * MT_N;
    uint32_t *e = ext + (size_t)blockIdx.x * MTJ_EXT_LEN;
    for (int i = t; i < MT_N; i += GEN_THREADS) {
        const uint32_t v = win[i];
        ring[48 + i] = v;
        e[i] = v;
    }
    const ModConst one = {1u, 0xffffffffu, 0u};
    const GenAddr g = make_gen_addr(t, one, one);
    __syncthreads();
    uint32_t *o = e + MT_N + t;
    for (int grp = 0; grp < MTJ_GROUPS; grp += 4) {
        o[0] = gen_raw<0>(ring, g);
        __syncthreads();
        if (grp + 1 < MTJ_GROUPS) o[GEN_THREADS] = gen_raw<1>(ring, g);
        __syncthreads();
        if (grp + 2 < MTJ_GROUPS) o[2 * GEN_THREADS] = gen_raw<2>(ring, g);
        __syncthreads();
        if (grp + 3 < MTJ_GROUPS) o[3 * GEN_THREADS] = gen_raw<3>(ring, g);
        __syncthreads();
        o += 4 * GEN_THREADS;
    }
}

/* dst[b][j] ^= XOR over this block's slice of the polynomial's set bits i of ext[b][i + j]. */
__global__ void __launch_bounds__(640) k_jump_conv(const uint32_t *__restrict__ ext, const uint16_t *__restrict__ pos,
                                                   uint32_t npos, uint32_t splits, uint32_t *dst) {
    const uint32_t b = blockIdx.x / splits, sp = blockIdx.x % splits;
    const uint32_t j = threadIdx.x;
    if (j >= MT_N) return;
    const uint32_t lo = (uint32_t)((uint64_t)npos * sp / splits), hi = (uint32_t)((uint64_t)npos * (sp + 1) / splits);
    const uint32_t *e = ext + (size_t)b * MTJ_EXT_LEN + j;
    uint32_t acc0 = 0, acc1 = 0;
    uint32_t q = lo;
    for (; q + 4 <= hi; q += 4) {
        const uint32_t p0 = pos[q], p1 = pos[q + 1], p2 = pos[q + 2], p3 = pos[q + 3];
        acc0 ^= e[p0] ^ e[p1];
        acc1 ^= e[p2] ^ e[p3];
    }
    for (; q < hi; ++q) acc0 ^= e[pos[q]];
    atomicXor(dst + (size_t)b * MT_N + j, acc0 ^ acc1);
}

int MtJump::run(cudaStream_t stream, const uint32_t *base_window, uint64_t count, uint64_t unit_words, int j,
                uint32_t *d_ck, uint64_t *launches, std::string *err) {
#define JCU(call)                                                     \
    do {                                                              \
        cudaError_t e__ = (call);                                     \
        if (e__ != cudaSuccess) {                                     \
            if (err) *err = std::string(#call) + ": " + cudaGetErrorString(e__); \
            return -1;                                                \
        }                                                             \
    } while (0)
    if (!phi().ok) {
        if (err) *err = "characteristic polynomial computation failed";
        return -1;
    }
    JCU(cudaMemcpyAsync(d_ck, base_window, sizeof(uint32_t) * MT_N, cudaMemcpyHostToDevice, stream));
    if (count <= 1) {
        JCU(cudaStreamSynchronize(stream)); /* base_window may be pageable */
        return 0;
    }
    JCU(cudaMemsetAsync(d_ck + MT_N, 0, sizeof(uint32_t) * MT_N * (count - 1), stream));
    int sms = 148;
    {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    }
    int levels = 0;
    while ((1ull << levels) < count) ++levels;
    uint64_t need_src = 1; /* the widest level of the doubling tree */
    for (int k = 0; k < levels; ++k) need_src = std::max<uint64_t>(need_src, std::min<uint64_t>(1ull << k, count - (1ull << k)));
    if (ext_cap < need_src) {
        if (d_ext) cudaFree(d_ext);
        d_ext = nullptr;
        ext_cap = 0;
        const uint64_t want = std::max<uint64_t>(need_src, 64);
        JCU(cudaMalloc(&d_ext, sizeof(uint32_t) * MTJ_EXT_LEN * want));
        ext_cap = want;
    }
    for (int k = 0; k < levels; ++k) {
        const uint64_t have = 1ull << k;
        const uint64_t nsrc = std::min<uint64_t>(have, count - have);
        /* polynomial x^(unit * 2^(j+k)) */
        DevPoly *dp = nullptr;
        for (auto &p : polys)
            if (p.unit == unit_words && p.j == j + k) dp = &p;
        if (!dp) {
            const std::vector<uint16_t> &pv = mt_jump_poly(unit_words, j + k);
            DevPoly np;
            np.unit = unit_words;
            np.j = j + k;
            np.count = (uint32_t)pv.size();
            np.pos = nullptr;
            JCU(cudaMalloc(&np.pos, sizeof(uint16_t) * (pv.size() + 8)));
            JCU(cudaMemcpyAsync(np.pos, pv.data(), sizeof(uint16_t) * pv.size(), cudaMemcpyHostToDevice, stream));
            JCU(cudaStreamSynchronize(stream));
            polys.push_back(np);
            dp = &polys.back();
        }
        k_jump_extend<<<(unsigned)nsrc, GEN_THREADS, 0, stream>>>(d_ck, d_ext);
        uint64_t splits = (2ull * sms + nsrc - 1) / nsrc;
        splits = std::max<uint64_t>(1, std::min<uint64_t>(splits, 64));
        k_jump_conv<<<(unsigned)(nsrc * splits), 640, 0, stream>>>(d_ext, dp->pos, dp->count, (uint32_t)splits,
                                                                  d_ck + (size_t)have * MT_N);
        if (launches) *launches += 2;
        JCU(cudaGetLastError());
    }
    JCU(cudaStreamSynchronize(stream)); /* base_window may be pageable; also surfaces kernel faults here */
    return 0;
#undef JCU
}

void MtJump::release() {
    for (auto &p : polys)
        if (p.pos) cudaFree(p.pos);
    polys.clear();
    if (d_ext) cudaFree(d_ext);
    d_ext = nullptr;
    ext_cap = 0;
}
