// MT19937 jump-ahead, device side: one extension of the base window and one launch that
// applies the polynomial x^(c * stride) mod phi to it for every checkpoint c.  The GF(2)
// polynomial arithmetic lives in mt_poly.cpp.  See mt_jump.h.
#include "mt_jump.h"

#include <algorithm>
#include <chrono>
#include <cstring>

#include "kernels.cuh"

/* ------------------------------------------------------------------ device side */

/* Extend the base window into MTJ_EXT_LEN raw words of its sequence (one CTA). */
__global__ void __launch_bounds__(GEN_THREADS) k_jump_extend(const uint32_t *base, uint32_t *ext, uint32_t *ck0) {
    __shared__ uint32_t ring[RING_WORDS];
    const int t = threadIdx.x;
    for (int i = t; i < MT_N; i += GEN_THREADS) {
        const uint32_t v = base[i];
        ring[48 + i] = v;
        ext[i] = v;
        if (ck0) ck0[i] = v; /* checkpoint 0 of an unsharded batch is the base window itself */
    }
    const ModConst one = {1u, 0xffffffffu, 0u};
    const GenAddr g = make_gen_addr(t, one, one);
    __syncthreads();
    uint32_t *o = ext + MT_N + t;
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

/* window_c[j] = XOR over set coefficients i of g_c of ext[i + j], j = 0..623.
 *
 * The extended sequence (82 KB) and g_c's bitmap sit in shared memory.  A thread owns
 * JA_OUT = 10 CONSECUTIVE outputs (63 threads cover 624; the loads of one instruction are 2-way bank
 * conflicted at stride 10, which the idle LSU pipe absorbs) and walks the bitmap one 32-bit word at a time:
 * 41 loads give it ext[32w + j0 .. 32w + j0 + 40], after which every set bit k of the word is
 * 10 register XORs with static indices.  The bitmap word is the same for the whole warp, so
 * the per-bit branch is warp-uniform.  The 624 bitmap words are split over `splits` CTAs x SEGS
 * two-warp segments; segments are folded through shared memory, CTAs through atomicXor. */
constexpr int JA_OUT = 10;
constexpr int JA_ACTIVE = (MT_N + JA_OUT - 1) / JA_OUT; /* 63 of the segment's 64 lanes (r1g: 13 x 48 left a quarter idle) */
constexpr int JA_SEG_THREADS = 64;
constexpr int JA_SEGS = 8;
constexpr int JA_EXT_WORDS = 20608; /* >= 32 * 623 + 13 * 47 + 44, multiple of 4 */
static_assert(JA_EXT_WORDS <= MTJ_EXT_LEN, "extension too short");
static_assert(JA_ACTIVE * JA_OUT >= MT_N && JA_ACTIVE <= JA_SEG_THREADS, "outputs must cover the window");
static_assert(JA_EXT_WORDS >= 32 * (MT_N - 1) + JA_OUT * (JA_ACTIVE - 1) + 32 + JA_OUT - 1, "extension too short for the last lane");
constexpr size_t JA_SMEM = sizeof(uint32_t) * (JA_EXT_WORDS + MTJ_POLY_WORDS + JA_SEGS * MT_N);

/* acc[m] ^= XOR of r[m + b] over the set bits b of the 4-bit pattern P: terms are folded two at a time
 * into three-input LOP3s (1, 1, 2, 2 instructions per output for 1..4 set bits; 1.25 per nibble on average
 * against 1.5 for two pairs), and one branch serves four coefficients instead of two. */
template <unsigned P>
__device__ __forceinline__ void ja_nibble(uint32_t (&acc)[JA_OUT], const uint32_t *r) {
    constexpr int b0 = (P & 1u) ? 0 : (P & 2u) ? 1 : (P & 4u) ? 2 : 3;
    constexpr unsigned P1 = P & (P - 1u);
    constexpr int b1 = (P1 & 2u) ? 1 : (P1 & 4u) ? 2 : 3;
    constexpr unsigned P2 = P1 & (P1 - 1u);
    constexpr int b2 = (P2 & 4u) ? 2 : 3;
    constexpr unsigned P3 = P2 & (P2 - 1u);
#pragma unroll
    for (int m = 0; m < JA_OUT; ++m) {
        if (P1 == 0)
            acc[m] ^= r[m + b0];
        else
            acc[m] ^= r[m + b0] ^ r[m + b1];
        if (P2 != 0) {
            if (P3 == 0)
                acc[m] ^= r[m + b2];
            else
                acc[m] ^= r[m + b2] ^ r[m + 3];
        }
    }
}

__global__ void __launch_bounds__(JA_SEGS *JA_SEG_THREADS, 2)
    k_jump_apply(const uint32_t *__restrict__ ext, const uint32_t *__restrict__ polys, uint32_t splits, uint32_t *d_ck,
                 uint32_t skip, uint32_t pstep, uint32_t regular, uint32_t extra_rel) {
    extern __shared__ __align__(16) uint32_t ja_smem[];
    uint32_t *sext = ja_smem;
    uint32_t *spoly = ja_smem + JA_EXT_WORDS;
    uint32_t *sred = spoly + MTJ_POLY_WORDS;
    /* window c <- bitmap c * pstep of the family (the one window behind the regular ones: bitmap extra_rel) */
    const uint32_t c = blockIdx.x / splits + skip, sp = blockIdx.x % splits;
    const uint32_t pc = c < regular ? c * pstep : extra_rel;
    const int tid = threadIdx.x;
    const uint32_t cta_lo = MTJ_POLY_WORDS * sp / splits, cta_hi = MTJ_POLY_WORDS * (sp + 1) / splits;
    {
        /* only the part of the sequence this CTA's bitmap words can reach */
        const int q_lo = (int)(32 * cta_lo) / 4, q_hi = min(JA_EXT_WORDS / 4, (int)(32 * cta_hi + JA_OUT * JA_ACTIVE + 32) / 4);
        const uint4 *src = reinterpret_cast<const uint4 *>(ext);
        uint4 *dst = reinterpret_cast<uint4 *>(sext);
        for (int i = q_lo + tid; i < q_hi; i += JA_SEGS * JA_SEG_THREADS) dst[i] = src[i];
        const uint32_t *pp = polys + (size_t)pc * MTJ_POLY_WORDS;
        for (int i = cta_lo + tid; i < (int)cta_hi; i += JA_SEGS * JA_SEG_THREADS) spoly[i] = pp[i];
    }
    __syncthreads();
    const int seg = tid / JA_SEG_THREADS, t = tid % JA_SEG_THREADS;
    const uint32_t len = cta_hi - cta_lo;
    const uint32_t w_lo = cta_lo + len * seg / JA_SEGS, w_hi = cta_lo + len * (seg + 1) / JA_SEGS;
    uint32_t acc[JA_OUT];
#pragma unroll
    for (int m = 0; m < JA_OUT; ++m) acc[m] = 0;
    if (t < JA_ACTIVE) {
        const uint32_t *e = sext + JA_OUT * t;
        for (uint32_t w = w_lo; w < w_hi; ++w) {
            const uint32_t pw = spoly[w];
            if (pw == 0) continue;
            const uint32_t *p = e + 32 * w;
            uint32_t r[32 + JA_OUT - 1];
#pragma unroll
            for (int k = 0; k < 32 + JA_OUT - 1; ++k) r[k] = p[k];
#pragma unroll
            for (int k = 0; k < 32; k += 4) { /* four coefficients at a time behind one warp-uniform 16-way branch */
                /* an explicit tree of single-bit tests on pw: one LOP3-to-predicate per level, no field extraction */
#define JA_B(b) (pw & ((1u << (b)) << k))
#define JA_L(P) ja_nibble<P>(acc, r + k)
                if (JA_B(3)) {
                    if (JA_B(2)) {
                        if (JA_B(1)) { if (JA_B(0)) JA_L(15); else JA_L(14); }
                        else { if (JA_B(0)) JA_L(13); else JA_L(12); }
                    } else {
                        if (JA_B(1)) { if (JA_B(0)) JA_L(11); else JA_L(10); }
                        else { if (JA_B(0)) JA_L(9); else JA_L(8); }
                    }
                } else {
                    if (JA_B(2)) {
                        if (JA_B(1)) { if (JA_B(0)) JA_L(7); else JA_L(6); }
                        else { if (JA_B(0)) JA_L(5); else JA_L(4); }
                    } else {
                        if (JA_B(1)) { if (JA_B(0)) JA_L(3); else JA_L(2); }
                        else { if (JA_B(0)) JA_L(1); }
                    }
                }
#undef JA_B
#undef JA_L
            }
        }
#pragma unroll
        for (int m = 0; m < JA_OUT; ++m)
            if (JA_OUT * t + m < MT_N) sred[seg * MT_N + JA_OUT * t + m] = acc[m];
    }
    __syncthreads();
    for (int j = tid; j < MT_N; j += JA_SEGS * JA_SEG_THREADS) {
        uint32_t v = 0;
#pragma unroll
        for (int s = 0; s < JA_SEGS; ++s) v ^= sred[s * MT_N + j];
        if (splits == 1)
            d_ck[(size_t)c * MT_N + j] = v;
        else
            atomicXor(d_ck + (size_t)c * MT_N + j, v);
    }
}

#define JCU(call)                                                     \
    do {                                                              \
        cudaError_t e__ = (call);                                     \
        if (e__ != cudaSuccess) {                                     \
            if (err) *err = std::string(#call) + ": " + cudaGetErrorString(e__); \
            return -1;                                                \
        }                                                             \
    } while (0)

/* The device copy of the polynomial family (stride; c = first .. first + count - 1), built and uploaded on a miss. */
static int family_on_device(MtJump &J, cudaStream_t stream, uint64_t first, uint64_t count, uint64_t stride,
                            MtJump::DevFamily **out, std::string *err) {
    J.last_build_ms = 0;
    if (!mt_jump_ready()) {
        if (err) *err = "characteristic polynomial computation failed";
        return -1;
    }
    MtJump::DevFamily *fam = nullptr;
    for (auto &f : J.families)
        if (f.stride == stride && f.first == first) fam = &f;
    if (!fam || fam->count < count) {
        const auto t0 = std::chrono::steady_clock::now();
        std::vector<uint32_t> bits;
        const uint64_t want = std::max<uint64_t>(count, fam ? fam->count : 0);
        if (!mt_jump_polys(stride, first, want, bits)) {
            if (err) *err = "jump polynomial computation failed";
            return -1;
        }
        if (!fam) {
            while (J.families.size() >= std::max<size_t>(1, J.max_families)) { /* evict the least recently used family */
                size_t lru = 0;
                for (size_t i = 1; i < J.families.size(); ++i)
                    if (J.families[i].last_use < J.families[lru].last_use) lru = i;
                JCU(cudaStreamSynchronize(stream));
                cudaFree(J.families[lru].bits);
                J.families.erase(J.families.begin() + lru);
            }
            J.families.push_back(MtJump::DevFamily{stride, first, 0, nullptr, 0});
            fam = &J.families.back();
        } else {
            JCU(cudaStreamSynchronize(stream));
            cudaFree(fam->bits);
            fam->bits = nullptr;
            fam->count = 0;
        }
        JCU(cudaMalloc(&fam->bits, sizeof(uint32_t) * bits.size()));
        JCU(cudaMemcpy(fam->bits, bits.data(), sizeof(uint32_t) * bits.size(), cudaMemcpyHostToDevice));
        fam->count = want;
        J.last_build_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    }
    fam->last_use = ++J.tick;
    *out = fam;
    return 0;
}

/* Bitmaps a run over (first, count, pstep, extra) touches: first .. the largest index it applies. */
static uint64_t family_span(uint64_t first, uint64_t count, uint64_t pstep, int64_t extra) {
    uint64_t hi = first + (count ? (count - 1) * pstep : 0);
    if (extra >= 0) hi = std::max<uint64_t>(hi, (uint64_t)extra);
    return hi - first + 1;
}

int MtJump::prepare(cudaStream_t stream, uint64_t first, uint64_t count, uint64_t stride, std::string *err, uint64_t pstep,
                    int64_t extra) {
    if (first == 0 && count <= 1 && extra < 0) return 0;
    DevFamily *fam = nullptr;
    return family_on_device(*this, stream, first, family_span(first, count, pstep, extra), stride, &fam, err);
}

int MtJump::run(cudaStream_t stream, int sms, const uint32_t *base_window, uint64_t first, uint64_t count, uint64_t stride,
                uint32_t *d_ck, uint64_t *launches, std::string *err, uint64_t pstep, int64_t extra) {
    if (first == 0 && count <= 1 && extra < 0) {
        JCU(cudaMemcpyAsync(d_ck, base_window, sizeof(uint32_t) * MT_N, cudaMemcpyHostToDevice, stream));
        last_build_ms = 0;
        return 0;
    }
    DevFamily *fam = nullptr;
    if (family_on_device(*this, stream, first, family_span(first, count, pstep, extra), stride, &fam, err)) return -1;
    if (!d_ext) {
        JCU(cudaMalloc(&d_ext, sizeof(uint32_t) * (MTJ_EXT_LEN + MT_N)));
        JCU(cudaFuncSetAttribute(k_jump_apply, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)JA_SMEM));
    }
    uint32_t *d_base = d_ext + MTJ_EXT_LEN;
    JCU(cudaMemcpyAsync(d_base, base_window, sizeof(uint32_t) * MT_N, cudaMemcpyHostToDevice, stream));
    const uint32_t skip = first == 0 ? 1u : 0u; /* window 0 of an unsharded batch needs no polynomial */
    const uint64_t windows = count + (extra >= 0 ? 1 : 0);
    const uint64_t apply = windows - skip;
    if (apply == 0) { /* a single window which is the base itself */
        JCU(cudaMemcpyAsync(d_ck, base_window, sizeof(uint32_t) * MT_N, cudaMemcpyHostToDevice, stream));
        return 0;
    }
    /* few checkpoints: split each polynomial over several CTAs so the launch still fills the GPU */
    uint64_t splits = (2ull * sms + apply - 1) / apply;
    splits = std::max<uint64_t>(1, std::min<uint64_t>(splits, 13));
    if (splits > 1) JCU(cudaMemsetAsync(d_ck + (size_t)MT_N * skip, 0, sizeof(uint32_t) * MT_N * apply, stream));
    k_jump_extend<<<1, GEN_THREADS, 0, stream>>>(d_base, d_ext, skip ? d_ck : nullptr);
    k_jump_apply<<<(unsigned)(apply * splits), JA_SEGS * JA_SEG_THREADS, JA_SMEM, stream>>>(
        d_ext, fam->bits, (uint32_t)splits, d_ck, skip, (uint32_t)pstep, (uint32_t)count,
        extra >= 0 ? (uint32_t)((uint64_t)extra - first) : 0u);
    if (launches) *launches += 2;
    JCU(cudaGetLastError());
    return 0;
}
#undef JCU

void MtJump::release() {
    for (auto &f : families)
        if (f.bits) cudaFree(f.bits);
    families.clear();
    if (d_ext) cudaFree(d_ext);
    d_ext = nullptr;
}
