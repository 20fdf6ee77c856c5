// sm_100a rollout kernels of libpinyon_cuda.  See DESIGN.md "Kernels".
//
// k_stream2     (stream.cuh) fused hot path: warp-per-stream MT19937 model generator
//               feeding consumer warps that play one rollout per lane against a
//               shared-memory table of the (shared) state stream.
// k_gen_codes2  (stream.cuh) the same generator writing action codes to global memory,
//               for batches the fused kernel does not take (ragged depth, private states).
// k_warp        general path: one rollout per warp, the state's private MT19937 in
//               shared memory, advanced 32 words per step by the warp.
// The 224-thread CTA-wide generator below (GenAddr, gen_raw) only serves the jump-ahead's
// k_jump_extend.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "mt19937.cuh"

#define PCU_FULL 0xffffffffu
#define PCU_HASH_INIT 2166136261u
#define PCU_HASH_MUL 16777619u

constexpr int GEN_THREADS = 224;           /* 7 warps; <= 227 so one group has no internal dependency */
constexpr int RING_WORDS = 4 * GEN_THREADS; /* 896 >= 624 + 224: four group slots */
constexpr int STEPS_PER_GROUP = GEN_THREADS / 4; /* a step = 2 random_int = 4 words */

enum { KIND_TREE = 0, KIND_MOLD = 1 };

__device__ __forceinline__ uint32_t hash_step(uint32_t h, uint32_t joint, uint32_t chance) {
    return (h ^ (joint | (chance << 16))) * PCU_HASH_MUL;
}

__device__ __forceinline__ unsigned long long warp_sum_u64(unsigned long long v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(PCU_FULL, v, o);
    return v;
}

/* Per-thread constants of the 224-wide generator: ring slots for the four phases. */
struct GenAddr {
    uint16_t ra[4], rb[4], rc[4], wa[4];
    uint32_t n, m, c32; /* modulus of this lane's word: rows for words 0,1 of a step, cols for 2,3 */
};

__device__ __forceinline__ GenAddr make_gen_addr(int t, const ModConst &mrow, const ModConst &mcol) {
    GenAddr g;
#pragma unroll
    for (int ph = 0; ph < 4; ++ph) {
        const int off = GEN_THREADS * ph + t;
        g.ra[ph] = (uint16_t)((off + 48) % RING_WORDS);  /* X[n-624] */
        g.rb[ph] = (uint16_t)((off + 49) % RING_WORDS);  /* X[n-623] */
        g.rc[ph] = (uint16_t)((off + 445) % RING_WORDS); /* X[n-227] */
        g.wa[ph] = (uint16_t)((off + 672) % RING_WORDS); /* X[n]     */
    }
    const ModConst &mc = (t & 2) ? mcol : mrow;
    g.n = mc.n;
    g.m = mc.m;
    g.c32 = mc.c32;
    return g;
}

__device__ __forceinline__ void load_checkpoint(uint32_t *ring, const uint32_t *ckpt, int t) {
    for (int i = t; i < MT_N; i += GEN_THREADS) ring[48 + i] = ckpt[i];
}

/* The raw (untempered) word only — used by the jump-ahead extension kernel. */
template <int PH>
__device__ __forceinline__ uint32_t gen_raw(uint32_t *ring, const GenAddr &g) {
    const uint32_t a = ring[g.ra[PH]], b = ring[g.rb[PH]], c = ring[g.rc[PH]];
    const uint32_t x = mt_twist(a, b, c);
    ring[g.wa[PH]] = x;
    return x;
}

#include "stream.cuh"

/* ----------------------------------------------------------------------- k_warp */

struct WarpParams {
    uint64_t n;
    uint32_t rows, cols, T, gen_words;
    ModConst mrow, mcol;
    const uint64_t *state_seed;
    uint64_t state_seed_u;
    const uint32_t *state_pos;
    uint32_t state_pos_u;
    const uint64_t *tseed;
    uint64_t tseed_u;
    const uint16_t *depth;
    uint16_t depth_u;
    const int32_t *bias;
    int32_t bias_u;
    int32_t thr_p, thr_q;
    const double *fq;
    uint32_t fq_w;
    const uint64_t *model_seeds;  /* mode P */
    const void *codes;            /* mode S: joint action codes */
    const uint64_t *step_off;     /* mode S: first step of rollout i (null: i * depth_u) */
    int32_t code_u16;
    uint8_t *payoff;
    uint8_t *traj;
    uint32_t traj_stride;
    uint32_t *traj_hash;
    unsigned long long *sums;
    /* slab of rollouts [i0, i1) seeded by k_seed_pre (seed.cuh): arrays indexed by i - i0 */
    uint64_t i0, i1;
    const uint32_t *xs_slab, *xm_slab;
    const double *pu_slab;
};

#include "lane.cuh"

/* A private MT19937 held in shared memory by one warp.  x[] is libstdc++'s state
 * array; base is the draw index of x[0] (-624 while x still holds the seeded array). */
struct WarpMT {
    uint32_t *x;
    long long base;
};

/* One refill of the warp's engine (bits/random.tcc _M_gen_rand), in place: x[w] <- x[w+397 mod 624]
 * ^ twist(x[w], x[w+1 mod 624]) for w = 0..623 in order.  Five groups of <= 128 words; a group reads
 * every operand before it writes anything (its writes would clobber the `b` operand of the word
 * before them), and since a group is shorter than 227 words the x[w-227] operands of later groups are
 * already NEW.  All addresses are lane + constant after unrolling; the only per-lane selects are the
 * sub-block where w+397 crosses 624 (w = 224..255) and the half sub-block at the end. */
__device__ __forceinline__ void wmt_twist(WarpMT &e, int lane) {
    uint32_t *xl = e.x + lane;
#pragma unroll
    for (int g = 0; g < 5; ++g) {
        uint32_t a[4], b[4], c[4];
#pragma unroll
        for (int m = 0; m < 4; ++m) {
            const int w0 = g * 128 + m * 32;
            if (w0 >= MT_N) continue;
            if (w0 + 32 <= MT_N) {
                a[m] = xl[w0];
                b[m] = xl[w0 + 1];
                if (w0 + 31 < MT_LAG)
                    c[m] = xl[w0 + MT_M];
                else if (w0 >= MT_LAG)
                    c[m] = xl[w0 - MT_LAG];
                else
                    c[m] = xl[(w0 + lane < MT_LAG) ? w0 + MT_M : w0 - MT_LAG];
            } else { /* w0 = 608: words 608..623 on lanes 0..15; b of word 623 is the new x[0] */
                a[m] = b[m] = c[m] = 0;
                if (lane < MT_N - w0) {
                    a[m] = xl[w0];
                    b[m] = (w0 + lane + 1 == MT_N) ? e.x[0] : xl[w0 + 1];
                    c[m] = xl[w0 - MT_LAG];
                }
            }
        }
        __syncwarp();
#pragma unroll
        for (int m = 0; m < 4; ++m) {
            const int w0 = g * 128 + m * 32;
            if (w0 >= MT_N) continue;
            if (w0 + 32 <= MT_N || lane < MT_N - w0) xl[w0] = mt_twist_dev(a[m], b[m], c[m]);
        }
        __syncwarp();
    }
    e.base += MT_N;
}

/* Tempered draw number `abs` for every active lane.  Draws are requested in
 * non-decreasing order, so lanes that cannot be served yet wait for the next block. */
__device__ __forceinline__ uint32_t wmt_fetch(WarpMT &e, long long abs, bool active, int lane) {
    uint32_t val = 0;
    bool done = !active;
    for (;;) {
        const long long off = abs - e.base;
        if (!done && off < MT_N) {
            val = mt_temper(e.x[off]);
            done = true;
        }
        if (__all_sync(PCU_FULL, done)) break;
        wmt_twist(e, lane);
    }
    return val;
}

/* Inputs of the small-batch kernel k_small (MODE 2), all passed BY VALUE in the launch parameters so that a
 * call of <= 32 rollouts is one launch and one synchronisation: the model engine in libstdc++'s form and the
 * per-rollout state fields.  Replaces Model::inference(State&&, ModelOutput&) (monte-carlo-model.hh:41-56). */
struct SmallExtra {
    uint32_t win[MT_N]; /* model engine words */
    uint32_t win_p;     /* next-word position 0..624 */
    uint32_t n;
    uint32_t mulc;      /* 1812433253 as data (lane.cuh) */
    uint32_t pad;
    uint64_t state_seed[32], tseed[32];
    uint32_t pos[32];
    int32_t bias[32];
    uint16_t depth[32];
};

/* MODE 0: mode S, action codes from k_gen_codes2.  MODE 1: mode P, model engines seeded per rollout.
 * MODE 2: mode S for <= 32 rollouts in ONE CTA (warp w = rollout w): every warp owns a copy of the model
 * engine and walks it to its rollout's offset; the state engines are seeded in the kernel, lane-per-rollout. */
template <int KIND, int MODE>
__device__ __forceinline__ void warp_body(const WarpParams &P, const SmallExtra *X) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31;
    const int wib = threadIdx.x >> 5;
    const int warps_per_block = blockDim.x >> 5;
    constexpr int ENGINES = (KIND == KIND_TREE ? 1 : 0) + (MODE >= 1 ? 1 : 0);
    /* MODE 2 pads a warp's region by one word: lane r then seeds engine r without bank conflicts */
    constexpr size_t WSTRIDE = (size_t)MT_N * (ENGINES ? ENGINES : 1) + (MODE == 2 ? 1 : 0);
    uint32_t *mine = reinterpret_cast<uint32_t *>(smem_raw) + (size_t)wib * WSTRIDE;
    uint32_t *xs = mine;                                       /* state engine */
    uint32_t *xm = mine + (KIND == KIND_TREE ? MT_N : 0);      /* model engine (modes P and 2) */
    const ModConst m10 = {10u, 429496729u, 6u};
    const ModConst m3 = {3u, 1431655765u, 1u};
    double *pu_s = reinterpret_cast<double *>(smem_raw + (((size_t)warps_per_block * WSTRIDE * 4 + 7) & ~(size_t)7));
    /* fq[n][S] = (double)(float)(n / S) (random-tree.hh:122-133 through math.hh:30-33) behind the engines */
    double *fq_s = pu_s + (MODE == 2 ? 32 : 0);
    if (KIND == KIND_TREE && P.T > 1)
        for (uint32_t k = threadIdx.x; k < 11u * P.fq_w; k += blockDim.x) fq_s[k] = P.fq[k];
    if (MODE == 2) {
        for (int k = lane; k < MT_N; k += 32) xm[k] = X->win[k];
        if (KIND == KIND_TREE && wib == 0 && lane < (int)X->n) {
            uint32_t *dst = reinterpret_cast<uint32_t *>(smem_raw) + (size_t)lane * WSTRIDE;
            uint32_t v = (uint32_t)X->state_seed[lane];
            dst[0] = v;
#pragma unroll 8
            for (uint32_t k = 1; k < MT_N; ++k) {
                v = lane_seed_step(v, X->mulc, k);
                dst[k] = v;
            }
        }
        if (KIND == KIND_TREE && P.T > 1 && wib == (warps_per_block > 1 ? 1 : 0) && lane < (int)X->n) {
            uint32_t v = (uint32_t)X->tseed[lane];
            const uint32_t t0 = v;
            uint32_t t1 = 0, t2 = 0, t397 = 0;
#pragma unroll 8
            for (uint32_t k = 1; k <= 398; ++k) {
                v = lane_seed_step(v, X->mulc, k);
                if (k == 1) t1 = v;
                if (k == 2) t2 = v;
                if (k == 397) t397 = v;
            }
            pu_s[lane] = prng_uniform_from_words(mt_temper(mt_twist(t0, t1, t397)), mt_temper(mt_twist(t1, t2, v)));
        }
    }
    if (MODE == 2 || (KIND == KIND_TREE && P.T > 1)) __syncthreads();

    unsigned long long acc_w = 0, acc_d = 0, acc_l = 0, acc_h = 0, acc_s = 0;
    const uint64_t warp0 = (uint64_t)blockIdx.x * warps_per_block + wib;
    const uint64_t stride = (uint64_t)gridDim.x * warps_per_block;

    for (uint64_t i = P.i0 + warp0; i < P.i1; i += stride) {
        uint32_t d;
        double pu = 0.0;
        long long moff = 0; /* MODE 2: model draws consumed by the rollouts before this one */
        if (MODE == 2) {
            d = X->depth[i];
            for (uint32_t j = 0; j < (uint32_t)i; ++j) moff += 4ll * X->depth[j];
            if (KIND == KIND_TREE && P.T > 1) pu = pu_s[i];
        } else {
            d = P.depth ? P.depth[i] : P.depth_u;
            /* ---- the seeded engines of this rollout, 20 coalesced 128-byte reads each */
            __syncwarp();
            if (KIND == KIND_TREE) {
                const uint32_t *src = P.xs_slab + (i - P.i0) * MT_N;
#pragma unroll 5
                for (int k = lane; k < MT_N; k += 32) xs[k] = src[k];
            }
            if (MODE == 1) {
                const uint32_t *src = P.xm_slab + (i - P.i0) * MT_N;
#pragma unroll 5
                for (int k = lane; k < MT_N; k += 32) xm[k] = src[k];
            }
            if (KIND == KIND_TREE && P.T > 1) pu = P.pu_slab[i - P.i0];
            __syncwarp();
        }
        WarpMT es = {xs, -(long long)MT_N};
        WarpMT em = {xm, MODE == 2 ? -(long long)X->win_p : -(long long)MT_N};

        long long cur = 0;
        int bias = 0;
        if (KIND == KIND_TREE) {
            if (MODE == 2) {
                cur = (long long)X->pos[i] - (long long)P.gen_words;
                bias = X->bias[i];
            } else {
                cur = (long long)(P.state_pos ? P.state_pos[i] : P.state_pos_u) - (long long)P.gen_words;
                bias = P.bias ? P.bias[i] : P.bias_u;
            }
        }
        const uint64_t soff = (MODE == 0) ? (P.step_off ? P.step_off[i] : i * (uint64_t)P.depth_u) : 0;
        uint32_t h = PCU_HASH_INIT;
        uint8_t *tr = P.traj ? P.traj + (size_t)i * P.traj_stride * 3 : nullptr;

        for (uint32_t s = 0; s < d; ++s) {
            uint32_t r = 0, c = 0, joint;
            if (MODE == 0) {
                /* the joint code is all a step needs; row and column are split off only for a trajectory */
                joint = P.code_u16 ? (uint32_t)((const uint16_t *)P.codes)[soff + s]
                                   : (uint32_t)((const uint8_t *)P.codes)[soff + s];
            } else {
                /* lanes 0..3 draw the four model words of this step (monte-carlo-model.hh:89-90) */
                const int mrel = (int)(moff + 4ll * s - em.base);
                uint32_t w;
                if (mrel >= 0 && mrel + 4 <= MT_N) /* the step's four draws lie in the current block */
                    w = mt_temper(xm[mrel + (lane & 3)]);
                else
                    w = wmt_fetch(em, moff + 4ll * s + lane, lane < 4, lane);
                const ModConst &mc = (lane & 2) ? P.mcol : P.mrow;
                const uint32_t m = mod32(w, mc.n, mc.m);
                const uint32_t lo = __shfl_down_sync(PCU_FULL, m, 1);
                const uint32_t vv = mod32(m * mc.c32 + lo, mc.n, mc.m);
                r = __shfl_sync(PCU_FULL, vv, 0);
                c = __shfl_sync(PCU_FULL, vv, 2);
                joint = r * P.cols + c;
            }
            uint32_t ch = 0;
            if (KIND == KIND_TREE) {
                const uint32_t T = P.T;
                const uint32_t st = joint * T;
                if (T > 1) {
                    /* lanes 0..2T-1 read the 2T words the chosen cell's strategies were drawn
                     * from (random-tree.hh:169); lane k then owns transition k. */
                    const int crel = (int)(cur + 2ll * st - es.base);
                    uint32_t w = 0;
                    if (crel >= 0 && crel + 2 * (int)T <= MT_N) { /* the cell's draws lie in the current block */
                        if (lane < (int)(2 * T)) w = mt_temper(xs[crel + lane]);
                    } else {
                        w = wmt_fetch(es, cur + 2ll * st + lane, lane < (int)(2 * T), lane);
                    }
                    const uint32_t hi = __shfl_sync(PCU_FULL, w, (2 * lane) & 31);
                    const uint32_t lo = __shfl_sync(PCU_FULL, w, (2 * lane + 1) & 31);
                    int nk = 0;
                    if (lane < (int)T) {
                        nk = (int)mod64(hi, lo, m10) + 1;
                        if (nk * P.thr_q < P.thr_p * 10) nk = 0;
                    }
                    int S = __reduce_add_sync(PCU_FULL, nk);
                    if (S == 0) {
                        if (lane == 0) nk = 1;
                        S = 1;
                    }
                    double pp = pu;
                    for (uint32_t k = 0; k < T; ++k) {
                        const int n_k = __shfl_sync(PCU_FULL, nk, k);
                        pp -= fq_s[(uint32_t)n_k * P.fq_w + (uint32_t)S];
                        if (pp <= 0) {
                            ch = k;
                            break;
                        }
                    }
                }
                cur += (long long)P.gen_words + st + ch;
                const int brel = (int)(cur - es.base);
                uint32_t w;
                if (brel >= 0 && brel + 2 <= MT_N)
                    w = mt_temper(xs[brel + (lane & 1)]);
                else
                    w = wmt_fetch(es, cur + lane, lane < 2, lane);
                const uint32_t hi = __shfl_sync(PCU_FULL, w, 0);
                const uint32_t lo = __shfl_sync(PCU_FULL, w, 1);
                bias += (int)mod64(hi, lo, m3) - 1;
                cur += 2;
            }
            h = hash_step(h, joint, ch);
            if (tr && lane == 0 && s < P.traj_stride) { /* later steps are played and hashed, not recorded */
                if (MODE == 0) {
                    r = joint / P.cols;
                    c = joint - r * P.cols;
                }
                tr[3 * s + 0] = (uint8_t)r;
                tr[3 * s + 1] = (uint8_t)c;
                tr[3 * s + 2] = (uint8_t)ch;
            }
        }
        uint32_t pay = 0;
        if (KIND == KIND_TREE) pay = (uint32_t)((bias > 0) - (bias < 0) + 1);
        if (lane == 0) {
            if (tr)
                for (uint32_t s = d; s < P.traj_stride; ++s) tr[3 * s] = tr[3 * s + 1] = tr[3 * s + 2] = 0;
            if (P.payoff) P.payoff[i] = (uint8_t)pay;
            if (P.traj_hash) P.traj_hash[i] = h;
            if (KIND == KIND_TREE) {
                acc_w += (pay == 2);
                acc_d += (pay == 1);
                acc_l += (pay == 0);
            } else {
                acc_d += 1;
            }
            acc_h += h;
            acc_s += d;
        }
        __syncwarp();
    }
    if (P.sums && lane == 0) {
        if (acc_w) atomicAdd(P.sums + 0, acc_w);
        if (acc_d) atomicAdd(P.sums + 1, acc_d);
        if (acc_l) atomicAdd(P.sums + 2, acc_l);
        if (acc_s) atomicAdd(P.sums + 3, acc_s);
        if (acc_h) atomicAdd(P.sums + 4, acc_h);
    }
}

template <int KIND, int MODE>
__global__ void __launch_bounds__(256) k_warp(const WarpParams P) {
    warp_body<KIND, MODE>(P, nullptr);
}

struct SmallParams {
    WarpParams w;
    SmallExtra x;
};

/* <= 32 rollouts, mode S, everything in the launch parameters: one launch, no copies in. */
template <int KIND>
__global__ void __launch_bounds__(1024) k_small(const __grid_constant__ SmallParams SP) {
    warp_body<KIND, 2>(SP.w, &SP.x);
}

#include "seed.cuh"
#include "leaf.cuh"
#include "thread.cuh"
#include "warp2.cuh"

/* ------------------------------------------------------- integer-issue peak probe */

/* VARIANT 0: LOP3 only; 1: LOP3 + SHF + IADD3 (all ALU pipe); 2: LOP3 + SHF + IMAD
 * (ALU + FMA pipes).  Eight chains per thread, cross-coupled so that no two
 * operations can be folded into one; 3 ops per chain per unrolled step. */
__device__ __forceinline__ uint32_t op_lop3(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t d;
    asm("lop3.b32 %0, %1, %2, %3, 0xE8;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ uint32_t op_shf(uint32_t a, uint32_t b) {
    uint32_t d;
    asm("shf.l.wrap.b32 %0, %1, %2, 7;" : "=r"(d) : "r"(a), "r"(b));
    return d;
}
__device__ __forceinline__ uint32_t op_iadd3(uint32_t a, uint32_t b, uint32_t c) { return a + b + c; }
__device__ __forceinline__ uint32_t op_imad(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t d;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

constexpr int INT_PEAK_OPS_PER_ITER = 4 * 8 * 3;

__device__ __forceinline__ uint32_t op_imadhi(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t d;
    asm("mad.hi.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

__device__ __forceinline__ uint32_t op_imadwide_hi(uint32_t a, uint32_t b, uint32_t c) {
    unsigned long long w;
    asm("mad.wide.u32 %0, %1, %2, %3;" : "=l"(w) : "r"(a), "r"(b | 0x10000u), "l"((unsigned long long)c << 20));
    return (uint32_t)(w >> 32);
}

/* Variants 0..2 feed the roofline; 3..6 are pipe probes used to design the generator:
 * 3: IMAD only; 4: IMAD.HI only; 5: LOP3, IMAD, LOP3 ... (2:1); 6: LOP3 + IMAD + IMAD.HI. */
template <int VARIANT>
__global__ void __launch_bounds__(256) k_int_peak(uint32_t *out, uint32_t seed, int iters) {
    uint32_t a[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) a[j] = seed + threadIdx.x * 8 + j + blockIdx.x * 2048;
    const uint32_t k1 = seed | 1u;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 4; ++u) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const uint32_t nb = a[(j + 3) & 7];
                if (VARIANT == 0) {
                    a[j] = op_lop3(a[j], nb, k1);
                    a[j] = op_lop3(a[j], k1, nb);
                    a[j] = op_lop3(nb, a[j], k1);
                } else if (VARIANT == 1) {
                    a[j] = op_lop3(a[j], nb, k1);
                    a[j] = op_shf(a[j], nb);
                    a[j] = op_iadd3(a[j], nb, k1);
                } else if (VARIANT == 2) {
                    a[j] = op_lop3(a[j], nb, k1);
                    a[j] = op_shf(a[j], nb);
                    a[j] = op_imad(a[j], k1, nb);
                } else if (VARIANT == 3) {
                    a[j] = op_imad(a[j], k1, nb);
                    a[j] = op_imad(a[j], nb, k1);
                    a[j] = op_imad(nb, a[j], k1);
                } else if (VARIANT == 4) {
                    a[j] = op_imadhi(a[j], k1, nb);
                    a[j] = op_imadhi(a[j], nb, k1);
                    a[j] = op_imadhi(nb, a[j], k1);
                } else if (VARIANT == 5) {
                    a[j] = op_lop3(a[j], nb, k1);
                    a[j] = op_imad(a[j], k1, nb);
                    a[j] = ((j + u) & 1) ? op_lop3(a[j], k1, nb) : op_imad(a[j], nb, k1);
                } else if (VARIANT == 6) {
                    a[j] = op_lop3(a[j], nb, k1);
                    a[j] = op_imadhi(a[j], k1, nb);
                    a[j] = ((j + u) & 1) ? op_lop3(a[j], k1, nb) : op_imad(a[j], nb, k1);
                } else if (VARIANT == 7) { /* IMAD.WIDE only: a right shift as the high half of a 64-bit product */
                    a[j] = op_imadwide_hi(a[j], k1, nb);
                    a[j] = op_imadwide_hi(a[j], nb, k1);
                    a[j] = op_imadwide_hi(nb, a[j], k1);
                } else if (VARIANT == 8) { /* LOP3 : IMAD.WIDE 2 : 1 */
                    a[j] = op_lop3(a[j], nb, k1);
                    a[j] = op_imadwide_hi(a[j], k1, nb);
                    a[j] = op_lop3(a[j], k1, nb);
                } else { /* 9: SHF only */
                    a[j] = op_shf(a[j], nb);
                    a[j] = op_shf(a[j], k1);
                    a[j] = op_shf(nb, a[j]);
                }
            }
        }
    }
    uint32_t r = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) r ^= a[j];
    if (r == 0x12345678u) out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}
