// Warp-per-stream MT19937 generator and the fused mode-S rollout kernel built on it.
//
// Generator ("gw"): ONE WARP owns one model stream.  MT19937 as a linear recurrence
//   X[n] = X[n-227] ^ f(X[n-624], X[n-623])
// allows 227 independent words per step; a warp produces 224 = 56 model steps at a time,
// lane l (of 28 active lanes) making the FOUR words of step l and of step 28 + l:
//   - operands come from the warp's private 896-word shared-memory ring as 128-bit loads
//     (consecutive lanes read consecutive 16 B: conflict-free), results go back as one
//     128-bit store per step; ring slots are compile-time constants plus 4 * lane;
//   - both random_int draws of a step (monte-carlo-model.hh:89-90; uniform_64 draws HIGH
//     word first, random.hh:26,28) live in one thread: no shuffles, no redundant lanes;
//   - no CTA barrier anywhere: a group ends with __syncwarp().
//
// Fused kernel (k_stream2): CTA = GW_PAIRS generator warps + GW_PAIRS consumer warps.
// Generator g streams joint action codes (row * cols + col) into a double-buffered
// shared-memory queue; consumer g (same SM sub-partition: warp id + GW_PAIRS) plays
// RB rollouts per buffer, one or two per lane, against the shared state-stream table.
// Hand-off is by named barriers (bar.sync / bar.arrive, 64 threads), so the consumer's
// latency-bound table walks overlap the generator's ALU work on the same scheduler.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "mt19937.cuh"

#ifndef PCU_FULL
#define PCU_FULL 0xffffffffu
#endif

constexpr int GW_ACTIVE = 28;             /* generator lanes in use */
constexpr int GW_STEPS = 2 * GW_ACTIVE;   /* model steps per group */
constexpr int GW_WORDS = 4 * GW_STEPS;    /* 224 <= 227 */
constexpr int GW_RING = 4 * GW_WORDS;     /* 896 >= 624 + 224: four group slots */
constexpr int GW_PAIRS = 4;               /* generator/consumer warp pairs per CTA */

/* Ring slot of (compile-time offset C) + 4 * lane; the wrap test is only emitted for the
 * few offsets whose 28 lanes straddle the end of the ring. */
template <int C>
__device__ __forceinline__ int gw_slot(int l4) {
    constexpr int c = C % GW_RING;
    if constexpr (c + 4 * (GW_ACTIVE - 1) + 3 < GW_RING) {
        return c + l4;
    } else {
        const int s = c + l4;
        return s >= GW_RING ? s - GW_RING : s;
    }
}

/* The four words of one model step: stream word n' (counted from the start of the current
 * 896-word pass) lives in ring slot (n' + 672) mod 896, so X[n-624] is at n' + 48 and
 * X[n-227] at n' + 445.  Returns the tempered words. */
template <int PH, int H>
__device__ __forceinline__ void gw_step_words(uint32_t *ring, int l4, uint32_t &z0, uint32_t &z1, uint32_t &z2,
                                              uint32_t &z3) {
    constexpr int off = GW_WORDS * PH + (GW_WORDS / 2) * H;
    const uint4 a = *reinterpret_cast<const uint4 *>(ring + gw_slot<off + 48>(l4));
    const uint32_t a4 = ring[gw_slot<off + 52>(l4)];
    const uint4 c = *reinterpret_cast<const uint4 *>(ring + gw_slot<off + 444>(l4)); /* X[n-228 .. n-225] */
    const uint32_t c4 = ring[gw_slot<off + 448>(l4)];
    uint4 x;
    x.x = mt_twist(a.x, a.y, c.y);
    x.y = mt_twist(a.y, a.z, c.z);
    x.z = mt_twist(a.z, a.w, c.w);
    x.w = mt_twist(a.w, a4, c4);
    *reinterpret_cast<uint4 *>(ring + gw_slot<off + 672>(l4)) = x;
    z0 = mt_temper(x.x);
    z1 = mt_temper(x.y);
    z2 = mt_temper(x.z);
    z3 = mt_temper(x.w);
}

template <int PH, int H>
__device__ __forceinline__ uint32_t gw_step_code(uint32_t *ring, int l4, const ModConst &mrow, const ModConst &mcol,
                                                 uint32_t cols) {
    uint32_t z0, z1, z2, z3;
    gw_step_words<PH, H>(ring, l4, z0, z1, z2, z3);
    const uint32_t row = mod64(z0, z1, mrow);
    const uint32_t col = mod64(z2, z3, mcol);
    return row * cols + col;
}

__device__ __forceinline__ void gw_load_window(uint32_t *ring, const uint32_t *ckpt, int lane) {
    for (int i = lane; i < MT_N; i += 32) ring[48 + i] = ckpt[i];
}

/* The 624 raw words preceding stream word `w_end` of the current 896-word pass. */
__device__ __forceinline__ void gw_dump_window(const uint32_t *ring, uint32_t w_end, uint32_t *dst, int lane) {
    for (int i = lane; i < MT_N; i += 32) dst[i] = ring[(w_end + 48 + i) % GW_RING];
}

__device__ __forceinline__ void nb_sync(int id) { asm volatile("bar.sync %0, 64;" ::"r"(id) : "memory"); }
__device__ __forceinline__ void nb_arrive(int id) {
    __threadfence_block();
    asm volatile("bar.arrive %0, 64;" ::"r"(id) : "memory");
}

/* ----------------------------------------------------------------- fused kernel */

struct Stream2Params {
    uint64_t n;       /* rollouts */
    uint64_t chunk;   /* rollouts per stream */
    uint32_t streams; /* ceil(n / chunk) */
    uint32_t depth;   /* uniform depth_bound / max_depth */
    uint32_t rb;      /* rollouts per queue buffer: 32 or 64 */
    uint32_t cols, T, gen_words;
    ModConst mrow, mcol;
    const uint32_t *checkpoints; /* [streams][624] raw words preceding each stream's first model word */
    const uint8_t *tab;          /* state-stream table: low nibble u64%10, bits 4-5 u64%3 at each draw index */
    uint32_t tab_base, tab_len;
    const uint32_t *state_pos;
    uint32_t state_pos_u;
    const int32_t *bias;
    int32_t bias_u;
    double pu;
    int32_t thr_p, thr_q;
    const double *fq; /* fq[n * fq_w + S] = (double)(float)(n / (double)S) */
    uint32_t fq_w, fq_len;
    uint8_t *payoff;
    uint8_t *traj;
    uint32_t traj_stride;
    uint32_t *traj_hash;
    unsigned long long *sums;
    uint32_t *end_window; /* [624] raw words preceding the first model word after the batch */
};

struct RolloutAcc {
    unsigned long long w = 0, d = 0, l = 0, h = 0, n = 0;
};

/* One generator warp: `total_steps` model steps from its checkpoint into the code queue
 * [2][B].  Buffer k (absolute) uses half k & 1; barriers: full[half] = bar_base + half,
 * empty[half] = bar_base + 2 + half. */
template <typename CodeT>
__device__ __forceinline__ void gw_run(uint32_t *ring, CodeT *queue, uint32_t B, uint64_t total_steps, int bar_base,
                                       const ModConst &mrow, const ModConst &mcol, uint32_t cols, int lane,
                                       uint32_t *end_window) {
    const int l4 = 4 * lane;
    const bool active = lane < GW_ACTIVE;
    const uint64_t groups = (total_steps + GW_STEPS - 1) / GW_STEPS;
    const uint64_t nbuf = (total_steps + B - 1) / B; /* buffers the consumer will read */
    uint32_t fill = 0, half = 0;
    uint64_t buf = 0;      /* absolute index of the buffer being filled */
    uint64_t produced = 0; /* steps emitted before the current group */

    auto emit = [&](uint32_t c0, uint32_t c1) {
        /* entering a buffer for the first time: wait until the consumer has drained that half
         * (buffers 0 and 1 start empty; nothing is waited for or written past the last buffer) */
        if (fill == 0 && buf >= 2) nb_sync(bar_base + 2 + half);
        if (fill + GW_STEPS > B && buf + 1 >= 2 && buf + 1 < nbuf) nb_sync(bar_base + 2 + (half ^ 1));
        if (active) {
            uint32_t i0 = fill + lane, i1 = fill + GW_ACTIVE + lane;
            CodeT *q0 = queue + half * B, *q1 = q0;
            if (i0 >= B) {
                i0 -= B;
                q0 = queue + (half ^ 1) * B;
            }
            if (i1 >= B) {
                i1 -= B;
                q1 = queue + (half ^ 1) * B;
            }
            if (produced + lane < total_steps) q0[i0] = (CodeT)c0;
            if (produced + GW_ACTIVE + lane < total_steps) q1[i1] = (CodeT)c1;
        }
        fill += GW_STEPS;
        produced += GW_STEPS;
        if (fill >= B) {
            nb_arrive(bar_base + half);
            fill -= B;
            half ^= 1;
            ++buf;
        }
    };

    uint64_t g = 0;
    for (; g + 4 <= groups; g += 4) {
        uint32_t c0 = 0, c1 = 0;
        if (active) {
            c0 = gw_step_code<0, 0>(ring, l4, mrow, mcol, cols);
            c1 = gw_step_code<0, 1>(ring, l4, mrow, mcol, cols);
        }
        __syncwarp();
        emit(c0, c1);
        if (active) {
            c0 = gw_step_code<1, 0>(ring, l4, mrow, mcol, cols);
            c1 = gw_step_code<1, 1>(ring, l4, mrow, mcol, cols);
        }
        __syncwarp();
        emit(c0, c1);
        if (active) {
            c0 = gw_step_code<2, 0>(ring, l4, mrow, mcol, cols);
            c1 = gw_step_code<2, 1>(ring, l4, mrow, mcol, cols);
        }
        __syncwarp();
        emit(c0, c1);
        if (active) {
            c0 = gw_step_code<3, 0>(ring, l4, mrow, mcol, cols);
            c1 = gw_step_code<3, 1>(ring, l4, mrow, mcol, cols);
        }
        __syncwarp();
        emit(c0, c1);
    }
    const uint32_t tail = (uint32_t)(groups - g); /* 0..3 groups in a last, partial pass */
    if (tail > 0) {
        uint32_t c0 = 0, c1 = 0;
        if (active) {
            c0 = gw_step_code<0, 0>(ring, l4, mrow, mcol, cols);
            c1 = gw_step_code<0, 1>(ring, l4, mrow, mcol, cols);
        }
        __syncwarp();
        emit(c0, c1);
        if (tail > 1) {
            if (active) {
                c0 = gw_step_code<1, 0>(ring, l4, mrow, mcol, cols);
                c1 = gw_step_code<1, 1>(ring, l4, mrow, mcol, cols);
            }
            __syncwarp();
            emit(c0, c1);
        }
        if (tail > 2) {
            if (active) {
                c0 = gw_step_code<2, 0>(ring, l4, mrow, mcol, cols);
                c1 = gw_step_code<2, 1>(ring, l4, mrow, mcol, cols);
            }
            __syncwarp();
            emit(c0, c1);
        }
    }
    /* a last, partly filled buffer; a group that overshot into a buffer beyond the last one the
     * consumer will read leaves fill > 0 there, which is harmless */
    if (fill > 0 && buf < nbuf) nb_arrive(bar_base + half);
    if (end_window) {
        /* words generated in the current pass before the stream's end: 4 * total_steps mod 896,
         * except that a stream ending exactly on a pass boundary sits at 896 */
        const uint64_t words = 4 * total_steps;
        uint32_t w_end = (uint32_t)(words % GW_RING);
        if (w_end == 0 && words > 0) w_end = GW_RING;
        gw_dump_window(ring, w_end, end_window, lane);
    }
}

/* Consumer side: the per-rollout walk over the shared state stream. */
struct TreeWalk {
    uint32_t p, gb, h;
    int bias;
};

template <int KIND, typename CodeT, bool TAB_SMEM>
__global__ void __launch_bounds__(GW_PAIRS * 64) k_stream2(const Stream2Params P) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t d = P.depth;
    const uint32_t B = P.rb * d; /* codes per queue buffer */
    /* layout: rings [PAIRS][896] u32 | fq doubles | queues [PAIRS][2][B] CodeT | table bytes */
    uint32_t *rings = reinterpret_cast<uint32_t *>(smem_raw);
    double *sfq = reinterpret_cast<double *>(rings + GW_PAIRS * GW_RING);
    CodeT *queues = reinterpret_cast<CodeT *>(sfq + P.fq_len);
    const uint8_t *tab = P.tab;
    if (KIND == KIND_TREE) {
        for (uint32_t i = tid; i < P.fq_len; i += blockDim.x) sfq[i] = P.fq[i];
        if (TAB_SMEM) {
            uint8_t *stab = reinterpret_cast<uint8_t *>(queues + (size_t)GW_PAIRS * 2 * B);
            stab = reinterpret_cast<uint8_t *>(((uintptr_t)stab + 15) & ~(uintptr_t)15);
            for (uint32_t i = tid; i < P.tab_len; i += blockDim.x) stab[i] = P.tab[i];
            tab = stab;
        }
    }
    const int pair = warp % GW_PAIRS;
    const bool is_gen = warp < GW_PAIRS;
    const uint32_t stream = blockIdx.x * GW_PAIRS + pair;
    uint32_t *ring = rings + pair * GW_RING;
    CodeT *queue = queues + (size_t)pair * 2 * B;
    if (stream < P.streams && is_gen) gw_load_window(ring, P.checkpoints + (size_t)stream * MT_N, lane);
    __syncthreads();
    if (stream >= P.streams) return;

    const uint64_t first = (uint64_t)stream * P.chunk;
    const uint64_t cnt = (first + P.chunk <= P.n) ? P.chunk : P.n - first;
    const uint64_t total_steps = cnt * d;
    /* four named barriers per pair: full[2], empty[2].  Id 0 doubles as __syncthreads, which is
     * only used above, before any pair starts. */
    const int bar_base = 4 * pair;

    if (is_gen) {
        uint32_t *endw = (stream + 1 == P.streams) ? P.end_window : nullptr;
        gw_run<CodeT>(ring, queue, B, total_steps, bar_base, P.mrow, P.mcol, P.cols, lane, endw);
        return;
    }

    /* ---- consumer warp */
    RolloutAcc acc;
    const uint64_t nbuf = (cnt + P.rb - 1) / P.rb;
    const uint32_t T = P.T;
    for (uint64_t k = 0; k < nbuf; ++k) {
        const uint32_t half = (uint32_t)(k & 1);
        nb_sync(bar_base + half);
        const CodeT *buf = queue + half * B;
        const uint64_t r0 = k * P.rb;
        const uint32_t live = (uint32_t)((cnt - r0 < P.rb) ? cnt - r0 : P.rb);
        for (uint32_t r = lane; r < live; r += 32) {
            const uint64_t i = first + r0 + r;
            const CodeT *my = buf + (size_t)r * d;
            uint32_t h = PCU_HASH_INIT;
            uint32_t pay = 0;
            uint8_t *tr = P.traj ? P.traj + (size_t)i * P.traj_stride * 3 : nullptr;
            if (KIND == KIND_MOLD) {
                for (uint32_t s = 0; s < d; ++s) {
                    const uint32_t code = my[s];
                    h = hash_step(h, code, 0);
                    if (tr) {
                        const uint32_t rr = code / P.cols;
                        tr[3 * s + 0] = (uint8_t)rr;
                        tr[3 * s + 1] = (uint8_t)(code - rr * P.cols);
                        tr[3 * s + 2] = 0;
                    }
                }
                acc.d += 1;
            } else {
                uint32_t p = (P.state_pos ? P.state_pos[i] : P.state_pos_u) - P.tab_base;
                int bias = P.bias ? P.bias[i] : P.bias_u;
                if (T == 1) {
                    const uint32_t adv = 2 + P.gen_words;
                    for (uint32_t s = 0; s < d; ++s) {
                        const uint32_t code = my[s];
                        p += code; /* discard(row*cols + col): random-tree.hh:88-89 */
                        bias += (int)((tab[p] >> 4) & 3u) - 1;
                        p += adv;
                        h = hash_step(h, code, 0);
                        if (tr) {
                            const uint32_t rr = code / P.cols;
                            tr[3 * s + 0] = (uint8_t)rr;
                            tr[3 * s + 1] = (uint8_t)(code - rr * P.cols);
                            tr[3 * s + 2] = 0;
                        }
                    }
                } else {
                    uint32_t gb = p - P.gen_words; /* where the current chance strategies were drawn */
                    for (uint32_t s = 0; s < d; ++s) {
                        const uint32_t code = my[s];
                        const uint32_t st = code * T;
                        const uint8_t *cell = tab + gb + 2 * st;
                        int S = 0;
                        for (uint32_t kk = 0; kk < T; ++kk) {
                            int nk = (int)(cell[2 * kk] & 15u) + 1;
                            if (nk * P.thr_q < P.thr_p * 10) nk = 0;
                            S += nk;
                        }
                        const bool fallback = (S == 0);
                        if (fallback) S = 1;
                        double pp = P.pu;
                        uint32_t ch = 0;
                        for (uint32_t kk = 0; kk < T; ++kk) {
                            int nk = (int)(cell[2 * kk] & 15u) + 1;
                            if (nk * P.thr_q < P.thr_p * 10) nk = 0;
                            if (fallback && kk == 0) nk = 1;
                            pp -= sfq[(uint32_t)nk * P.fq_w + (uint32_t)S];
                            if (pp <= 0) {
                                ch = kk;
                                break;
                            }
                        }
                        p += st + ch;
                        bias += (int)((tab[p] >> 4) & 3u) - 1;
                        p += 2;
                        gb = p;
                        p += P.gen_words;
                        h = hash_step(h, code, ch);
                        if (tr) {
                            const uint32_t rr = code / P.cols;
                            tr[3 * s + 0] = (uint8_t)rr;
                            tr[3 * s + 1] = (uint8_t)(code - rr * P.cols);
                            tr[3 * s + 2] = (uint8_t)ch;
                        }
                    }
                }
                pay = (uint32_t)((bias > 0) - (bias < 0) + 1);
                acc.w += (pay == 2);
                acc.d += (pay == 1);
                acc.l += (pay == 0);
            }
            if (tr)
                for (uint32_t s = d; s < P.traj_stride; ++s) tr[3 * s] = tr[3 * s + 1] = tr[3 * s + 2] = 0;
            if (P.payoff) P.payoff[i] = (uint8_t)pay;
            if (P.traj_hash) P.traj_hash[i] = h;
            acc.h += h;
            acc.n += 1;
        }
        __syncwarp();
        if (k + 2 < nbuf) nb_arrive(bar_base + 2 + half);
    }
    if (P.sums) {
        acc.w = warp_sum_u64(acc.w);
        acc.d = warp_sum_u64(acc.d);
        acc.l = warp_sum_u64(acc.l);
        acc.h = warp_sum_u64(acc.h);
        acc.n = warp_sum_u64(acc.n);
        if (lane == 0) {
            if (acc.w) atomicAdd(P.sums + 0, acc.w);
            if (acc.d) atomicAdd(P.sums + 1, acc.d);
            if (acc.l) atomicAdd(P.sums + 2, acc.l);
            atomicAdd(P.sums + 3, acc.n * d);
            atomicAdd(P.sums + 4, acc.h);
        }
    }
}

/* ------------------------------------------------------------------ k_gen_codes2 */
/* The same generator writing joint action codes to global memory, for batches the fused
 * kernel does not take (ragged depth, private state engines). */

struct Gen2Params {
    uint64_t total_steps;
    uint64_t chunk_steps; /* steps per stream */
    uint32_t streams;
    uint32_t cols;
    ModConst mrow, mcol;
    const uint32_t *checkpoints;
    void *codes; /* CodeT[total_steps] */
    uint32_t *end_window;
};

template <typename CodeT>
__global__ void __launch_bounds__(128) k_gen_codes2(const Gen2Params P) {
    __shared__ __align__(16) uint32_t rings[4 * GW_RING];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t stream = blockIdx.x * 4 + warp;
    if (stream >= P.streams) return;
    uint32_t *ring = rings + warp * GW_RING;
    const uint64_t first = (uint64_t)stream * P.chunk_steps;
    const uint64_t last = (first + P.chunk_steps < P.total_steps) ? first + P.chunk_steps : P.total_steps;
    gw_load_window(ring, P.checkpoints + (size_t)stream * MT_N, lane);
    __syncwarp();
    CodeT *codes = reinterpret_cast<CodeT *>(P.codes);
    const int l4 = 4 * lane;
    const bool active = lane < GW_ACTIVE;
    uint64_t q = first;
#define GEN2_GROUP(PH)                                                              \
    {                                                                               \
        if (active) {                                                               \
            const uint32_t c0 = gw_step_code<PH, 0>(ring, l4, P.mrow, P.mcol, P.cols); \
            const uint32_t c1 = gw_step_code<PH, 1>(ring, l4, P.mrow, P.mcol, P.cols); \
            if (q + lane < last) codes[q + lane] = (CodeT)c0;                       \
            if (q + GW_ACTIVE + lane < last) codes[q + GW_ACTIVE + lane] = (CodeT)c1; \
        }                                                                           \
        __syncwarp();                                                               \
        q += GW_STEPS;                                                              \
    }
    while (q < last) {
        GEN2_GROUP(0)
        if (q >= last) break;
        GEN2_GROUP(1)
        if (q >= last) break;
        GEN2_GROUP(2)
        if (q >= last) break;
        GEN2_GROUP(3)
    }
#undef GEN2_GROUP
    if (stream + 1 == P.streams && P.end_window) {
        const uint64_t words = 4 * (last - first);
        uint32_t w_end = (uint32_t)(words % GW_RING);
        if (w_end == 0 && words > 0) w_end = GW_RING;
        gw_dump_window(ring, w_end, P.end_window, lane);
    }
}
