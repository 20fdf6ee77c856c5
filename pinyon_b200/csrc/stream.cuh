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
constexpr int GW_PASS_STEPS = 4 * GW_STEPS; /* model steps per pass (one turn of the ring) */
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

/* Operands of one group for one lane: both of its model steps (H = 0: step lane, H = 1: step
 * 28 + lane).  Stream word n' (counted from the start of the current 896-word pass) lives in
 * ring slot (n' + 672) mod 896, so X[n-624] is at n' + 48 and X[n-227] at n' + 445. */
struct GwOps {
    uint4 a0, c0, a1, c1; /* X[n-624 .. n-621], X[n-228 .. n-225] */
    uint32_t a04, c04, a14, c14; /* X[n-620], X[n-224] */
};

template <int PH>
__device__ __forceinline__ void gw_load(const uint32_t *ring, int l4, GwOps &o) {
    constexpr int off0 = GW_WORDS * PH, off1 = off0 + GW_WORDS / 2;
    o.a0 = *reinterpret_cast<const uint4 *>(ring + gw_slot<off0 + 48>(l4));
    o.a1 = *reinterpret_cast<const uint4 *>(ring + gw_slot<off1 + 48>(l4));
    o.c0 = *reinterpret_cast<const uint4 *>(ring + gw_slot<off0 + 444>(l4));
    o.c1 = *reinterpret_cast<const uint4 *>(ring + gw_slot<off1 + 444>(l4));
    o.a04 = ring[gw_slot<off0 + 52>(l4)];
    o.a14 = ring[gw_slot<off1 + 52>(l4)];
    o.c04 = ring[gw_slot<off0 + 448>(l4)];
    o.c14 = ring[gw_slot<off1 + 448>(l4)];
}

__device__ __forceinline__ uint4 gw_twist4(const uint4 &a, uint32_t a4, const uint4 &c, uint32_t c4) {
    uint4 x;
    x.x = mt_twist_dev(a.x, a.y, c.y);
    x.y = mt_twist_dev(a.y, a.z, c.z);
    x.z = mt_twist_dev(a.z, a.w, c.w);
    x.w = mt_twist_dev(a.w, a4, c4);
    return x;
}

template <int PH>
__device__ __forceinline__ void gw_store(uint32_t *ring, int l4, bool active, const uint4 &x0, const uint4 &x1) {
    constexpr int off0 = GW_WORDS * PH, off1 = off0 + GW_WORDS / 2;
    if (active) {
        *reinterpret_cast<uint4 *>(ring + gw_slot<off0 + 672>(l4)) = x0;
        *reinterpret_cast<uint4 *>(ring + gw_slot<off1 + 672>(l4)) = x1;
    }
}

/* Raw words of one step -> joint action code row * cols + col: temper, then the two
 * random_int draws (uniform_64 takes the HIGH word first, random.hh:26,28). */
__device__ __forceinline__ uint32_t gw_code(const uint4 &x, const FoldConst &mrow, const FoldConst &mcol, uint32_t cols) {
    const uint32_t row = mod64_fold(mt_temper(x.x), mt_temper(x.y), mrow);
    const uint32_t col = mod64_fold(mt_temper(x.z), mt_temper(x.w), mcol);
    return row * cols + col;
}

__device__ __forceinline__ void gw_load_window(uint32_t *ring, const uint32_t *ckpt, int lane) {
    for (int i = lane; i < MT_N; i += 32) ring[48 + i] = ckpt[i];
}

/* The 624 raw words preceding stream word `w_end` of the current 896-word pass. */
__device__ __forceinline__ void gw_dump_window(const uint32_t *ring, uint32_t w_end, uint32_t *dst, int lane) {
    for (int i = lane; i < MT_N; i += 32) dst[i] = ring[(w_end + 48 + i) % GW_RING];
}

/* Named barriers between a generator warp and its consumer warp (64 threads each): a waiting
 * warp takes no issue slots, unlike a polling loop on the same scheduler. */
__device__ __forceinline__ void nb_sync(int id) { asm volatile("bar.sync %0, 64;" ::"r"(id) : "memory"); }
__device__ __forceinline__ void nb_arrive(int id) {
    __threadfence_block();
    asm volatile("bar.arrive %0, 64;" ::"r"(id) : "memory");
}

/* The generator loop, software pipelined: while the 8 words of group g are tempered and reduced
 * to action codes (ALU work with no memory dependence), the operands of group g + 1 are already
 * in flight from the ring; its twist and store close the iteration.  A pass is four groups
 * (896 words: one turn of the ring).  `sink` receives the codes:
 *   sink.begin_pass(j)            before the first group of pass j (may wait for queue space)
 *   sink.put(g, c0, c1, active)   codes of steps 56 g + lane and 56 g + 28 + lane
 *   sink.end_pass(j)              after the last group of pass j (publishes progress)
 * Lanes 28..31 shadow lane 27 (same loads, stores predicated off): no divergence. */
template <typename Sink>
__device__ __forceinline__ void gw_generate(uint32_t *ring, uint64_t groups, int lane, const FoldConst &mrow,
                                            const FoldConst &mcol, uint32_t cols, Sink &sink) {
    const bool active = lane < GW_ACTIVE;
    const int l4 = 4 * (active ? lane : GW_ACTIVE - 1);
    if (groups == 0) return;
    GwOps o;
    uint4 x0, x1;
    gw_load<0>(ring, l4, o);
    x0 = gw_twist4(o.a0, o.a04, o.c0, o.c04);
    x1 = gw_twist4(o.a1, o.a14, o.c1, o.c14);
    gw_store<0>(ring, l4, active, x0, x1);
    __syncwarp();

#define GW_BODY(PH, NEXT)                                                  \
    {                                                                      \
        if (NEXT) gw_load<((PH) + 1) & 3>(ring, l4, o);                    \
        const uint32_t c0 = gw_code(x0, mrow, mcol, cols);                 \
        const uint32_t c1 = gw_code(x1, mrow, mcol, cols);                 \
        sink.put(g + (PH), c0, c1, active);                                \
        if (NEXT) {                                                        \
            x0 = gw_twist4(o.a0, o.a04, o.c0, o.c04);                      \
            x1 = gw_twist4(o.a1, o.a14, o.c1, o.c14);                      \
            gw_store<((PH) + 1) & 3>(ring, l4, active, x0, x1);            \
            __syncwarp();                                                  \
        }                                                                  \
    }

    uint64_t g = 0, pass = 0;
    for (; g + 4 < groups; g += 4, ++pass) { /* every group of this pass has a successor */
        sink.begin_pass(pass);
        GW_BODY(0, true)
        GW_BODY(1, true)
        GW_BODY(2, true)
        GW_BODY(3, true)
        sink.end_pass(pass);
    }
    const uint32_t rest = (uint32_t)(groups - g); /* 1..4 groups left; (x0, x1) is group g, phase 0 */
    sink.begin_pass(pass);
    GW_BODY(0, rest > 1)
    if (rest > 1) GW_BODY(1, rest > 2)
    if (rest > 2) GW_BODY(2, rest > 3)
    if (rest > 3) GW_BODY(3, false)
    sink.end_pass(pass);
#undef GW_BODY
}

/* ----------------------------------------------------------------- fused kernel */

struct Stream2Params {
    uint64_t n;       /* rollouts */
    uint64_t chunk;   /* rollouts per stream */
    uint32_t streams; /* ceil(n / chunk) */
    uint32_t depth;   /* uniform depth_bound / max_depth */
    uint32_t rb;      /* rollouts per consumer batch */
    uint32_t qlen;    /* codes per queue ring: power of two >= 2 * rb * depth + 448 */
    uint32_t cols, T, gen_words;
    FoldConst mrow, mcol;
    const uint32_t *checkpoints; /* [streams][624] raw words preceding each stream's first model word */
    const uint8_t *tab;          /* [tab_len] int8 bias steps, then (T > 1) [tab_len] chance actions; see
                                    host_build_table */
    uint32_t tab_base, tab_len;
    const uint32_t *state_pos;
    uint32_t state_pos_u;
    const int32_t *bias;
    int32_t bias_u;
    uint8_t *payoff;
    uint8_t *payoff_packed2; /* optional 2-bit payoffs, four rollouts per byte (rb = 128 only) */
    uint32_t vec_ok;         /* chunk % 4 == 0 and payoff / traj_hash suitably aligned: vector stores allowed */
    uint8_t *traj;
    uint32_t traj_stride;
    uint32_t *traj_hash;
    unsigned long long *sums;
    uint32_t *end_window; /* [624] raw words preceding the first model word after the batch */
};

struct RolloutAcc {
    unsigned long long w = 0, d = 0, l = 0, h = 0, n = 0;
};

/* Code queue between a generator warp and its consumer: a ring of Q (power of two) codes
 * indexed by absolute step number.  Hand-off is per pass (224 codes) over four named barriers,
 * full[2] = bar + {0,1} and empty[2] = bar + {2,3}, indexed by pass parity: the generator
 * arrives on full[j & 1] after pass j and, from pass 2 on, first waits on empty[j & 1], which the
 * consumer arrives on once it has played every rollout completed by pass j - 2.  Unplayed codes
 * are then less than one batch plus two passes: Q >= 2 * rb * depth + 448 holds them. */
template <typename CodeT>
struct QueueSink {
    CodeT *queue; /* Q codes, then a dummy slot for the shadow lanes */
    uint32_t qmask;
    int bar;
    int lane;
    uint32_t qpos = 0; /* (56 * g) & qmask for the next group */

    __device__ __forceinline__ void begin_pass(uint64_t pass) {
        if (pass >= 2) nb_sync(bar + 2 + (int)(pass & 1));
    }
    __device__ __forceinline__ void put(uint64_t, uint32_t c0, uint32_t c1, bool active) {
        const uint32_t i0 = active ? ((qpos + lane) & qmask) : qmask + 1;
        const uint32_t i1 = active ? ((qpos + GW_ACTIVE + lane) & qmask) : qmask + 1;
        queue[i0] = (CodeT)c0;
        queue[i1] = (CodeT)c1;
        qpos = (qpos + GW_STEPS) & qmask;
    }
    __device__ __forceinline__ void end_pass(uint64_t pass) {
        __syncwarp();
        nb_arrive(bar + (int)(pass & 1));
    }
};

/* Consumer side: one rollout = `d` codes of the queue walked over the shared state stream.
 * Codes of a rollout are contiguous in the ring (index & qmask).  All return the trajectory
 * hash; TRAJ also writes (row, col, chance) bytes. */
template <bool TRAJ>
__device__ __forceinline__ void put_traj(uint8_t *tr, uint32_t s, uint32_t code, uint32_t cols, uint32_t ch,
                                         uint32_t stride) {
    if (TRAJ && s < stride) { /* steps past traj_stride are played and hashed, not recorded (as the oracle) */
        const uint32_t rr = code / cols;
        tr[3 * s + 0] = (uint8_t)rr;
        tr[3 * s + 1] = (uint8_t)(code - rr * cols);
        tr[3 * s + 2] = (uint8_t)ch;
    }
}

template <bool TRAJ, typename CodeT>
__device__ __forceinline__ uint32_t walk_mold(const CodeT *queue, uint32_t qmask, uint32_t qb, uint32_t d, uint32_t cols,
                                              uint8_t *tr, uint32_t tstride) {
    uint32_t h = PCU_HASH_INIT;
#pragma unroll 5
    for (uint32_t s = 0; s < d; ++s) {
        const uint32_t code = queue[(qb + s) & qmask];
        h = hash_step(h, code, 0);
        put_traj<TRAJ>(tr, s, code, cols, 0, tstride);
    }
    return h;
}

/* T == 1: chance strategies are all 1, chance action 0 (random-tree.hh:115-135 with p - 1 <= 0);
 * a step is discard(row * cols + col), one bias draw, then 2 * rows * cols strategy draws. */
template <bool TRAJ, typename CodeT>
__device__ __forceinline__ uint32_t walk_tree1(const CodeT *queue, uint32_t qmask, uint32_t qb, uint32_t d, uint32_t cols,
                                               const int8_t *tb, uint32_t p, uint32_t adv, int &bias, uint8_t *tr,
                                               uint32_t tstride) {
    uint32_t h = PCU_HASH_INIT;
    int b = bias;
#pragma unroll 5
    for (uint32_t s = 0; s < d; ++s) {
        const uint32_t code = queue[(qb + s) & qmask];
        const uint32_t t = p + code; /* discard(row*cols + col): random-tree.hh:88-89 */
        b += tb[t];
        p = t + adv;
        h = hash_step(h, code, 0);
        put_traj<TRAJ>(tr, s, code, cols, 0, tstride);
    }
    bias = b;
    return h;
}

/* T > 1: the chance action of the chosen cell comes from the per-position table. */
template <bool TRAJ, typename CodeT>
__device__ __forceinline__ uint32_t walk_treeT(const CodeT *queue, uint32_t qmask, uint32_t qb, uint32_t d, uint32_t cols,
                                               uint32_t T, const int8_t *tb, const uint8_t *tch, uint32_t p,
                                               uint32_t gen_words, int &bias, uint8_t *tr, uint32_t tstride) {
    uint32_t h = PCU_HASH_INIT;
    int b = bias;
    uint32_t gb = p - gen_words; /* where the current chance strategies were drawn */
#pragma unroll 5
    for (uint32_t s = 0; s < d; ++s) {
        const uint32_t code = queue[(qb + s) & qmask];
        const uint32_t st = code * T;
        const uint32_t ch = tch[gb + 2 * st];
        const uint32_t t = gb + gen_words + st + ch; /* discard(get_transition_idx): random-tree.hh:88-89,149-152 */
        b += tb[t];
        gb = t + 2;
        h = hash_step(h, code, ch);
        put_traj<TRAJ>(tr, s, code, cols, ch, tstride);
    }
    bias = b;
    return h;
}

/* One batch of rb = 32 Q rollouts whose codes are complete.  Lane l plays the Q CONSECUTIVE rollouts
 * Q l .. Q l + Q - 1 of the batch, their table walks interleaved (Q independent chains of dependent
 * shared-memory loads), and stores their outputs as ONE vector: Q payoff bytes (Q = 4: a warp writes one
 * contiguous 128-byte line, which is also one PCIe write when the payoffs go straight to pinned host memory),
 * or one byte of 2-bit payoffs (Q = 4, pcu_out.payoff_packed2), and Q trajectory hashes.  Lanes past the end of
 * a partial batch replay their last live rollout and store nothing for it. */
template <int KIND, typename CodeT, int Q>
__device__ __forceinline__ void consume_batch(const Stream2Params &P, const CodeT *queue, uint32_t qmask, const uint8_t *tab,
                                              uint64_t i0, uint32_t kbase, uint32_t live, int lane, RolloutAcc &acc) {
    const uint32_t d = P.depth, T = P.T;
    const uint32_t rl = (uint32_t)Q * lane;
    if (rl >= live) return;
    const uint32_t mine = (live - rl < (uint32_t)Q) ? live - rl : (uint32_t)Q; /* 1..Q rollouts of this lane */
    const uint64_t il = i0 + rl;
    uint32_t h[Q], pay[Q];
    if (P.traj) { /* trajectories: one rollout at a time (tests and small batches) */
#pragma unroll
        for (int j = 0; j < Q; ++j) {
            h[j] = PCU_HASH_INIT;
            pay[j] = 0;
            if ((uint32_t)j >= mine) continue;
            const uint64_t i = il + j;
            const uint32_t qb = kbase + (rl + j) * d;
            uint8_t *tr = P.traj + (size_t)i * P.traj_stride * 3;
            if (KIND == KIND_MOLD) {
                h[j] = walk_mold<true>(queue, qmask, qb, d, P.cols, tr, P.traj_stride);
            } else {
                const uint32_t p = (P.state_pos ? P.state_pos[i] : P.state_pos_u) - P.tab_base;
                int bias = P.bias ? P.bias[i] : P.bias_u;
                const int8_t *tb = reinterpret_cast<const int8_t *>(tab);
                if (T == 1)
                    h[j] = walk_tree1<true>(queue, qmask, qb, d, P.cols, tb, p, 2 + P.gen_words, bias, tr, P.traj_stride);
                else
                    h[j] = walk_treeT<true>(queue, qmask, qb, d, P.cols, T, tb, tab + P.tab_len, p, P.gen_words, bias, tr,
                                            P.traj_stride);
                pay[j] = (uint32_t)((bias > 0) - (bias < 0) + 1);
            }
            for (uint32_t s = d; s < P.traj_stride; ++s) tr[3 * s] = tr[3 * s + 1] = tr[3 * s + 2] = 0;
        }
    } else {
        uint32_t qb[Q], p[Q];
        int bias[Q];
#pragma unroll
        for (int j = 0; j < Q; ++j) {
            const uint32_t jj = ((uint32_t)j < mine) ? (uint32_t)j : mine - 1;
            h[j] = PCU_HASH_INIT;
            pay[j] = 0;
            qb[j] = kbase + (rl + jj) * d;
            if (KIND == KIND_TREE) {
                p[j] = (P.state_pos ? P.state_pos[il + jj] : P.state_pos_u) - P.tab_base;
                bias[j] = P.bias ? P.bias[il + jj] : P.bias_u;
            }
        }
        if (KIND == KIND_MOLD) {
#pragma unroll 2
            for (uint32_t s = 0; s < d; ++s) {
#pragma unroll
                for (int j = 0; j < Q; ++j) h[j] = hash_step(h[j], queue[(qb[j] + s) & qmask], 0);
            }
        } else if (T == 1) {
            /* chance strategies are all 1, chance action 0 (random-tree.hh:115-135 with p - 1 <= 0): a step is
             * discard(row * cols + col), one bias draw, then 2 * rows * cols strategy draws */
            const int8_t *tb = reinterpret_cast<const int8_t *>(tab);
            const uint32_t adv = 2 + P.gen_words;
#pragma unroll 2
            for (uint32_t s = 0; s < d; ++s) {
#pragma unroll
                for (int j = 0; j < Q; ++j) {
                    const uint32_t code = queue[(qb[j] + s) & qmask];
                    const uint32_t t = p[j] + code; /* discard(row*cols + col): random-tree.hh:88-89 */
                    bias[j] += tb[t];
                    p[j] = t + adv;
                    h[j] = hash_step(h[j], code, 0);
                }
            }
        } else {
            const int8_t *tb = reinterpret_cast<const int8_t *>(tab);
            const uint8_t *tch = tab + P.tab_len;
            const uint32_t gen = P.gen_words;
#pragma unroll
            for (int j = 0; j < Q; ++j) p[j] -= gen; /* where the current chance strategies were drawn */
#pragma unroll 2
            for (uint32_t s = 0; s < d; ++s) {
#pragma unroll
                for (int j = 0; j < Q; ++j) {
                    const uint32_t code = queue[(qb[j] + s) & qmask];
                    const uint32_t st = code * T;
                    const uint32_t ch = tch[p[j] + 2 * st];
                    const uint32_t t = p[j] + gen + st + ch; /* discard(get_transition_idx): random-tree.hh:88-89,149-152 */
                    bias[j] += tb[t];
                    p[j] = t + 2;
                    h[j] = hash_step(h[j], code, ch);
                }
            }
        }
        if (KIND == KIND_TREE) {
#pragma unroll
            for (int j = 0; j < Q; ++j) pay[j] = (uint32_t)((bias[j] > 0) - (bias[j] < 0) + 1);
        }
    }
    /* ---- outputs and counters */
    uint32_t packed = 0, bytes = 0;
#pragma unroll
    for (int j = 0; j < Q; ++j) {
        if ((uint32_t)j < mine) {
            bytes |= pay[j] << (8 * j);
            packed |= pay[j] << (2 * j);
            if (KIND == KIND_MOLD) {
                acc.d += 1;
            } else {
                acc.w += (pay[j] == 2);
                acc.d += (pay[j] == 1);
                acc.l += (pay[j] == 0);
            }
            acc.h += h[j];
            acc.n += 1;
        }
    }
    if (P.payoff) {
        if (Q == 4 && P.vec_ok && mine == 4) {
            *reinterpret_cast<uint32_t *>(P.payoff + il) = bytes;
        } else if (Q == 2 && P.vec_ok && mine == 2) {
            *reinterpret_cast<uint16_t *>(P.payoff + il) = (uint16_t)bytes;
        } else {
#pragma unroll
            for (int j = 0; j < Q; ++j)
                if ((uint32_t)j < mine) P.payoff[il + j] = (uint8_t)pay[j];
        }
    }
    if (Q == 4 && P.payoff_packed2) P.payoff_packed2[il >> 2] = (uint8_t)packed; /* il is a multiple of 4 (host-checked) */
    if (P.traj_hash) {
        if (Q == 4 && P.vec_ok && mine == 4) {
            *reinterpret_cast<uint4 *>(P.traj_hash + il) = make_uint4(h[0], h[1], h[2], h[3]);
        } else {
#pragma unroll
            for (int j = 0; j < Q; ++j)
                if ((uint32_t)j < mine) P.traj_hash[il + j] = h[j];
        }
    }
}

/* The consumer warp of one stream: waits for each pass of its generator and plays every batch the pass completed. */
template <int KIND, typename CodeT, int Q>
__device__ __forceinline__ void consume_stream(const Stream2Params &P, const CodeT *queue, uint32_t qmask, const uint8_t *tab,
                                               int bar, uint64_t first, uint64_t cnt, int lane, RolloutAcc &acc) {
    constexpr uint32_t RB = 32u * Q;
    const uint32_t d = P.depth;
    const uint64_t total_steps = cnt * d;
    const uint64_t nbuf = (cnt + RB - 1) / RB;
    const uint64_t npass = (total_steps + GW_PASS_STEPS - 1) / GW_PASS_STEPS;
    uint64_t k = 0; /* next batch of RB rollouts */
    for (uint64_t j = 0; j < npass; ++j) {
        nb_sync(bar + (int)(j & 1));
        const uint64_t avail = ((j + 1) * GW_PASS_STEPS < total_steps) ? (j + 1) * GW_PASS_STEPS : total_steps;
        for (; k < nbuf; ++k) { /* every batch whose codes pass j completed */
            const uint64_t r0 = k * RB;
            const uint32_t live = (uint32_t)((cnt - r0 < RB) ? cnt - r0 : RB);
            if ((r0 + live) * d > avail) break;
            consume_batch<KIND, CodeT, Q>(P, queue, qmask, tab, first + r0, (uint32_t)((r0 * d) & qmask), live, lane, acc);
        }
        if (j + 2 < npass) {
            __syncwarp();
            nb_arrive(bar + 2 + (int)(j & 1));
        }
    }
}

template <int KIND, typename CodeT, bool TAB_SMEM>
__global__ void __launch_bounds__(GW_PAIRS * 64) k_stream2(const Stream2Params P) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t d = P.depth;
    /* layout: rings [PAIRS][896] u32 | queues [PAIRS][Q + 16] CodeT | tables */
    uint32_t *rings = reinterpret_cast<uint32_t *>(smem_raw);
    CodeT *queues = reinterpret_cast<CodeT *>(rings + GW_PAIRS * GW_RING);
    const uint32_t Q = P.qlen;
    const uint8_t *tab = P.tab;
    if (KIND == KIND_TREE && TAB_SMEM) {
        uint8_t *stab = reinterpret_cast<uint8_t *>(queues + (size_t)GW_PAIRS * (Q + 16));
        stab = reinterpret_cast<uint8_t *>(((uintptr_t)stab + 15) & ~(uintptr_t)15);
        const uint32_t bytes = P.tab_len * (P.T > 1 ? 2u : 1u);
        for (uint32_t i = tid; i < bytes; i += blockDim.x) stab[i] = P.tab[i];
        tab = stab;
    }
    const int pair = warp % GW_PAIRS;
    const bool is_gen = warp < GW_PAIRS;
    const uint32_t stream = blockIdx.x * GW_PAIRS + pair;
    uint32_t *ring = rings + pair * GW_RING;
    CodeT *queue = queues + (size_t)pair * (Q + 16);
    /* four named barriers per pair; id 0 doubles as __syncthreads, used only below, before any
     * pair starts */
    const int bar = 4 * pair;
    if (stream < P.streams && is_gen) gw_load_window(ring, P.checkpoints + (size_t)stream * MT_N, lane);
    __syncthreads();
    if (stream >= P.streams) return;

    const uint64_t first = (uint64_t)stream * P.chunk;
    const uint64_t cnt = (first + P.chunk <= P.n) ? P.chunk : P.n - first;
    const uint64_t total_steps = cnt * d;
    const uint32_t qmask = Q - 1;

    if (is_gen) {
        QueueSink<CodeT> sink{queue, qmask, bar, lane};
        gw_generate(ring, (total_steps + GW_STEPS - 1) / GW_STEPS, lane, P.mrow, P.mcol, P.cols, sink);
        if (stream + 1 == P.streams && P.end_window) {
            /* words generated in the current pass before the stream's end: 4 * total_steps mod 896,
             * except that a stream ending exactly on a pass boundary sits at 896 */
            const uint64_t words = 4 * total_steps;
            uint32_t w_end = (uint32_t)(words % GW_RING);
            if (w_end == 0 && words > 0) w_end = GW_RING;
            gw_dump_window(ring, w_end, P.end_window, lane);
        }
        return;
    }

    /* ---- consumer warp */
    RolloutAcc acc;
    if (P.rb == 128)
        consume_stream<KIND, CodeT, 4>(P, queue, qmask, tab, bar, first, cnt, lane, acc);
    else if (P.rb == 64)
        consume_stream<KIND, CodeT, 2>(P, queue, qmask, tab, bar, first, cnt, lane, acc);
    else
        consume_stream<KIND, CodeT, 1>(P, queue, qmask, tab, bar, first, cnt, lane, acc);
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
    FoldConst mrow, mcol;
    const uint32_t *checkpoints;
    void *codes; /* CodeT[total_steps] */
    uint32_t *end_window;
};

template <typename CodeT>
struct GlobalSink {
    CodeT *codes; /* first code of this stream */
    uint64_t count;
    int lane;
    __device__ __forceinline__ void begin_pass(uint64_t) {}
    __device__ __forceinline__ void put(uint64_t g, uint32_t c0, uint32_t c1, bool active) {
        const uint64_t q = g * GW_STEPS + lane;
        if (active && q < count) codes[q] = (CodeT)c0;
        if (active && q + GW_ACTIVE < count) codes[q + GW_ACTIVE] = (CodeT)c1;
    }
    __device__ __forceinline__ void end_pass(uint64_t) {}
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
    GlobalSink<CodeT> sink{reinterpret_cast<CodeT *>(P.codes) + first, last - first, lane};
    gw_generate(ring, (last - first + GW_STEPS - 1) / GW_STEPS, lane, P.mrow, P.mcol, P.cols, sink);
    if (stream + 1 == P.streams && P.end_window) {
        const uint64_t words = 4 * (last - first);
        uint32_t w_end = (uint32_t)(words % GW_RING);
        if (w_end == 0 && words > 0) w_end = GW_RING;
        gw_dump_window(ring, w_end, P.end_window, lane);
    }
}
