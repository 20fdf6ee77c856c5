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
// Fused kernel (k_stream2): CTA = SP_STREAMS streams x 3 warps.  Two generator warps share a
// stream and alternate its 224-word groups (the recurrence of one group runs beside the
// tempering of the previous one); they put joint action codes (row * cols + col) into a
// shared-memory queue, and the stream's consumer warp plays the rollouts whose codes are
// complete, one to four per lane, against the shared state-stream table.  Hand-off is by named
// barriers (bar.sync): a waiting warp takes no issue slots.
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

/* The single-warp generator loop (k_gen_codes2), software pipelined: while the 8 words of group g are tempered and reduced
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
    uint32_t rb;      /* rollouts per consumer batch: 32, 64 or 128 */
    uint32_t bcodes;  /* rb * depth: codes of one batch */
    uint32_t ring_codes; /* code queue length: a whole number of batches, >= bcodes + 2 passes */
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

/* One model stream = THREE warps on one SM sub-partition (warp ids s, 4 + s, 8 + s of the CTA):
 *   generator halves A and B   share the stream's 896-word ring and alternate its 224-word groups: A makes groups
 *        0, 2, 4, ..., B groups 1, 3, 5, ...  A group is (1) the recurrence — operands from the ring, twist, store —
 *        and (2) temper + uniform_int reduction of its eight words per lane to two joint action codes, pure register
 *        arithmetic.  Group g + 1 needs group g's words (X[n-227]), so part (1) is a serial chain through the ring;
 *        part (2) is three quarters of the instructions and depends on nothing.  After storing its group a half meets
 *        its partner at a named barrier (64 threads) and then does part (2) WHILE the partner runs part (1) of the next
 *        group: the scheduler always has an independent instruction stream beside the serial one, which one
 *        generator warp per stream did not give it (r1 ncu: one warp issued every 2.4 cycles, ALU pipe 62 % busy),
 *        and a second STREAM per scheduler would cost a second jump-ahead checkpoint.
 *   consumer                    plays the rollouts whose codes are complete, once per pass.
 * Hand-off generator -> consumer: ONE named barrier per stream and pass (96 threads).  The consumer plays the
 * batches completed by pass j while the halves generate pass j + 1; the code queue holds a batch plus two passes. */
constexpr int SP_STREAMS = 4;               /* streams per CTA */
constexpr int SP_THREADS = SP_STREAMS * 96; /* 12 warps */

__device__ __forceinline__ void bar_sync(int id, int threads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}

/* Joint action codes of one group into the stream's code queue: a ring of `ring_codes` codes indexed by the absolute
 * step number modulo its length.  Lane l holds steps 56 g + l and 56 g + 28 + l. */
template <typename CodeT>
struct PairQueue {
    CodeT *queue;
    uint32_t ring_codes;
    uint32_t qpos; /* (56 g) mod ring_codes for this half's next group */
    __device__ __forceinline__ void put(uint32_t c0, uint32_t c1, int lane, bool active) {
        uint32_t i0 = qpos + lane, i1 = qpos + GW_ACTIVE + lane;
        i0 = min(i0, i0 - ring_codes); /* unsigned: wraps when i0 >= ring_codes */
        i1 = min(i1, i1 - ring_codes);
        if (active) {
            queue[i0] = (CodeT)c0;
            queue[i1] = (CodeT)c1;
        }
        qpos += 2 * GW_STEPS; /* this half's next group is two groups on */
        qpos = min(qpos, qpos - ring_codes);
    }
};

template <int HALF, int PH, typename CodeT>
__device__ __forceinline__ void pair_phase(uint32_t *ring, int l4, int lane, bool active, const FoldConst &mrow,
                                           const FoldConst &mcol, uint32_t cols, PairQueue<CodeT> &q, int pair_bar) {
    constexpr bool MINE = (PH & 1) == HALF;
    uint4 x0 = make_uint4(0, 0, 0, 0), x1 = x0;
    if (MINE) { /* part (1): the recurrence for this group */
        GwOps o;
        gw_load<PH>(ring, l4, o);
        x0 = gw_twist4(o.a0, o.a04, o.c0, o.c04);
        x1 = gw_twist4(o.a1, o.a14, o.c1, o.c14);
        gw_store<PH>(ring, l4, active, x0, x1);
    }
    bar_sync(pair_bar, 64); /* the group is in the ring: the partner may start the next one */
    if (MINE) {             /* part (2), beside the partner's part (1) */
        const uint32_t c0 = gw_code(x0, mrow, mcol, cols);
        const uint32_t c1 = gw_code(x1, mrow, mcol, cols);
        q.put(c0, c1, lane, active);
    }
}

template <int HALF, typename CodeT>
__device__ __forceinline__ void pair_generate(uint32_t *ring, CodeT *queue, uint32_t ring_codes, uint32_t groups, int lane,
                                              const FoldConst &mrow, const FoldConst &mcol, uint32_t cols, int pair_bar,
                                              int pass_bar) {
    const bool active = lane < GW_ACTIVE;
    const int l4 = 4 * (active ? lane : GW_ACTIVE - 1); /* lanes 28..31 shadow lane 27, stores predicated off */
    PairQueue<CodeT> q{queue, ring_codes, (uint32_t)(GW_STEPS * HALF)};
    for (uint32_t g = 0; g < groups; g += 4) {
        const uint32_t rest = groups - g; /* >= 1; the same in both halves */
        pair_phase<HALF, 0>(ring, l4, lane, active, mrow, mcol, cols, q, pair_bar);
        if (rest > 1) pair_phase<HALF, 1>(ring, l4, lane, active, mrow, mcol, cols, q, pair_bar);
        if (rest > 2) pair_phase<HALF, 2>(ring, l4, lane, active, mrow, mcol, cols, q, pair_bar);
        if (rest > 3) pair_phase<HALF, 3>(ring, l4, lane, active, mrow, mcol, cols, q, pair_bar);
        bar_sync(pass_bar, 96); /* the pass is in the queue (both halves) and the consumer is done with the pass
                                   before the previous one */
    }
}

/* Per-lane counters of a consumer warp: payoffs are 0, 1, 2, so the three outcome counts come from two sums. */
struct RolloutAcc {
    uint32_t pay_sum = 0, draws = 0, n = 0;
    unsigned long long h = 0;
};

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

/* One batch of rb = 32 Q rollouts whose codes are complete; `q` points at the batch's first code (a batch never wraps
 * in the queue: its length is a whole number of batches).  Lane l plays the Q CONSECUTIVE rollouts Q l .. Q l + Q - 1,
 * their table walks interleaved (Q independent chains of dependent shared-memory loads), and stores their outputs as
 * ONE vector: Q payoff bytes (Q = 4: a warp writes one contiguous 128-byte line, which is also one PCIe write when the
 * payoffs go straight to pinned host memory), or one byte of 2-bit payoffs (Q = 4, pcu_out.payoff_packed2), and Q
 * trajectory hashes.  Lanes past the end of a partial batch replay their last live rollout and store nothing for it.
 * A step of the walk: T == 1: chance strategies are all 1, chance action 0 (random-tree.hh:115-135 with p - 1 <= 0):
 * discard(row cols + col) (:88-89), one bias draw (:97), then 2 rows cols strategy draws; T > 1: the chance action
 * of the chosen cell comes from the per-position table. */
/* The state-stream table: in shared memory (32-bit shared addresses, LDS) when it fits, else read-only global
 * memory.  A generic pointer would cost a 64-bit add and a generic load per table access. */
template <bool SMEM>
struct TabAccess;
template <>
struct TabAccess<true> {
    using Addr = uint32_t;
    static __device__ __forceinline__ Addr at(const uint8_t *tab, uint32_t idx) {
        return (uint32_t)__cvta_generic_to_shared(tab) + idx;
    }
    static __device__ __forceinline__ int s8(Addr a) {
        int v;
        asm("ld.shared.s8 %0, [%1];" : "=r"(v) : "r"(a));
        return v;
    }
    static __device__ __forceinline__ uint32_t u8(Addr a) {
        uint32_t v;
        asm("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(a));
        return v;
    }
};
template <>
struct TabAccess<false> {
    using Addr = const uint8_t *;
    static __device__ __forceinline__ Addr at(const uint8_t *tab, uint32_t idx) { return tab + idx; }
    static __device__ __forceinline__ int s8(Addr a) { return (int)(int8_t)__ldg(a); }
    static __device__ __forceinline__ uint32_t u8(Addr a) { return __ldg(a); }
};

template <int KIND, typename CodeT, int Q, bool TAB_SMEM>
__device__ __forceinline__ void consume_batch(const Stream2Params &P, const CodeT *q, const uint8_t *tab, uint64_t i0,
                                              uint32_t live, int lane, RolloutAcc &acc) {
    using TA = TabAccess<TAB_SMEM>;
    const uint32_t d = P.depth, T = P.T;
    const uint32_t rl = (uint32_t)Q * lane;
    if (rl >= live) return;
    const uint32_t mine = (live - rl < (uint32_t)Q) ? live - rl : (uint32_t)Q; /* 1..Q rollouts of this lane */
    const uint64_t il = i0 + rl;
    const int8_t *tb = reinterpret_cast<const int8_t *>(tab);
    const uint8_t *tch = tab + P.tab_len;
    uint32_t h[Q], pay[Q];
    if (P.traj) { /* trajectories: one rollout at a time (tests and small batches) */
#pragma unroll
        for (int j = 0; j < Q; ++j) {
            h[j] = PCU_HASH_INIT;
            pay[j] = 0;
            if ((uint32_t)j >= mine) continue;
            const uint64_t i = il + j;
            const CodeT *qc = q + (rl + j) * d;
            uint8_t *tr = P.traj + (size_t)i * P.traj_stride * 3;
            if (KIND == KIND_MOLD) {
                for (uint32_t s = 0; s < d; ++s) {
                    const uint32_t code = qc[s];
                    h[j] = hash_step(h[j], code, 0);
                    put_traj<true>(tr, s, code, P.cols, 0, P.traj_stride);
                }
            } else {
                uint32_t g = (P.state_pos ? P.state_pos[i] : P.state_pos_u) - P.tab_base - P.gen_words;
                int bias = P.bias ? P.bias[i] : P.bias_u;
                for (uint32_t s = 0; s < d; ++s) {
                    const uint32_t code = qc[s];
                    const uint32_t st = code * T;
                    const uint32_t ch = T > 1 ? (uint32_t)tch[g + 2 * st] : 0u;
                    const uint32_t t = g + P.gen_words + st + ch; /* discard(get_transition_idx): :88-89,149-152 */
                    bias += tb[t];
                    g = t + 2;
                    h[j] = hash_step(h[j], code, ch);
                    put_traj<true>(tr, s, code, P.cols, ch, P.traj_stride);
                }
                pay[j] = (uint32_t)((bias > 0) - (bias < 0) + 1);
            }
            for (uint32_t s = d; s < P.traj_stride; ++s) tr[3 * s] = tr[3 * s + 1] = tr[3 * s + 2] = 0;
        }
    } else {
        const CodeT *qc[Q];
        typename TA::Addr p[Q]; /* table address of the next discard: one add per step instead of base + index */
        int bias[Q];
#pragma unroll
        for (int j = 0; j < Q; ++j) {
            const uint32_t jj = ((uint32_t)j < mine) ? (uint32_t)j : mine - 1;
            h[j] = PCU_HASH_INIT;
            pay[j] = 0;
            qc[j] = q + (rl + jj) * d;
            if (KIND == KIND_TREE) {
                p[j] = TA::at(tab, (P.state_pos ? P.state_pos[il + jj] : P.state_pos_u) - P.tab_base);
                bias[j] = P.bias ? P.bias[il + jj] : P.bias_u;
            }
        }
        if (KIND == KIND_MOLD) {
#pragma unroll 2
            for (uint32_t s = 0; s < d; ++s) {
#pragma unroll
                for (int j = 0; j < Q; ++j) h[j] = hash_step(h[j], qc[j][s], 0);
            }
        } else if (T == 1) {
            const uint32_t adv = 2 + P.gen_words;
#pragma unroll 2
            for (uint32_t s = 0; s < d; ++s) {
#pragma unroll
                for (int j = 0; j < Q; ++j) {
                    const uint32_t code = qc[j][s];
                    const typename TA::Addr t = p[j] + code;
                    bias[j] += TA::s8(t);
                    p[j] = t + adv;
                    h[j] = hash_step(h[j], code, 0);
                }
            }
        } else {
            const uint32_t gen = P.gen_words, tl = P.tab_len;
#pragma unroll
            for (int j = 0; j < Q; ++j) p[j] -= gen; /* where the current chance strategies were drawn */
#pragma unroll 2
            for (uint32_t s = 0; s < d; ++s) {
#pragma unroll
                for (int j = 0; j < Q; ++j) {
                    const uint32_t code = qc[j][s];
                    const uint32_t st = code * T;
                    const uint32_t ch = TA::u8(p[j] + (tl + 2 * st)); /* the chance table follows the bias table */
                    const typename TA::Addr t = p[j] + (gen + st + ch);
                    bias[j] += TA::s8(t);
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
            acc.pay_sum += pay[j];
            acc.draws += pay[j] & 1u;
            acc.h += h[j];
        }
    }
    acc.n += mine;
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

/* The consumer warp of one stream: meets the generator halves at the pass barrier and plays every batch the pass
 * completed.  All stream-local arithmetic is 32-bit (a stream is < 2^32 steps: host-checked). */
template <int KIND, typename CodeT, int Q, bool TAB_SMEM>
__device__ __forceinline__ void consume_stream(const Stream2Params &P, const CodeT *queue, const uint8_t *tab, int pass_bar,
                                               uint64_t first, uint32_t cnt, int lane, RolloutAcc &acc) {
    constexpr uint32_t RB = 32u * Q;
    const uint32_t d = P.depth;
    const uint32_t total_steps = cnt * d;
    const uint32_t npass = (total_steps + GW_PASS_STEPS - 1) / GW_PASS_STEPS;
    uint32_t r0 = 0;    /* first rollout of the next batch */
    uint32_t qoff = 0;  /* its first code in the queue: (r0 * d) mod ring_codes, always a batch boundary */
    uint32_t avail = 0; /* steps generated so far */
    for (uint32_t j = 0; j < npass; ++j) {
        bar_sync(pass_bar, 96);
        avail = (total_steps - avail > (uint32_t)GW_PASS_STEPS) ? avail + GW_PASS_STEPS : total_steps;
        while (r0 < cnt) { /* every batch whose codes pass j completed */
            const uint32_t live = (cnt - r0 < RB) ? cnt - r0 : RB;
            if ((r0 + live) * d > avail) break;
            consume_batch<KIND, CodeT, Q, TAB_SMEM>(P, queue + qoff, tab, first + r0, live, lane, acc);
            r0 += RB;
            qoff += P.bcodes;
            if (qoff >= P.ring_codes) qoff = 0;
        }
    }
}

template <int KIND, typename CodeT, bool TAB_SMEM>
__global__ void __launch_bounds__(SP_THREADS) k_stream2(const Stream2Params P) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t d = P.depth;
    /* layout: rings [SP_STREAMS][896] u32 | queues [SP_STREAMS][ring_codes + 16] CodeT | tables */
    uint32_t *rings = reinterpret_cast<uint32_t *>(smem_raw);
    CodeT *queues = reinterpret_cast<CodeT *>(rings + SP_STREAMS * GW_RING);
    const uint32_t R = P.ring_codes;
    const uint8_t *tab = P.tab;
    if (KIND == KIND_TREE && TAB_SMEM) {
        uint8_t *stab = reinterpret_cast<uint8_t *>(queues + (size_t)SP_STREAMS * (R + 16));
        stab = reinterpret_cast<uint8_t *>(((uintptr_t)stab + 15) & ~(uintptr_t)15);
        const uint32_t bytes = P.tab_len * (P.T > 1 ? 2u : 1u);
        for (uint32_t i = tid; i < bytes; i += blockDim.x) stab[i] = P.tab[i];
        tab = stab;
    }
    const int sl = warp % SP_STREAMS;   /* stream of this warp within the CTA: its three warps share a sub-partition */
    const int role = warp / SP_STREAMS; /* 0, 1: generator halves A, B; 2: consumer */
    const uint32_t stream = blockIdx.x * SP_STREAMS + sl;
    uint32_t *ring = rings + sl * GW_RING;
    CodeT *queue = queues + (size_t)sl * (R + 16);
    const int pair_bar = 1 + sl, pass_bar = 1 + SP_STREAMS + sl; /* barrier 0 is __syncthreads, used once below */
    if (stream < P.streams && role == 0) gw_load_window(ring, P.checkpoints + (size_t)stream * MT_N, lane);
    __syncthreads();
    if (stream >= P.streams) return;

    const uint64_t first = (uint64_t)stream * P.chunk;
    const uint32_t cnt = (uint32_t)((first + P.chunk <= P.n) ? P.chunk : P.n - first);
    const uint32_t total_steps = cnt * d;
    const uint32_t groups = (total_steps + GW_STEPS - 1) / GW_STEPS;

    if (role == 0) {
        pair_generate<0, CodeT>(ring, queue, R, groups, lane, P.mrow, P.mcol, P.cols, pair_bar, pass_bar);
        if (stream + 1 == P.streams && P.end_window) {
            /* words generated in the current pass before the stream's end: 4 * total_steps mod 896,
             * except that a stream ending exactly on a pass boundary sits at 896 */
            const uint64_t words = 4ull * total_steps;
            uint32_t w_end = (uint32_t)(words % GW_RING);
            if (w_end == 0 && words > 0) w_end = GW_RING;
            gw_dump_window(ring, w_end, P.end_window, lane);
        }
        return;
    }
    if (role == 1) {
        pair_generate<1, CodeT>(ring, queue, R, groups, lane, P.mrow, P.mcol, P.cols, pair_bar, pass_bar);
        return;
    }

    /* ---- consumer warp */
    RolloutAcc acc;
    if (P.rb == 128)
        consume_stream<KIND, CodeT, 4, TAB_SMEM>(P, queue, tab, pass_bar, first, cnt, lane, acc);
    else if (P.rb == 64)
        consume_stream<KIND, CodeT, 2, TAB_SMEM>(P, queue, tab, pass_bar, first, cnt, lane, acc);
    else
        consume_stream<KIND, CodeT, 1, TAB_SMEM>(P, queue, tab, pass_bar, first, cnt, lane, acc);
    if (P.sums) {
        unsigned long long pay_sum = warp_sum_u64(acc.pay_sum), draws = warp_sum_u64(acc.draws), n = warp_sum_u64(acc.n);
        const unsigned long long h = warp_sum_u64(acc.h);
        if (lane == 0) {
            if (KIND == KIND_MOLD) { /* MoldState never writes a payoff: every rollout counts in the draw slot */
                atomicAdd(P.sums + 1, n);
            } else {
                const unsigned long long wins = (pay_sum - draws) / 2; /* payoff_x2 2 = row win, 1 = draw, 0 = loss */
                if (wins) atomicAdd(P.sums + 0, wins);
                if (draws) atomicAdd(P.sums + 1, draws);
                if (n - wins - draws) atomicAdd(P.sums + 2, n - wins - draws);
            }
            atomicAdd(P.sums + 3, n * d);
            atomicAdd(P.sums + 4, h);
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
