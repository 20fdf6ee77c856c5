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
// Generator g streams joint action codes (row * cols + col) into a shared-memory queue;
// consumer g (same SM sub-partition: warp id + GW_PAIRS) plays the rollouts whose codes are
// complete, one to four per lane, against the shared state-stream table.  Streams come in
// pairs around one jump-ahead checkpoint, one generated forwards and one BACKWARDS from it.
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

/* ---- the same stream generated BACKWARDS.
 * MT19937 is invertible: from X[n] ^ X[n-227] = (y >> 1) ^ (lsb(y) ? 0x9908b0df : 0), y = msb(X[n-624]) | low31(X[n-623]),
 * the top bit of the left side is lsb(y) (the constant has its top bit set, y >> 1 has not) and bit 30 is msb(y) (the
 * constant's bit 30 is clear).  So a window also yields the words BEFORE it, and ONE jump-ahead checkpoint serves TWO
 * model streams: one generated forwards from it, one backwards — half the polynomial applications per stream.
 * In reversed indexing Z[i] = X[P - 1 - i] (Z[0..623] is the window, newest word first) the inverse is again a forward
 * recurrence,  Z[j] = G(Z[j-624], Z[j-623], Z[j-397], Z[j-396])  for j >= 624:
 *     v1 = Z[j-624] ^ Z[j-397]     msb(Z[j])   = bit 30 of v1
 *     v2 = Z[j-623] ^ Z[j-396]     low31(Z[j]) = low31(((v2 ^ (msb(v2) ? 0x9908b0df : 0)) << 1) | msb(v2))
 * with the SAME ring layout as the forward generator (word n' of the current pass in slot (n' + 672) mod 896, the
 * window in slots 48..671) and a shortest lag of 396 words, so 224-word groups are independent of their predecessor.
 * v2 of word j is v1 of word j + 1: five XORs serve a lane's four consecutive words. */
template <int PH>
__device__ __forceinline__ void gw_load_b(const uint32_t *ring, int l4, GwOps &o, uint4 &d0, uint4 &d1) {
    constexpr int off0 = GW_WORDS * PH, off1 = off0 + GW_WORDS / 2;
    o.a0 = *reinterpret_cast<const uint4 *>(ring + gw_slot<off0 + 48>(l4)); /* Z[j-624 .. j-621] */
    o.a1 = *reinterpret_cast<const uint4 *>(ring + gw_slot<off1 + 48>(l4));
    o.a04 = ring[gw_slot<off0 + 52>(l4)];                                    /* Z[j-620] */
    o.a14 = ring[gw_slot<off1 + 52>(l4)];
    o.c0 = *reinterpret_cast<const uint4 *>(ring + gw_slot<off0 + 272>(l4)); /* Z[j-400 .. j-397]: .w is Z[j-397] */
    o.c1 = *reinterpret_cast<const uint4 *>(ring + gw_slot<off1 + 272>(l4));
    d0 = *reinterpret_cast<const uint4 *>(ring + gw_slot<off0 + 276>(l4));   /* Z[j-396 .. j-393] */
    d1 = *reinterpret_cast<const uint4 *>(ring + gw_slot<off1 + 276>(l4));
}

__device__ __forceinline__ uint32_t mt_untwist_dev(uint32_t v1, uint32_t v2) {
    const uint32_t sgn = (uint32_t)((int32_t)v2 >> 31);
    uint32_t w, x;
    asm("lop3.b32 %0, %1, %2, %3, 0x78;" : "=r"(w) : "r"(v2), "r"(sgn), "r"(0x9908b0dfu)); /* v2 ^ (sgn & matrix) */
    const uint32_t y = __funnelshift_l(v2, w, 1);                                            /* (w << 1) | msb(v2) */
    asm("lop3.b32 %0, %1, %2, %3, 0xE4;" : "=r"(x) : "r"(v1 << 1), "r"(y), "r"(0x80000000u)); /* msb from v1's bit 30 */
    return x;
}

__device__ __forceinline__ uint4 gw_untwist4(const uint4 &a, uint32_t a4, uint32_t c3, const uint4 &d) {
    /* V[k] = Z[j-624+k] ^ Z[j-397+k], k = 0..4 */
    const uint32_t v0 = a.x ^ c3, v1 = a.y ^ d.x, v2 = a.z ^ d.y, v3 = a.w ^ d.z, v4 = a4 ^ d.w;
    uint4 x;
    x.x = mt_untwist_dev(v0, v1);
    x.y = mt_untwist_dev(v1, v2);
    x.z = mt_untwist_dev(v2, v3);
    x.w = mt_untwist_dev(v3, v4);
    return x;
}

/* A backward lane holds its step's words newest-index first: (w3, w2, w1, w0). */
__device__ __forceinline__ uint32_t gw_code_b(const uint4 &x, const FoldConst &mrow, const FoldConst &mcol, uint32_t cols) {
    const uint32_t row = mod64_fold(mt_temper(x.w), mt_temper(x.z), mrow);
    const uint32_t col = mod64_fold(mt_temper(x.y), mt_temper(x.x), mcol);
    return row * cols + col;
}

/* Z[i] = window[623 - i] into slots 48..671.
 * The low 31 bits of a window's OLDEST word are not part of the generator's 19937-bit state (the forward recurrence
 * reads only its top bit), and the jump-ahead does not reproduce them: x^D = g(x) mod phi holds for the state, and the
 * polynomial's constant term drags the equally dead bits of the base window's oldest word into them.  Generating
 * backwards DOES read them, so they are rebuilt here from the window itself by the inverse recurrence:
 * low31(X[m]) comes from X[m+623] ^ X[m+396], i.e. low31(Z[623]) from Z[0] ^ Z[227]. */
__device__ __forceinline__ void gw_load_window_b(uint32_t *ring, const uint32_t *ckpt, int lane) {
    for (int i = lane; i < MT_N; i += 32) ring[48 + i] = ckpt[MT_N - 1 - i];
    __syncwarp();
    if (lane == 0) {
        const uint32_t old = ring[48 + MT_N - 1];
        ring[48 + MT_N - 1] = mt_untwist_dev(old >> 1, ring[48] ^ ring[48 + MT_LAG]); /* keeps old's top bit */
    }
    __syncwarp();
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
template <int DIR, typename Sink>
__device__ __forceinline__ void gw_generate(uint32_t *ring, uint64_t groups, int lane, const FoldConst &mrow,
                                            const FoldConst &mcol, uint32_t cols, Sink &sink) {
    const bool active = lane < GW_ACTIVE;
    const int l4 = 4 * (active ? lane : GW_ACTIVE - 1);
    if (groups == 0) return;
    GwOps o;
    uint4 d0, d1, x0, x1;
#define GW_LOAD(PH)                                     \
    {                                                   \
        if (DIR == 0)                                   \
            gw_load<(PH)>(ring, l4, o);                 \
        else                                            \
            gw_load_b<(PH)>(ring, l4, o, d0, d1);       \
    }
#define GW_TWIST()                                             \
    {                                                          \
        if (DIR == 0) {                                        \
            x0 = gw_twist4(o.a0, o.a04, o.c0, o.c04);          \
            x1 = gw_twist4(o.a1, o.a14, o.c1, o.c14);          \
        } else {                                               \
            x0 = gw_untwist4(o.a0, o.a04, o.c0.w, d0);         \
            x1 = gw_untwist4(o.a1, o.a14, o.c1.w, d1);         \
        }                                                      \
    }
    GW_LOAD(0)
    GW_TWIST()
    gw_store<0>(ring, l4, active, x0, x1);
    __syncwarp();

#define GW_BODY(PH, NEXT)                                                                                  \
    {                                                                                                      \
        if (NEXT) GW_LOAD(((PH) + 1) & 3)                                                                  \
        if constexpr (Sink::WORDS) {                                                                       \
            sink.put_words(g + (PH), x0, x1, active);                                                      \
        } else {                                                                                           \
            const uint32_t c0 = DIR == 0 ? gw_code(x0, mrow, mcol, cols) : gw_code_b(x0, mrow, mcol, cols); \
            const uint32_t c1 = DIR == 0 ? gw_code(x1, mrow, mcol, cols) : gw_code_b(x1, mrow, mcol, cols); \
            sink.put(g + (PH), c0, c1, active);                                                            \
        }                                                                                                  \
        if (NEXT) {                                                                                        \
            GW_TWIST()                                                                                     \
            gw_store<((PH) + 1) & 3>(ring, l4, active, x0, x1);                                            \
            __syncwarp();                                                                                  \
        }                                                                                                  \
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
#undef GW_LOAD
#undef GW_TWIST
}

/* ----------------------------------------------------------------- fused kernel */

struct Stream2Params {
    uint64_t n;       /* rollouts */
    uint64_t chunk;   /* rollouts per stream */
    uint32_t streams; /* ceil(n / chunk) */
    uint32_t ck_regular; /* windows at even stream positions: (streams + 1) / 2; one extra follows when streams is even */
    uint32_t depth;   /* uniform depth_bound / max_depth */
    uint32_t rb;      /* rollouts per consumer batch: 32, 64 or 128 */
    uint32_t bcodes;  /* rb * depth: codes of one batch */
    uint32_t ring_codes; /* code queue length: a whole number of batches, >= bcodes + 2 passes + 156 */
    uint32_t cols, T, gen_words;
    FoldConst mrow, mcol;
    const uint32_t *checkpoints; /* [ck_regular (+ 1)][624]: window k holds the raw words preceding model word
                                    2 k * (4 * depth * chunk); the extra one those preceding the last stream */
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

/* Fused kernel: CTA = GW_PAIRS generator warps + GW_PAIRS consumer warps; generator g and consumer g (warp id
 * g + GW_PAIRS: the same SM sub-partition) serve one model stream.  Streams come in PAIRS around one jump-ahead
 * checkpoint: stream 2c runs FORWARDS from the window at position 2c * stride, stream 2c - 1 BACKWARDS from the same
 * window (gw_load_b / gw_untwist4), so S streams cost S / 2 polynomial applications.  The last stream of a batch is
 * always a forward one (the only stream that may be shorter than a chunk, and the one that hands back the post-batch
 * engine window); when its index is odd it gets a checkpoint of its own. */
constexpr int GW_PAIRS = 4;                    /* generator/consumer warp pairs per CTA */
constexpr int GW_WINDOW_STEPS = MT_N / 4;      /* 156: the model steps a window itself holds (a backward stream's first) */

/* Named barriers between a generator warp and its consumer warp (64 threads each): a waiting
 * warp takes no issue slots, unlike a polling loop on the same scheduler. */
__device__ __forceinline__ void nb_sync(int id) { asm volatile("bar.sync %0, 64;" ::"r"(id) : "memory"); }
__device__ __forceinline__ void nb_arrive(int id) {
    __threadfence_block();
    asm volatile("bar.arrive %0, 64;" ::"r"(id) : "memory");
}

/* Code queue between a generator warp and its consumer: a ring of `ring_codes` codes (a whole number of batches, so a
 * batch never wraps) indexed by the step number in GENERATION order.  Hand-off is per pass (224 codes) over four named
 * barriers, full[2] = bar + {0,1} and empty[2] = bar + {2,3}, indexed by pass parity: the generator arrives on
 * full[j & 1] after pass j and, from pass 2 on, first waits on empty[j & 1], which the consumer arrives on once it has
 * played every rollout completed by pass j - 2.  Unplayed codes are then less than one batch plus two passes (plus the
 * 156 codes a backward stream takes from its window before pass 0). */
template <typename CodeT>
struct QueueSink {
    static constexpr bool WORDS = false;
    CodeT *queue;
    uint32_t ring_codes;
    int bar;
    int lane;
    uint32_t qpos; /* queue slot of the next group's first code */

    __device__ __forceinline__ void begin_pass(uint64_t pass) {
        if (pass >= 2) nb_sync(bar + 2 + (int)(pass & 1));
    }
    __device__ __forceinline__ void put(uint64_t, uint32_t c0, uint32_t c1, bool active) {
        uint32_t i0 = qpos + lane, i1 = qpos + GW_ACTIVE + lane;
        i0 = min(i0, i0 - ring_codes); /* unsigned: wraps when i0 >= ring_codes */
        i1 = min(i1, i1 - ring_codes);
        if (active) {
            queue[i0] = (CodeT)c0;
            queue[i1] = (CodeT)c1;
        }
        qpos += GW_STEPS;
        qpos = min(qpos, qpos - ring_codes);
    }
    __device__ __forceinline__ void end_pass(uint64_t pass) {
        __syncwarp();
        nb_arrive(bar + (int)(pass & 1));
    }
};

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

/* DIR = 1: a backward stream — the batch's rollouts come in DESCENDING index order from i0 (rollout r of the batch is
 * i0 - r) and a rollout's codes are stored last step first. */
template <int KIND, typename CodeT, int Q, bool TAB_SMEM, int DIR>
__device__ __forceinline__ void consume_batch(const Stream2Params &P, const CodeT *q, const uint8_t *tab, uint64_t i0,
                                              uint32_t live, int lane, RolloutAcc &acc) {
    using TA = TabAccess<TAB_SMEM>;
    const uint32_t d = P.depth, T = P.T;
    const uint32_t rl = (uint32_t)Q * lane;
    if (rl >= live) return;
    const uint32_t mine = (live - rl < (uint32_t)Q) ? live - rl : (uint32_t)Q; /* 1..Q rollouts of this lane */
    const int cstep = DIR ? -1 : 1;                         /* direction of a rollout's codes in the queue */
    const uint32_t cfirst = DIR ? d - 1 : 0;                /* queue offset of its first step */
    auto index_of = [&](uint32_t j) -> uint64_t { return DIR ? i0 - (rl + j) : i0 + rl + j; };
    const uint64_t il = DIR ? i0 - (rl + mine - 1) : i0 + rl; /* lowest rollout index of this lane */
    const int8_t *tb = reinterpret_cast<const int8_t *>(tab);
    const uint8_t *tch = tab + P.tab_len;
    uint32_t h[Q], pay[Q];
    if (P.traj) { /* trajectories: one rollout at a time (tests and small batches) */
#pragma unroll
        for (int j = 0; j < Q; ++j) {
            h[j] = PCU_HASH_INIT;
            pay[j] = 0;
            if ((uint32_t)j >= mine) continue;
            const uint64_t i = index_of(j);
            const CodeT *qc = q + (rl + j) * d + cfirst;
            uint8_t *tr = P.traj + (size_t)i * P.traj_stride * 3;
            if (KIND == KIND_MOLD) {
                for (uint32_t s = 0; s < d; ++s) {
                    const uint32_t code = qc[cstep * (int)s];
                    h[j] = hash_step(h[j], code, 0);
                    put_traj<true>(tr, s, code, P.cols, 0, P.traj_stride);
                }
            } else {
                uint32_t g = (P.state_pos ? P.state_pos[i] : P.state_pos_u) - P.tab_base - P.gen_words;
                int bias = P.bias ? P.bias[i] : P.bias_u;
                for (uint32_t s = 0; s < d; ++s) {
                    const uint32_t code = qc[cstep * (int)s];
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
            qc[j] = q + (rl + jj) * d + cfirst;
            if (KIND == KIND_TREE) {
                const uint64_t i = index_of(jj);
                p[j] = TA::at(tab, (P.state_pos ? P.state_pos[i] : P.state_pos_u) - P.tab_base);
                bias[j] = P.bias ? P.bias[i] : P.bias_u;
            }
        }
        if (KIND == KIND_MOLD) {
#pragma unroll 2
            for (uint32_t s = 0; s < d; ++s) {
#pragma unroll
                for (int j = 0; j < Q; ++j) h[j] = hash_step(h[j], qc[j][cstep * (int)s], 0);
            }
        } else if (T == 1) {
            const uint32_t adv = 2 + P.gen_words;
#pragma unroll 2
            for (uint32_t s = 0; s < d; ++s) {
#pragma unroll
                for (int j = 0; j < Q; ++j) {
                    const uint32_t code = qc[j][cstep * (int)s];
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
                    const uint32_t code = qc[j][cstep * (int)s];
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
            const int slot = DIR ? (int)mine - 1 - j : j; /* position of rollout j in the lane's ascending vector */
            bytes |= pay[j] << (8 * slot);
            packed |= pay[j] << (2 * slot);
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
                if ((uint32_t)j < mine) P.payoff[index_of(j)] = (uint8_t)pay[j];
        }
    }
    /* packed: a forward lane's il is a multiple of 4 (host-checked); a backward stream only has whole batches, so its
     * lanes hold four rollouts and il is a multiple of 4 as well */
    if (Q == 4 && P.payoff_packed2) P.payoff_packed2[il >> 2] = (uint8_t)packed;
    if (P.traj_hash) {
        if (Q == 4 && P.vec_ok && mine == 4) {
            *reinterpret_cast<uint4 *>(P.traj_hash + il) =
                DIR ? make_uint4(h[3], h[2], h[1], h[0]) : make_uint4(h[0], h[1], h[2], h[3]);
        } else {
#pragma unroll
            for (int j = 0; j < Q; ++j)
                if ((uint32_t)j < mine) P.traj_hash[index_of(j)] = h[j];
        }
    }
}

/* The consumer warp of one stream: waits for each pass of its generator and plays every batch the pass completed.
 * All stream-local arithmetic is 32-bit (a stream is < 2^32 steps: host-checked).  `pro`: codes already in the queue
 * before pass 0 (a backward stream's window steps). */
template <int KIND, typename CodeT, int Q, bool TAB_SMEM, int DIR>
__device__ __forceinline__ void consume_stream(const Stream2Params &P, const CodeT *queue, const uint8_t *tab, int bar,
                                               uint64_t first, uint32_t cnt, uint32_t pro, uint32_t npass, int lane,
                                               RolloutAcc &acc) {
    constexpr uint32_t RB = 32u * Q;
    const uint32_t d = P.depth;
    const uint32_t total_steps = cnt * d;
    uint32_t r0 = 0;      /* first rollout (in generation order) of the next batch */
    uint32_t qoff = 0;    /* its first code in the queue: always a batch boundary */
    uint32_t avail = pro; /* steps generated so far */
    for (uint32_t j = 0; j < npass; ++j) {
        nb_sync(bar + (int)(j & 1));
        avail = (total_steps - avail > (uint32_t)GW_PASS_STEPS) ? avail + GW_PASS_STEPS : total_steps;
        while (r0 < cnt) { /* every batch whose codes pass j completed */
            const uint32_t live = (cnt - r0 < RB) ? cnt - r0 : RB;
            if ((r0 + live) * d > avail) break;
            consume_batch<KIND, CodeT, Q, TAB_SMEM, DIR>(P, queue + qoff, tab, DIR ? first + (cnt - 1 - r0) : first + r0, live,
                                                         lane, acc);
            r0 += RB;
            qoff += P.bcodes;
            if (qoff >= P.ring_codes) qoff = 0;
        }
        if (j + 2 < npass) {
            __syncwarp();
            nb_arrive(bar + 2 + (int)(j & 1));
        }
    }
}

template <int KIND, typename CodeT, bool TAB_SMEM, int DIR>
__device__ __forceinline__ void consume_any(const Stream2Params &P, const CodeT *queue, const uint8_t *tab, int bar,
                                            uint64_t first, uint32_t cnt, uint32_t pro, uint32_t npass, int lane,
                                            RolloutAcc &acc) {
    if (P.rb == 128)
        consume_stream<KIND, CodeT, 4, TAB_SMEM, DIR>(P, queue, tab, bar, first, cnt, pro, npass, lane, acc);
    else if (P.rb == 64)
        consume_stream<KIND, CodeT, 2, TAB_SMEM, DIR>(P, queue, tab, bar, first, cnt, pro, npass, lane, acc);
    else
        consume_stream<KIND, CodeT, 1, TAB_SMEM, DIR>(P, queue, tab, bar, first, cnt, pro, npass, lane, acc);
}

template <int KIND, typename CodeT, bool TAB_SMEM>
__global__ void __launch_bounds__(GW_PAIRS * 64, 2) k_stream2(const Stream2Params P) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t d = P.depth;
    /* layout: rings [PAIRS][896] u32 | queues [PAIRS][ring_codes + 16] CodeT | tables */
    uint32_t *rings = reinterpret_cast<uint32_t *>(smem_raw);
    CodeT *queues = reinterpret_cast<CodeT *>(rings + GW_PAIRS * GW_RING);
    const uint32_t R = P.ring_codes;
    const uint8_t *tab = P.tab;
    if (KIND == KIND_TREE && TAB_SMEM) {
        uint8_t *stab = reinterpret_cast<uint8_t *>(queues + (size_t)GW_PAIRS * (R + 16));
        stab = reinterpret_cast<uint8_t *>(((uintptr_t)stab + 15) & ~(uintptr_t)15);
        const uint32_t bytes = P.tab_len * (P.T > 1 ? 2u : 1u);
        for (uint32_t i = tid; i < bytes; i += blockDim.x) stab[i] = P.tab[i];
        tab = stab;
    }
    const int pair = warp % GW_PAIRS;
    const bool is_gen = warp < GW_PAIRS;
    const uint32_t stream = blockIdx.x * GW_PAIRS + pair;
    uint32_t *ring = rings + pair * GW_RING;
    CodeT *queue = queues + (size_t)pair * (R + 16);
    /* four named barriers per pair; id 0 doubles as __syncthreads, used only below, before any pair starts */
    const int bar = 4 * pair;
    /* stream sigma: even -> forwards from window sigma / 2; odd -> backwards from window (sigma + 1) / 2, except the
     * batch's last stream, which runs forwards from the extra window behind the regular ones */
    const bool last = stream + 1 == P.streams;
    const bool backward = (stream & 1u) && !last;
    const uint32_t ck = (stream & 1u) ? (last ? P.ck_regular : (stream + 1) / 2) : stream / 2;
    if (stream < P.streams && is_gen) {
        if (backward)
            gw_load_window_b(ring, P.checkpoints + (size_t)ck * MT_N, lane);
        else
            gw_load_window(ring, P.checkpoints + (size_t)ck * MT_N, lane);
    }
    __syncthreads();
    if (stream >= P.streams) return;

    const uint64_t first = (uint64_t)stream * P.chunk;
    const uint32_t cnt = (uint32_t)((first + P.chunk <= P.n) ? P.chunk : P.n - first);
    const uint32_t total_steps = cnt * d;
    /* a backward stream's first 156 steps are the window itself */
    const uint32_t pro = backward ? (total_steps < (uint32_t)GW_WINDOW_STEPS ? total_steps : (uint32_t)GW_WINDOW_STEPS) : 0u;
    const uint32_t groups = (total_steps - pro + GW_STEPS - 1) / GW_STEPS;
    const uint32_t npass = groups ? (groups + 3) / 4 : 1u; /* a stream without a generated group still hands over once */

    if (is_gen) {
        QueueSink<CodeT> sink{queue, R, bar, lane, pro};
        if (backward) {
            for (uint32_t t = lane; t < pro; t += 32) /* codes of the window's own steps: temper + reduce only */
                queue[t] = (CodeT)gw_code_b(*reinterpret_cast<const uint4 *>(ring + 48 + 4 * t), P.mrow, P.mcol, P.cols);
            __syncwarp();
            gw_generate<1>(ring, groups, lane, P.mrow, P.mcol, P.cols, sink);
        } else {
            gw_generate<0>(ring, groups, lane, P.mrow, P.mcol, P.cols, sink);
        }
        if (groups == 0) sink.end_pass(0);
        if (last && P.end_window) {
            /* words generated in the current pass before the stream's end: 4 * total_steps mod 896,
             * except that a stream ending exactly on a pass boundary sits at 896 */
            const uint64_t words = 4ull * total_steps;
            uint32_t w_end = (uint32_t)(words % GW_RING);
            if (w_end == 0 && words > 0) w_end = GW_RING;
            gw_dump_window(ring, w_end, P.end_window, lane);
        }
        return;
    }

    /* ---- consumer warp */
    RolloutAcc acc;
    if (backward)
        consume_any<KIND, CodeT, TAB_SMEM, 1>(P, queue, tab, bar, first, cnt, pro, npass, lane, acc);
    else
        consume_any<KIND, CodeT, TAB_SMEM, 0>(P, queue, tab, bar, first, cnt, pro, npass, lane, acc);
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
    static constexpr bool WORDS = false;
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
    gw_generate<0>(ring, (last - first + GW_STEPS - 1) / GW_STEPS, lane, P.mrow, P.mcol, P.cols, sink);
    if (stream + 1 == P.streams && P.end_window) {
        const uint64_t words = 4 * (last - first);
        uint32_t w_end = (uint32_t)(words % GW_RING);
        if (w_end == 0 && words > 0) w_end = GW_RING;
        gw_dump_window(ring, w_end, P.end_window, lane);
    }
}

/* ------------------------------------------------------------------ k_gen_words */
/* The model stream itself — four tempered words per step — to global memory, for the variant kernel (thread.cuh),
 * whose action counts change from step to step and cannot be folded into the generator. */
struct WordSink {
    static constexpr bool WORDS = true;
    uint4 *words; /* first step of this stream */
    uint64_t count;
    int lane;
    __device__ __forceinline__ void begin_pass(uint64_t) {}
    __device__ __forceinline__ void put_words(uint64_t g, const uint4 &x0, const uint4 &x1, bool active) {
        const uint64_t q = g * GW_STEPS + lane;
        if (active && q < count) words[q] = make_uint4(mt_temper(x0.x), mt_temper(x0.y), mt_temper(x0.z), mt_temper(x0.w));
        if (active && q + GW_ACTIVE < count)
            words[q + GW_ACTIVE] = make_uint4(mt_temper(x1.x), mt_temper(x1.y), mt_temper(x1.z), mt_temper(x1.w));
    }
    __device__ __forceinline__ void put(uint64_t, uint32_t, uint32_t, bool) {}
    __device__ __forceinline__ void end_pass(uint64_t) {}
};

static __global__ void __launch_bounds__(128) k_gen_words(const Gen2Params P) {
    __shared__ __align__(16) uint32_t rings[4 * GW_RING];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t stream = blockIdx.x * 4 + warp;
    if (stream >= P.streams) return;
    uint32_t *ring = rings + warp * GW_RING;
    const uint64_t first = (uint64_t)stream * P.chunk_steps;
    const uint64_t last = (first + P.chunk_steps < P.total_steps) ? first + P.chunk_steps : P.total_steps;
    gw_load_window(ring, P.checkpoints + (size_t)stream * MT_N, lane);
    __syncwarp();
    WordSink sink{reinterpret_cast<uint4 *>(P.codes) + first, last - first, lane};
    gw_generate<0>(ring, (last - first + GW_STEPS - 1) / GW_STEPS, lane, P.mrow, P.mcol, P.cols, sink);
    if (stream + 1 == P.streams && P.end_window) {
        const uint64_t words = 4 * (last - first);
        uint32_t w_end = (uint32_t)(words % GW_RING);
        if (w_end == 0 && words > 0) w_end = GW_RING;
        gw_dump_window(ring, w_end, P.end_window, lane);
    }
}
