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
__device__ __forceinline__ void put_traj(uint8_t *tr, uint32_t s, uint32_t code, uint32_t cols, uint32_t ch) {
    if (TRAJ) {
        const uint32_t rr = code / cols;
        tr[3 * s + 0] = (uint8_t)rr;
        tr[3 * s + 1] = (uint8_t)(code - rr * cols);
        tr[3 * s + 2] = (uint8_t)ch;
    }
}

template <bool TRAJ, typename CodeT>
__device__ __forceinline__ uint32_t walk_mold(const CodeT *queue, uint32_t qmask, uint32_t qb, uint32_t d, uint32_t cols,
                                              uint8_t *tr) {
    uint32_t h = PCU_HASH_INIT;
#pragma unroll 5
    for (uint32_t s = 0; s < d; ++s) {
        const uint32_t code = queue[(qb + s) & qmask];
        h = hash_step(h, code, 0);
        put_traj<TRAJ>(tr, s, code, cols, 0);
    }
    return h;
}

/* T == 1: chance strategies are all 1, chance action 0 (random-tree.hh:115-135 with p - 1 <= 0);
 * a step is discard(row * cols + col), one bias draw, then 2 * rows * cols strategy draws. */
template <bool TRAJ, typename CodeT>
__device__ __forceinline__ uint32_t walk_tree1(const CodeT *queue, uint32_t qmask, uint32_t qb, uint32_t d, uint32_t cols,
                                               const int8_t *tb, uint32_t p, uint32_t adv, int &bias, uint8_t *tr) {
    uint32_t h = PCU_HASH_INIT;
    int b = bias;
#pragma unroll 5
    for (uint32_t s = 0; s < d; ++s) {
        const uint32_t code = queue[(qb + s) & qmask];
        const uint32_t t = p + code; /* discard(row*cols + col): random-tree.hh:88-89 */
        b += tb[t];
        p = t + adv;
        h = hash_step(h, code, 0);
        put_traj<TRAJ>(tr, s, code, cols, 0);
    }
    bias = b;
    return h;
}

/* T > 1: the chance action of the chosen cell comes from the per-position table. */
template <bool TRAJ, typename CodeT>
__device__ __forceinline__ uint32_t walk_treeT(const CodeT *queue, uint32_t qmask, uint32_t qb, uint32_t d, uint32_t cols,
                                               uint32_t T, const int8_t *tb, const uint8_t *tch, uint32_t p,
                                               uint32_t gen_words, int &bias, uint8_t *tr) {
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
        put_traj<TRAJ>(tr, s, code, cols, ch);
    }
    bias = b;
    return h;
}

template <int KIND, typename CodeT, bool TAB_SMEM>
__global__ void __launch_bounds__(GW_PAIRS * 64) k_stream2(const Stream2Params P) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t d = P.depth;
    const uint32_t B = P.rb * d; /* codes per queue buffer */
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
    const uint64_t nbuf = (cnt + P.rb - 1) / P.rb;
    const uint32_t T = P.T;
    const uint64_t npass = (total_steps + GW_PASS_STEPS - 1) / GW_PASS_STEPS;
    uint64_t k = 0; /* next batch of rb rollouts */
    for (uint64_t j = 0; j < npass; ++j) {
      nb_sync(bar + (int)(j & 1));
      const uint64_t avail = ((j + 1) * GW_PASS_STEPS < total_steps) ? (j + 1) * GW_PASS_STEPS : total_steps;
      for (; k < nbuf; ++k) { /* every batch whose codes pass j completed */
        const uint64_t r0 = k * P.rb;
        const uint32_t live = (uint32_t)((cnt - r0 < P.rb) ? cnt - r0 : P.rb);
        if ((r0 + live) * d > avail) break;
        const uint32_t kbase = (uint32_t)((r0 * d) & qmask);
        for (uint32_t r = lane; r < live; r += 32) {
            const uint64_t i = first + r0 + r;
            const uint32_t qb = kbase + r * d;
            uint32_t h = PCU_HASH_INIT, pay = 0;
            uint8_t *tr = P.traj ? P.traj + (size_t)i * P.traj_stride * 3 : nullptr;
            if (KIND == KIND_MOLD) {
                if (tr)
                    h = walk_mold<true>(queue, qmask, qb, d, P.cols, tr);
                else
                    h = walk_mold<false>(queue, qmask, qb, d, P.cols, tr);
                acc.d += 1;
            } else {
                const uint32_t p = (P.state_pos ? P.state_pos[i] : P.state_pos_u) - P.tab_base;
                int bias = P.bias ? P.bias[i] : P.bias_u;
                const int8_t *tb = reinterpret_cast<const int8_t *>(tab);
                if (T == 1) {
                    if (tr)
                        h = walk_tree1<true>(queue, qmask, qb, d, P.cols, tb, p, 2 + P.gen_words, bias, tr);
                    else
                        h = walk_tree1<false>(queue, qmask, qb, d, P.cols, tb, p, 2 + P.gen_words, bias, tr);
                } else {
                    if (tr)
                        h = walk_treeT<true>(queue, qmask, qb, d, P.cols, T, tb, tab + P.tab_len, p, P.gen_words, bias, tr);
                    else
                        h = walk_treeT<false>(queue, qmask, qb, d, P.cols, T, tb, tab + P.tab_len, p, P.gen_words, bias, tr);
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
      }
      if (j + 2 < npass) {
          __syncwarp();
          nb_arrive(bar + 2 + (int)(j & 1));
      }
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

/* ------------------------------------------------------------------ k_stream3 */
/* One CTA = one model stream = FIVE warps, so that the serial part of the generator (the recurrence) is
 * no longer in the same in-order instruction stream as the embarrassingly parallel part (temper, reduce):
 *   G   (warp 0)  the recurrence only: loads, twist, ring store, and a copy of the raw words into a
 *                 double-buffered raw queue, one 896-word pass at a time
 *   T0  (warp 1)  tempers and reduces the FIRST step of every lane (x0) to action codes
 *   T1  (warp 2)  the same for the SECOND step of every lane (x1)
 *   W0, W1 (warps 3, 4)  play the rollouts a pass completed against the state-stream table, alternate batches
 * Eight mbarriers (two per edge, by parity of the pass): RAWFULL G -> T0,T1; RAWEMPTY T0,T1 -> G; CODES T0,T1 -> W0,W1;
 * QEMPTY W0,W1 -> T0,T1.  r1c ncu of k_stream2: one generator warp per scheduler issues at 2.4 cycles per instruction
 * (fixed-latency waits with nothing else to issue); the ALU pipe was 62 % busy. */
constexpr int S3_THREADS = 160;
constexpr int S3_RAWQ_U4 = 2 * 4 * 2 * 32; /* [parity][group][half][lane] uint4 */

/* Shared-memory mbarriers instead of named barriers: the 16 hardware barriers are an SM-wide pool, and a CTA that
 * addresses them by register reserves all 16 -> one CTA per SM (first k_stream3 build: 1.68 ms instead of 1.11).
 * One elected lane arrives (release) after __syncwarp; waiters spin on try_wait (acquire), which suspends in
 * hardware for a bounded time rather than burning issue slots. */
__device__ __forceinline__ void mb_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mb_arrive(uint64_t *bar, int lane) {
    __syncwarp();
    if (lane == 0)
        asm volatile("mbarrier.arrive.release.cta.shared::cta.b64 _, [%0];" ::"r"((uint32_t)__cvta_generic_to_shared(bar)) : "memory");
}
__device__ __forceinline__ void mb_wait(uint64_t *bar, uint32_t parity) {
    const uint32_t a = (uint32_t)__cvta_generic_to_shared(bar);
    uint32_t done;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.acquire.cta.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(a), "r"(parity)
            : "memory");
    } while (!done);
}

template <int PH>
__device__ __forceinline__ void s3_group(uint32_t *ring, int l4, bool active, uint4 *rq, int lane) {
    GwOps o;
    gw_load<PH>(ring, l4, o);
    const uint4 x0 = gw_twist4(o.a0, o.a04, o.c0, o.c04);
    const uint4 x1 = gw_twist4(o.a1, o.a14, o.c1, o.c14);
    gw_store<PH>(ring, l4, active, x0, x1);
    rq[(PH * 2 + 0) * 32 + lane] = x0;
    rq[(PH * 2 + 1) * 32 + lane] = x1;
    __syncwarp();
}

template <int KIND, typename CodeT, bool TAB_SMEM>
__global__ void __launch_bounds__(S3_THREADS) k_stream3(const Stream2Params P) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t d = P.depth;
    /* layout: mbarriers [8] u64 | ring [896] u32 | raw queue [512] uint4 | code queue [Q + 16] CodeT | tables */
    uint64_t *mbar = reinterpret_cast<uint64_t *>(smem_raw);
    uint32_t *ring = reinterpret_cast<uint32_t *>(smem_raw + 64);
    uint4 *rawq = reinterpret_cast<uint4 *>(ring + GW_RING);
    CodeT *queue = reinterpret_cast<CodeT *>(rawq + S3_RAWQ_U4);
    const uint32_t Q = P.qlen;
    const uint8_t *tab = P.tab;
    if (KIND == KIND_TREE && TAB_SMEM) {
        uint8_t *stab = reinterpret_cast<uint8_t *>(queue + (size_t)(Q + 16));
        stab = reinterpret_cast<uint8_t *>(((uintptr_t)stab + 15) & ~(uintptr_t)15);
        const uint32_t bytes = P.tab_len * (P.T > 1 ? 2u : 1u);
        for (uint32_t i = tid; i < bytes; i += blockDim.x) stab[i] = P.tab[i];
        tab = stab;
    }
    const uint32_t stream = blockIdx.x;
    if (warp == 0) gw_load_window(ring, P.checkpoints + (size_t)stream * MT_N, lane);
    enum { RAWFULL = 0, RAWEMPTY = 2, CODES = 4, QEMPTY = 6 };
    if (tid < 8) mb_init(mbar + tid, tid >= RAWEMPTY ? 2u : 1u); /* all but RAWFULL: two warps arrive */
    __syncthreads();

    const uint64_t first = (uint64_t)stream * P.chunk;
    const uint64_t cnt = (first + P.chunk <= P.n) ? P.chunk : P.n - first;
    const uint64_t total_steps = cnt * d;
    const uint32_t qmask = Q - 1;
    const uint64_t groups = (total_steps + GW_STEPS - 1) / GW_STEPS;
    const uint64_t npass = (groups + 3) / 4;
    const bool active = lane < GW_ACTIVE;

    if (warp == 0) { /* ---- G: the recurrence */
        const int l4 = 4 * (active ? lane : GW_ACTIVE - 1);
        for (uint64_t j = 0; j < npass; ++j) {
            const int par = (int)(j & 1);
            if (j >= 2) mb_wait(mbar + RAWEMPTY + par, (uint32_t)(((j >> 1) - 1) & 1));
            uint4 *rq = rawq + par * (S3_RAWQ_U4 / 2);
            const uint64_t g = 4 * j;
            s3_group<0>(ring, l4, active, rq, lane);
            if (g + 1 < groups) s3_group<1>(ring, l4, active, rq, lane);
            if (g + 2 < groups) s3_group<2>(ring, l4, active, rq, lane);
            if (g + 3 < groups) s3_group<3>(ring, l4, active, rq, lane);
            mb_arrive(mbar + RAWFULL + par, lane);
        }
        if (stream + 1 == P.streams && P.end_window) {
            const uint64_t words = 4 * total_steps;
            uint32_t w_end = (uint32_t)(words % GW_RING);
            if (w_end == 0 && words > 0) w_end = GW_RING;
            gw_dump_window(ring, w_end, P.end_window, lane);
        }
        return;
    }

    /* ---- T0 / T1 (warps 1, 2): temper + reduce one step per lane and group */
    if (warp < 3) {
        const int half = warp - 1;
        uint32_t qpos = 0;
        for (uint64_t j = 0; j < npass; ++j) {
            const int par = (int)(j & 1);
            mb_wait(mbar + RAWFULL + par, (uint32_t)((j >> 1) & 1));
            const uint4 *rq = rawq + par * (S3_RAWQ_U4 / 2);
            /* all four groups at once: 16 independent temper chains in flight.  Groups past the end of the stream
             * hold stale words; their codes land beyond the last step anyone reads (two passes of slack). */
            uint4 xs[4];
#pragma unroll
            for (int ph = 0; ph < 4; ++ph) xs[ph] = rq[(ph * 2 + half) * 32 + lane];
            uint32_t cs[4];
#pragma unroll
            for (int ph = 0; ph < 4; ++ph) cs[ph] = gw_code(xs[ph], P.mrow, P.mcol, P.cols);
            if (j + 2 < npass) mb_arrive(mbar + RAWEMPTY + par, lane);
            if (j >= 2) mb_wait(mbar + QEMPTY + par, (uint32_t)(((j >> 1) - 1) & 1));
#pragma unroll
            for (int ph = 0; ph < 4; ++ph) {
                const uint32_t idx = active ? ((qpos + ph * GW_STEPS + half * GW_ACTIVE + lane) & qmask) : qmask + 1;
                queue[idx] = (CodeT)cs[ph];
            }
            qpos = (qpos + GW_PASS_STEPS) & qmask;
            mb_arrive(mbar + CODES + par, lane);
        }
        return;
    }

    /* ---- W0 / W1 (warps 3, 4): play the rollouts each pass completes, alternate batches of rb (a walk is a chain
     * of dependent shared-memory loads: one walker could not keep up with the split generator) */
    const int wid = warp - 3;
    RolloutAcc acc;
    const uint64_t nbuf = (cnt + P.rb - 1) / P.rb;
    const uint32_t T = P.T;
    uint64_t k = 0; /* next batch of rb rollouts */
    for (uint64_t j = 0; j < npass; ++j) {
        const int par = (int)(j & 1);
        mb_wait(mbar + CODES + par, (uint32_t)((j >> 1) & 1));
        const uint64_t avail = ((j + 1) * GW_PASS_STEPS < total_steps) ? (j + 1) * GW_PASS_STEPS : total_steps;
        for (; k < nbuf; ++k) { /* every batch whose codes pass j completed */
            const uint64_t r0 = k * P.rb;
            const uint32_t live = (uint32_t)((cnt - r0 < P.rb) ? cnt - r0 : P.rb);
            if ((r0 + live) * d > avail) break;
            if ((int)(k & 1) != wid) continue;
            const uint32_t kbase = (uint32_t)((r0 * d) & qmask);
            for (uint32_t r = lane; r < live; r += 32) {
                const uint64_t i = first + r0 + r;
                const uint32_t qb = kbase + r * d;
                uint32_t h = PCU_HASH_INIT, pay = 0;
                uint8_t *tr = P.traj ? P.traj + (size_t)i * P.traj_stride * 3 : nullptr;
                if (KIND == KIND_MOLD) {
                    if (tr)
                        h = walk_mold<true>(queue, qmask, qb, d, P.cols, tr);
                    else
                        h = walk_mold<false>(queue, qmask, qb, d, P.cols, tr);
                    acc.d += 1;
                } else {
                    const uint32_t p = (P.state_pos ? P.state_pos[i] : P.state_pos_u) - P.tab_base;
                    int bias = P.bias ? P.bias[i] : P.bias_u;
                    const int8_t *tb = reinterpret_cast<const int8_t *>(tab);
                    if (T == 1) {
                        if (tr)
                            h = walk_tree1<true>(queue, qmask, qb, d, P.cols, tb, p, 2 + P.gen_words, bias, tr);
                        else
                            h = walk_tree1<false>(queue, qmask, qb, d, P.cols, tb, p, 2 + P.gen_words, bias, tr);
                    } else {
                        if (tr)
                            h = walk_treeT<true>(queue, qmask, qb, d, P.cols, T, tb, tab + P.tab_len, p, P.gen_words, bias, tr);
                        else
                            h = walk_treeT<false>(queue, qmask, qb, d, P.cols, T, tb, tab + P.tab_len, p, P.gen_words, bias, tr);
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
        }
        if (j + 2 < npass) mb_arrive(mbar + QEMPTY + par, lane);
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
