// k_lane — distinct roots, ONE ROLLOUT PER LANE.
//
// Every rollout owns a private state prng (random-tree.hh:14, seeded by prng(seed) ->
// std::mt19937::seed, random.hh:16) and, in mode P, a private model prng.  The 623-step
// seeding chain x[k] = 1812433253 (x[k-1] ^ x[k-1] >> 30) + k is strictly serial, so the
// only parallelism is ACROSS rollouts: a warp runs 32 chains per instruction.  A full
// engine (624 words) per lane does not fit in shared memory at any useful occupancy, so
// the kernel never stores the seeded array at all:
//
//   phase 1  chain L runs alone from x[0] to x[397]  (mode P: the model's too) — its own kernel,
//            k_lane_pre, at full occupancy, because a lone chain is latency-bound
//   phase 2  chains E (at k) and L (at k+397) run in lock-step; the first refill's word
//            X[k] = x[k+397] ^ twist(x[k], x[k+1]) (bits/random.tcc _M_gen_rand) falls out
//            of registers and is the only thing stored, [word][lane] (conflict-free)
//   phase 3  for k >= 227, X[k] = X[k-227] ^ twist(x[k], x[k+1]) from the lane's own column
//   phase 4  the lane plays its rollout: the draw positions it reads (random-tree.hh:89,97)
//            are a function of the action codes alone when transitions == 1
//
// Only X[0 .. words) is produced, words = 2 + the largest draw index any rollout of the
// batch can read (host-checked <= 623: everything stays inside the first refill).
// Control flow is warp-uniform throughout; lanes differ only in data and smem columns.
#pragma once

struct LaneParams {
    WarpParams w;
    uint32_t words_pad; /* rows X[0 .. words_pad) are produced per lane (multiple of 8, <= 624); 0 for MoldState */
    uint32_t keep;      /* rows [0, keep) have their own shared-memory row (multiple of 8): X[k-227] operands */
    uint32_t ring;      /* rows >= keep live in a ring of this many rows (multiple of 8; 0: none, keep = words_pad) */
    uint32_t ring_m;    /* floor(2^32 / ring) */
    uint32_t hi0;       /* highest draw index step 0 can read, over the batch */
    uint32_t hstep;     /* hi of step s+1 - hi of step s = gen + rows cols - 1 + 2 */
    uint32_t dmax;      /* largest depth in the batch */
    uint32_t mulc;      /* 1812433253, passed as data so that it lives in a register: IMAD then takes the
                         * chain index as its immediate addend and a chain step is SHF + LOP3 + IMAD */
    uint32_t one;       /* 1, as data: index increments written k * one + c issue on the FMA pipe */
    uint32_t *l397;     /* [n] x[397] of every state engine, written by k_lane_pre */
    uint32_t *lm397;    /* [n] the same for the model engines (mode P) */
};

/* x[k] = mulc (x[k-1] ^ x[k-1] >> 30) + k — std::mt19937::seed (bits/random.tcc), reached from
 * prng(seed) (random.hh:16). */
__device__ __forceinline__ uint32_t lane_seed_step(uint32_t v, uint32_t mulc, uint32_t k) {
    uint32_t r;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(v ^ (v >> 30)), "r"(mulc), "r"(k));
    return r;
}
/* k + c on the FMA pipe (ptxas cannot fold a multiplier it does not know) */
__device__ __forceinline__ uint32_t lane_fma_add(uint32_t k, uint32_t one, uint32_t c) {
    uint32_t r;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(k), "r"(one), "r"(c));
    return r;
}

/* One draw of the lane's model engine in mode P from the lock-step chains: em = x[k], lm = x[k+397],
 * valid while k < 227; k1 = k + 1 and k398 = k + 398 are carried in registers. */
__device__ __forceinline__ uint32_t lane_model_word(uint32_t &em, uint32_t &lm, uint32_t &k1, uint32_t &k398,
                                                    uint32_t mulc, uint32_t one) {
    const uint32_t en = lane_seed_step(em, mulc, k1);
    const uint32_t x = mt_twist_dev(em, en, lm);
    em = en;
    lm = lane_seed_step(lm, mulc, k398);
    k1 = lane_fma_add(k1, one, 1);
    k398 = lane_fma_add(k398, one, 1);
    return mt_temper(x);
}

/* Phase 1 as its own kernel: the late chains run alone from x[0] to x[397].  No shared memory, full
 * occupancy, one rollout per thread; fully unrolled so every chain index is an immediate. */
template <int KIND, int MODE>
__global__ void __launch_bounds__(256) k_lane_pre(const LaneParams LP) {
    const WarpParams &P = LP.w;
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P.n) return;
    const uint32_t mulc = LP.mulc;
    uint32_t l = 0, lm = 0;
    if (KIND == KIND_TREE) l = (uint32_t)(P.state_seed ? P.state_seed[i] : P.state_seed_u);
    if (MODE == 1) lm = (uint32_t)P.model_seeds[i];
#pragma unroll
    for (uint32_t k = 1; k <= (uint32_t)MT_M; ++k) {
        if (KIND == KIND_TREE) l = lane_seed_step(l, mulc, k);
        if (MODE == 1) lm = lane_seed_step(lm, mulc, k);
    }
    if (KIND == KIND_TREE) LP.l397[i] = l;
    if (MODE == 1) LP.lm397[i] = lm;
}

template <int KIND, int MODE>
__global__ void __launch_bounds__(32) k_lane(const LaneParams LP) {
    extern __shared__ __align__(16) uint32_t lane_smem[];
    const WarpParams &P = LP.w;
    const int lane = threadIdx.x;
    uint32_t *col = lane_smem + lane; /* X[k] of this lane's rollout: row k < keep at col[32 k], else in the ring */
    const ModConst m3 = {3u, 1431655765u, 1u};
    const uint32_t words_pad = LP.words_pad, keep = LP.keep, ring = LP.ring;
    const uint32_t mulc = LP.mulc, one = LP.one;
    uint32_t *const ring_begin = col + 32 * keep, *const ring_end = ring_begin + 32 * ring;

    unsigned long long acc_w = 0, acc_d = 0, acc_l = 0, acc_h = 0, acc_s = 0;
    const uint64_t chunks = (P.n + 31) / 32;
    /* seeds of the NEXT chunk are fetched one chunk ahead */
    uint32_t e_nx = 0, l_nx = 0, em_nx = 0, lm_nx = 0;
    auto fetch = [&](uint64_t chunk) {
        if (chunk >= chunks) return;
        const uint64_t j = chunk * 32 + lane < P.n ? chunk * 32 + lane : P.n - 1;
        if (KIND == KIND_TREE) {
            e_nx = (uint32_t)(P.state_seed ? P.state_seed[j] : P.state_seed_u);
            l_nx = LP.l397[j]; /* phase 1 ran in k_lane_pre */
        }
        if (MODE == 1) {
            em_nx = (uint32_t)P.model_seeds[j];
            lm_nx = LP.lm397[j];
        }
    };
    fetch(blockIdx.x);
    for (uint64_t chunk = blockIdx.x; chunk < chunks; chunk += gridDim.x) {
        const uint64_t i = chunk * 32 + lane;
        const bool live = i < P.n;
        const uint64_t ii = live ? i : P.n - 1; /* idle lanes replay the last rollout, write nothing */
        uint32_t e = e_nx, l = l_nx, em = em_nx, lm = lm_nx;
        fetch(chunk + gridDim.x);
        const uint32_t d = P.depth ? P.depth[ii] : P.depth_u;
        uint32_t pos = 0;
        int bias = 0;
        if (KIND == KIND_TREE) {
            pos = (P.state_pos ? P.state_pos[ii] : P.state_pos_u) - P.gen_words;
            bias = P.bias ? P.bias[ii] : P.bias_u;
        }
        const uint64_t soff = (MODE == 0) ? (P.step_off ? P.step_off[ii] : ii * (uint64_t)P.depth_u) : 0;
        uint32_t h = PCU_HASH_INIT;
        uint8_t *tr = (P.traj && live) ? P.traj + (size_t)i * P.traj_stride * 3 : nullptr;
        uint32_t mk1 = 1, mk398 = 398;

        /* ---- one step of every rollout of the chunk (phase 4); warp-uniform s */
        auto play = [&](uint32_t s) {
            uint32_t joint = 0;
            if (MODE == 0) {
                if (s < d)
                    joint = P.code_u16 ? (uint32_t)((const uint16_t *)P.codes)[soff + s]
                                       : (uint32_t)((const uint8_t *)P.codes)[soff + s];
            } else { /* the model chains advance in lock-step for every lane */
                const uint32_t a0 = lane_model_word(em, lm, mk1, mk398, mulc, one);
                const uint32_t a1 = lane_model_word(em, lm, mk1, mk398, mulc, one);
                const uint32_t b0 = lane_model_word(em, lm, mk1, mk398, mulc, one);
                const uint32_t b1 = lane_model_word(em, lm, mk1, mk398, mulc, one);
                joint = mod64(a0, a1, P.mrow) * P.cols + mod64(b0, b1, P.mcol);
            }
            if (s >= d) return;
            if (KIND == KIND_TREE) {
                pos += P.gen_words + joint; /* discard(idx), T == 1: chance action 0 (random-tree.hh:89) */
                uint32_t q0 = pos, q1 = pos + 1;
                if (q0 >= keep) q0 = keep + mod32(q0 - keep, ring, LP.ring_m);
                if (q1 >= keep) q1 = keep + mod32(q1 - keep, ring, LP.ring_m);
                const uint32_t hi = mt_temper(col[32 * q0]);
                const uint32_t lo = mt_temper(col[32 * q1]);
                bias += (int)mod64(hi, lo, m3) - 1; /* :97 */
                pos += 2;
            }
            h = hash_step(h, joint, 0);
            if (tr && s < P.traj_stride) { /* later steps are played and hashed, not recorded */
                const uint32_t r = joint / P.cols;
                tr[3 * s + 0] = (uint8_t)r;
                tr[3 * s + 1] = (uint8_t)(joint - r * P.cols);
                tr[3 * s + 2] = 0;
            }
        };

        if (KIND == KIND_TREE) {
            __syncwarp();
            /* ---- phases 2-3 in blocks of 8 rows; a play step runs as soon as every row it can read exists,
             * after which the rows below the next step's lowest reachable draw are dead and the ring reuses them */
            uint32_t k1 = 1, k398 = 398;
            uint32_t *dst = col;
            const uint32_t *src3 = col; /* X[k - 227] for phase 3: always a kept row */
            uint32_t s = 0, hi = LP.hi0; /* hi = highest draw index step s can read, over the whole batch */
            auto row2 = [&]() { /* X[k], k < 227, from the two chains in lock-step */
                const uint32_t en = lane_seed_step(e, mulc, k1);
                *dst = mt_twist_dev(e, en, l);
                dst += 32;
                e = en;
                l = lane_seed_step(l, mulc, k398);
                k1 = lane_fma_add(k1, one, 1);
                k398 = lane_fma_add(k398, one, 1);
            };
            auto row3 = [&]() { /* X[k], 227 <= k <= 622, from the lane's own kept rows */
                const uint32_t en = lane_seed_step(e, mulc, k1);
                *dst = mt_twist_dev(e, en, *src3);
                dst += 32;
                src3 += 32;
                e = en;
                k1 = lane_fma_add(k1, one, 1);
            };
            auto after_block = [&](uint32_t kb) {
                if (kb + 8 > keep && dst == ring_end) dst = ring_begin;
                while (s < LP.dmax && hi < kb + 8) {
                    __syncwarp();
                    play(s);
                    ++s;
                    hi += LP.hstep;
                }
            };
            const uint32_t n2 = words_pad < 224u ? words_pad : 224u;
            for (uint32_t kb = 0; kb < n2; kb += 8) {
#pragma unroll
                for (int r = 0; r < 8; ++r) row2();
                after_block(kb);
            }
            if (words_pad > 224u) {
                row2(); row2(); row2();                 /* 224..226 */
                row3(); row3(); row3(); row3(); row3(); /* 227..231 */
                after_block(224u);
                for (uint32_t kb = 232; kb < words_pad; kb += 8) {
#pragma unroll
                    for (int r = 0; r < 8; ++r) row3();
                    after_block(kb);
                }
            }
            for (; s < LP.dmax; ++s) play(s); /* not reached when hi0/hstep cover the batch; kept for safety */
        } else {
            for (uint32_t s = 0; s < LP.dmax; ++s) play(s);
        }

        if (live) {
            uint32_t pay = 0;
            if (KIND == KIND_TREE) pay = (uint32_t)((bias > 0) - (bias < 0) + 1);
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
    if (P.sums) {
        acc_w = warp_sum_u64(acc_w);
        acc_d = warp_sum_u64(acc_d);
        acc_l = warp_sum_u64(acc_l);
        acc_s = warp_sum_u64(acc_s);
        acc_h = warp_sum_u64(acc_h);
        if (lane == 0) {
            if (acc_w) atomicAdd(P.sums + 0, acc_w);
            if (acc_d) atomicAdd(P.sums + 1, acc_d);
            if (acc_l) atomicAdd(P.sums + 2, acc_l);
            if (acc_s) atomicAdd(P.sums + 3, acc_s);
            if (acc_h) atomicAdd(P.sums + 4, acc_h);
        }
    }
}

/* Largest draw index any rollout of the batch reads, + 2 (= LaneParams.words), for batches
 * whose positions / depths are per-rollout arrays: pos - gen + d (gen + rows cols - 1 + 2). */
static __global__ void k_lane_bound(uint64_t n, const uint32_t *pos, uint32_t pos_u, const uint16_t *depth, uint32_t depth_u,
                             uint32_t gen_words, uint32_t per_step, unsigned long long *out /* [0] words [1] dmax [2] max pos [3] ~min pos */) {
    unsigned long long mw = 0, md = 0, mp = 0, mq = 0; /* mq = max of ~pos: the smallest position */
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        const unsigned long long p = pos ? pos[i] : pos_u, d = depth ? depth[i] : depth_u;
        const unsigned long long w = p - gen_words + d * per_step;
        mw = w > mw ? w : mw;
        md = d > md ? d : md;
        mp = p > mp ? p : mp;
        mq = (0xffffffffull - p) > mq ? (0xffffffffull - p) : mq;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const unsigned long long a = __shfl_xor_sync(PCU_FULL, mw, o), b = __shfl_xor_sync(PCU_FULL, md, o);
        mw = a > mw ? a : mw;
        md = b > md ? b : md;
        const unsigned long long c = __shfl_xor_sync(PCU_FULL, mp, o), e = __shfl_xor_sync(PCU_FULL, mq, o);
        mp = c > mp ? c : mp;
        mq = e > mq ? e : mq;
    }
    if ((threadIdx.x & 31) == 0) {
        atomicMax(out + 0, mw);
        atomicMax(out + 1, md);
        atomicMax(out + 2, mp);
        atomicMax(out + 3, mq);
    }
}
