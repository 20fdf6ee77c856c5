// k_warp2 — RandomTree rollouts with PRIVATE state engines whose draws leave the first refill (distinct roots with
// several chance transitions: BASELINE config 3 with distinct roots), model-rng modes S and W (joint action codes from
// k_gen_codes2 / k_words_to_codes).  Replaces k_seed_pre + k_warp<TREE, 0> for that case (round 2);
// mode P and MoldState stay on k_warp.
//
// What a rollout costs here is its state engine: get_chance_strategies (random-tree.hh:154-199) draws
// 2*rows*cols*T words per step of which a rollout reads 2T + 2, so on config 3 the engine is refilled ~20 times per
// rollout and every word of every refill has to be made (X[n] = X[n-227] ^ f(X[n-624], X[n-623]) reaches all of
// them; evaluating only the dependency cone of the words read is sparser for one step and denser than the full
// stream after three).  The kernel is therefore a twist loop with a few reads, one persistent CTA per SM:
//
//  * a pool of W2_SLOTS rings of 640 words in shared memory, used in FIFO order by the CTA's rollouts;
//  * TWO seeder warps (on different schedulers) run std::mt19937::seed (623 dependent steps, bits/random.tcc)
//    lane-per-rollout, 16 rollouts per pass, straight into free rings with 128-bit stores (ring stride 644 words =
//    4 * odd: a quarter warp's stores cover all 32 banks), and finish prng{transition_seed}.uniform()
//    (random-tree.hh:118) beside it; the seeded engines never visit HBM (k_seed_pre wrote and k_warp re-read 2.5 KB
//    per rollout);
//  * the player warps each take the next seeded rollout(s) (shared-memory counter) and play them in place: draw j
//    lives in slot j mod 640 (the seeded array sits in slots 16..639, so draw 0 lands in slot 0); a group is 128
//    draws, lane l makes draws 4l..4l+3 of it from two 128-bit loads, two shuffles and one 128-bit store at
//    compile-time slots + 4*lane (five phases laid out as one unrolled cycle; the ring is 16 words longer than the
//    state, so a group never overwrites an operand it still needs): 38 instructions and 14 shared-memory wavefronts
//    per 128 draws against ~44 and 16 of k_warp's in-place refill; positions are 32-bit (the host keeps batches whose
//    draw indices could pass 2^31 on k_warp), a rollout's codes are fetched several steps at a time;
//  * LOCK (1, 2 or 4; option warp2_lock) rollouts per player warp in lock-step: the engines advance one after the
//    other with the whole warp, the rest of a step runs on 32 / LOCK lanes per rollout.  Fewer instructions per
//    rollout (r2q ncu: 8955 / 7708 / 6985) but fewer warps to hide the twist chain's latency (issue slots busy
//    86 % / 79 % / 66 %): 1.22e8 / 1.20e8 / 1.11e8 rollouts/s on config 3 with distinct roots, k_warp 1.04e8.
#pragma once

#define W2_RING 640
#define W2_GROUP 128
#define W2_STRIDE 644 /* words between two rings: 4 * 161; words 640..641 hold the rollout's uniform() as a double */
#define W2_SLOTS 80   /* multiple of 8 (bank pattern of the seeder's stores across the pool's wrap) */
/* LOCK = rollouts a player warp steps in lock-step (32 / LOCK lanes each; transitions <= 8 needs LOCK <= 4);
 * players * LOCK rollouts in flight + one seeder pass <= W2_SLOTS */
template <int LOCK>
struct W2Cfg {
    static constexpr int PLAYERS = LOCK == 1 ? 30 : LOCK == 2 ? 24 : 16;
    static constexpr int NSEED = 2;                   /* seeder warps (on different schedulers) */
    static constexpr int SEEDB = LOCK == 4 ? 8 : 16;  /* rollouts per seeder pass */
    static constexpr int THREADS = (PLAYERS + NSEED) * 32;
    static constexpr int GL = 32 / LOCK; /* lanes per rollout */
};

struct Warp2Params {
    WarpParams w;
    double pu_u;   /* prng{tseed_u}.uniform() when the transition seed is batch-uniform (w.tseed == null) */
    uint32_t mulc; /* 1812433253 as data (lane.cuh) */
};

struct W2Ctl {
    uint32_t seeded; /* rollouts of this CTA seeded so far (written by the seeder's lane 0) */
    uint32_t taken;  /* next rollout to play (atomicAdd by the players) */
    uint32_t released[W2_SLOTS]; /* how many rollouts have finished with ring s */
};

/* slot of (LO + 4 * lane) in the 640-word ring; a per-lane select only where the range straddles the end */
template <int LO>
__device__ __forceinline__ uint32_t w2_slot(int lane) {
    if constexpr (LO + 127 < W2_RING) {
        return (uint32_t)(LO + 4 * lane);
    } else if constexpr (LO >= W2_RING) {
        return (uint32_t)(LO - W2_RING + 4 * lane);
    } else {
        const int s = LO + 4 * lane;
        return (uint32_t)(s >= W2_RING ? s - W2_RING : s);
    }
}

/* Draws 128 * g + 4 * lane + {0..3}, g = PH (mod 5).  Draw j sits in slot j mod 640; its operands X[j-624], X[j-623],
 * X[j-227] in slots j + 16, j + 17, j + 413 (mod 640). */
template <int PH>
__device__ __forceinline__ void w2_group(uint32_t *ring, int lane) {
    constexpr int P0 = PH * W2_GROUP;
    const uint4 a = *reinterpret_cast<const uint4 *>(ring + w2_slot<P0 + 16>(lane));
    const uint4 c = *reinterpret_cast<const uint4 *>(ring + w2_slot<P0 + 412>(lane));
    /* the fifth word of each operand run is the next lane's first: a shuffle instead of a 32-bit load whose 16-byte
     * lane stride makes it a 4-way bank conflict (r2m ncu: 20 wavefronts per group, 78 % of the shared-memory pipe);
     * lane 31's lies past the warp's 128 words */
    uint32_t a4 = __shfl_down_sync(PCU_FULL, a.x, 1);
    uint32_t c4 = __shfl_down_sync(PCU_FULL, c.x, 1);
    if (lane == 31) {
        a4 = ring[(P0 + 16 + W2_GROUP) % W2_RING];
        c4 = ring[(P0 + 412 + W2_GROUP) % W2_RING];
    }
    uint4 o;
    o.x = mt_twist_dev(a.x, a.y, c.y);
    o.y = mt_twist_dev(a.y, a.z, c.z);
    o.z = mt_twist_dev(a.z, a.w, c.w);
    o.w = mt_twist_dev(a.w, a4, c4);
    __syncwarp(); /* every lane has read before anyone writes (lanes l+4, l+5 write what lane l read as X[j-624]) */
    *reinterpret_cast<uint4 *>(ring + P0 + 4 * lane) = o;
    __syncwarp();
}

/* Make draws until `need` of them exist; G = draws made so far (multiple of 128), ph = (G / 128) mod 5.  The five
 * phases are laid out as one unrolled cycle entered at the current phase (a step of config 3 needs ~5 groups = one
 * cycle): per group one add, one compare and one branch beside the group itself. */
__device__ __forceinline__ void w2_gen_inline(uint32_t *ring, uint32_t &G, uint32_t &ph, uint32_t need, int lane) {
    if (G >= need) return;
    switch (ph) {
        do {
            case 0:
                w2_group<0>(ring, lane);
                G += W2_GROUP;
                if (G >= need) {
                    ph = 1;
                    goto done;
                }
            case 1:
                w2_group<1>(ring, lane);
                G += W2_GROUP;
                if (G >= need) {
                    ph = 2;
                    goto done;
                }
            case 2:
                w2_group<2>(ring, lane);
                G += W2_GROUP;
                if (G >= need) {
                    ph = 3;
                    goto done;
                }
            case 3:
                w2_group<3>(ring, lane);
                G += W2_GROUP;
                if (G >= need) {
                    ph = 4;
                    goto done;
                }
            default:
                w2_group<4>(ring, lane);
                G += W2_GROUP;
                if (G >= need) {
                    ph = 0;
                    goto done;
                }
        } while (true);
    }
done:;
}

/* one copy for the lock-step players: inlined at 2 * LOCK call sites it outgrew the instruction cache (r2n ncu) */
static __device__ __noinline__ uint2 w2_gen_call(uint32_t *ring, uint32_t G, uint32_t ph, uint32_t need, int lane) {
    w2_gen_inline(ring, G, ph, need, lane);
    return make_uint2(G, ph);
}

template <int LOCK>
__device__ __forceinline__ void w2_gen(uint32_t *ring, uint32_t &G, uint32_t &ph, uint32_t need, int lane) {
    if constexpr (LOCK == 1) {
        w2_gen_inline(ring, G, ph, need, lane);
    } else {
        if (G < need) {
            const uint2 t = w2_gen_call(ring, G, ph, need, lane);
            G = t.x;
            ph = t.y;
        }
    }
}

__device__ __forceinline__ uint32_t w2_draw(const uint32_t *ring, uint32_t j) { return mt_temper(ring[j % W2_RING]); }

/* A seeder warp: passes si, si + NSEED, ... of SEEDB rollouts each over [c0, c0 + count), lane-per-rollout.
 * Seeding must stay ahead of the players although its 623-step chain is pure latency and shares a scheduler with
 * seven busy warps (r2p ncu: with one 32-lane seeder the players spun 23 times per rollout step): two seeders on
 * different schedulers, a loop of nothing but the chain (17 instructions per four words), passes published in order. */
template <int SEEDB, int NSEED, bool WANT_PU>
__device__ __forceinline__ void w2_seeder(const Warp2Params &P, W2Ctl *ctl, uint32_t *rings, uint64_t c0, uint32_t count,
                                          int lane, int si) {
    const WarpParams &W = P.w;
    const uint32_t mulc = P.mulc;
    static_assert(SEEDB <= 16, "lanes SEEDB .. 2 SEEDB - 1 run the transition-seed chains");
    for (uint32_t r0 = si * SEEDB; r0 < count; r0 += NSEED * SEEDB) {
        /* lanes 0 .. SEEDB-1 run the state chains of the pass's rollouts; with per-rollout transition seeds lanes
         * SEEDB .. 2 SEEDB-1 run prng{transition_seed}'s chain of the same rollouts in the same instructions (only
         * x[0..2], x[397], x[398] matter for its uniform(), random-tree.hh:118) */
        const bool pu_lane = WANT_PU && lane >= SEEDB && lane < 2 * SEEDB;
        const uint32_t r = r0 + (pu_lane ? lane - SEEDB : lane);
        const bool active = lane < SEEDB && r < count;
        const uint32_t slot = r % W2_SLOTS, use = r / W2_SLOTS;
        if (active) {
            const volatile uint32_t *rel = &ctl->released[slot];
            while (*rel != use) __nanosleep(500); /* the ring's previous rollout is still being played */
        }
        __syncwarp();
        __threadfence_block();
        const uint64_t i = c0 + (r < count ? r : count - 1);
        uint32_t *ring = rings + (size_t)slot * W2_STRIDE;
        if (!WANT_PU && active && W.T > 1) *reinterpret_cast<double *>(ring + W2_RING) = P.pu_u;
        /* x[4q .. 4q+3] -> slots 16 + 4q ..: one 128-bit store per four chain steps */
        uint4 o;
        o.x = pu_lane ? (uint32_t)W.tseed[i] : (uint32_t)(W.state_seed ? W.state_seed[i] : W.state_seed_u);
        o.y = lane_seed_step(o.x, mulc, 1);
        o.z = lane_seed_step(o.y, mulc, 2);
        o.w = lane_seed_step(o.z, mulc, 3);
        const uint32_t t0 = o.x, t1 = o.y, t2 = o.z;
        uint32_t t397 = 0, t398 = 0;
        uint32_t *dst = ring + 16;
        if (active) *reinterpret_cast<uint4 *>(dst) = o;
#pragma unroll 2
        for (uint32_t k = 4; k < MT_N; k += 4) {
            o.x = lane_seed_step(o.w, mulc, k);
            o.y = lane_seed_step(o.x, mulc, k + 1);
            o.z = lane_seed_step(o.y, mulc, k + 2);
            o.w = lane_seed_step(o.z, mulc, k + 3);
            if (WANT_PU && k == 396) {
                t397 = o.y;
                t398 = o.z;
            }
            if (active) *reinterpret_cast<uint4 *>(dst + k) = o;
        }
        if (pu_lane && r < count)
            *reinterpret_cast<double *>(ring + W2_RING) =
                prng_uniform_from_words(mt_temper(mt_twist(t0, t1, t397)), mt_temper(mt_twist(t1, t2, t398)));
        __threadfence_block();
        __syncwarp();
        if (lane == 0) {
            volatile uint32_t *seeded = &ctl->seeded;
            while (*seeded != r0) __nanosleep(200); /* publish in order */
            *seeded = (r0 + SEEDB < count) ? r0 + SEEDB : count;
        }
        __syncwarp();
    }
}

/* LOCK rollouts (r0 .. r0 + nr - 1 of the CTA) played in lock-step by one warp: the engines are advanced one after
 * the other with the whole warp (w2_gen_until), everything else of a step (random-tree.hh:87-135) runs on eight lanes
 * per rollout — lane k of a group owns chance transition k — so its ~150 instructions serve four rollouts. */
template <int LOCK>
__device__ __forceinline__ void w2_play(const WarpParams &W, uint32_t *rings, const double *fq_s, uint64_t c0, uint32_t r0,
                                        uint32_t nr, int lane, unsigned long long &acc_w, unsigned long long &acc_d,
                                        unsigned long long &acc_l, unsigned long long &acc_h, unsigned long long &acc_s) {
    const ModConst m10 = {10u, 429496729u, 6u};
    const ModConst m3 = {3u, 1431655765u, 1u};
    constexpr int GL = W2Cfg<LOCK>::GL, CB = GL < 8 ? GL : 8; /* codes fetched CB steps at a time */
    const int g = lane / GL, k = lane % GL, g0 = lane - k;
    const bool valid = (uint32_t)g < nr;
    const uint32_t r = r0 + (valid ? g : 0);
    const uint64_t i = c0 + r;
    uint32_t *myring = rings + (size_t)(r % W2_SLOTS) * W2_STRIDE;
    uint32_t *ring_of[LOCK];
#pragma unroll
    for (int gg = 0; gg < LOCK; ++gg) ring_of[gg] = rings + (size_t)((r0 + gg) % W2_SLOTS) * W2_STRIDE;
    const uint32_t d = valid ? (W.depth ? W.depth[i] : W.depth_u) : 0u;
    const uint32_t dmax = __reduce_max_sync(PCU_FULL, d);
    const uint32_t T = W.T, gen = W.gen_words;
    const double pu = T > 1 ? *reinterpret_cast<const double *>(myring + W2_RING) : 0.0;
    uint32_t cur = (W.state_pos ? W.state_pos[i] : W.state_pos_u) - gen; /* where the current strategies were drawn */
    int bias = W.bias ? W.bias[i] : W.bias_u;
    const uint64_t soff = W.step_off ? W.step_off[i] : i * (uint64_t)W.depth_u;
    uint32_t h = PCU_HASH_INIT;
    uint8_t *tr = (W.traj && valid && k == 0) ? W.traj + (size_t)i * W.traj_stride * 3 : nullptr;
    uint32_t G[LOCK], ph[LOCK];
#pragma unroll
    for (int gg = 0; gg < LOCK; ++gg) G[gg] = ph[gg] = 0;
    uint32_t mycode = 0;

    for (uint32_t s = 0; s < dmax; ++s) {
        const bool act = LOCK == 1 ? true : s < d; /* one rollout per warp: dmax == d */
        if (s % CB == 0) { /* the next CB joint action codes of this group's rollout, one per lane */
            mycode = 0;
            if (k < CB && s + k < d)
                mycode = W.code_u16 ? (uint32_t)((const uint16_t *)W.codes)[soff + s + k]
                                    : (uint32_t)((const uint8_t *)W.codes)[soff + s + k];
        }
        const uint32_t joint = __shfl_sync(PCU_FULL, mycode, g0 + (s % CB));
        const uint32_t st = joint * T;
        uint32_t ch = 0;
        if (T > 1) {
            /* lane k < T reads the two words strategy entry k of the chosen cell was drawn from (:169) */
            const uint32_t q = cur + 2 * st;
            const uint32_t need = act ? q + 2 * T : 0u;
#pragma unroll
            for (int gg = 0; gg < LOCK; ++gg)
                w2_gen<LOCK>(ring_of[gg], G[gg], ph[gg], LOCK == 1 ? need : __shfl_sync(PCU_FULL, need, GL * gg), lane);
            int nk = 0;
            if constexpr (GL >= 16) { /* one lane per word: entry e on lane 2e (branch-free: idle lanes read a
                                         valid slot and drop the result) */
                const uint32_t w = w2_draw(myring, q + k);
                const uint32_t lo = __shfl_down_sync(PCU_FULL, w, 1);
                int v = (int)mod64(w, lo, m10) + 1;
                if (v * W.thr_q < W.thr_p * 10) v = 0;
                nk = (act && k < (int)(2 * T) && !(k & 1)) ? v : 0;
            } else { /* eight lanes per rollout: entry e on lane e, both words */
                const uint32_t hi = w2_draw(myring, q + 2 * k), lo = w2_draw(myring, q + 2 * k + 1);
                int v = (int)mod64(hi, lo, m10) + 1;
                if (v * W.thr_q < W.thr_p * 10) v = 0;
                nk = (act && k < (int)T) ? v : 0;
            }
            int S;
            if constexpr (LOCK == 1) {
                S = __reduce_add_sync(PCU_FULL, nk);
            } else {
                S = nk;
#pragma unroll
                for (int o = 1; o < (GL < 16 ? GL : 16); o <<= 1) S += __shfl_xor_sync(PCU_FULL, S, o);
            }
            if (S == 0) {
                if (k == 0) nk = 1;
                S = 1;
            }
            double pp = pu;
            bool found = false;
            for (uint32_t kk = 0; kk < T; ++kk) {
                const int n_k = __shfl_sync(PCU_FULL, nk, g0 + (GL >= 16 ? 2 * kk : kk));
                pp -= fq_s[(uint32_t)n_k * W.fq_w + (uint32_t)S];
                if (!found && pp <= 0) {
                    ch = kk;
                    found = true;
                }
            }
        }
        if (act) cur += gen + st + ch; /* discard(get_transition_idx) :88-89 */
        {
            const uint32_t need = act ? cur + 2 : 0u;
#pragma unroll
            for (int gg = 0; gg < LOCK; ++gg)
                w2_gen<LOCK>(ring_of[gg], G[gg], ph[gg], LOCK == 1 ? need : __shfl_sync(PCU_FULL, need, GL * gg), lane);
        }
        {
            uint32_t hi, lo;
            if constexpr (GL >= 16) { /* lanes 0 and 1 of the group temper one word each */
                const uint32_t w = w2_draw(myring, cur + (k & 1));
                hi = __shfl_sync(PCU_FULL, w, g0);
                lo = __shfl_sync(PCU_FULL, w, g0 + 1);
            } else {
                hi = w2_draw(myring, cur);
                lo = w2_draw(myring, cur + 1);
            }
            const int b1 = (int)mod64(hi, lo, m3) - 1; /* :97 */
            const uint32_t h1 = hash_step(h, joint, ch);
            if (act) {
                bias += b1;
                cur += 2;
                h = h1;
                if (tr && s < W.traj_stride) {
                    const uint32_t row = joint / W.cols;
                    tr[3 * s + 0] = (uint8_t)row;
                    tr[3 * s + 1] = (uint8_t)(joint - row * W.cols);
                    tr[3 * s + 2] = (uint8_t)ch;
                }
            }
        }
    }
    if (valid && k == 0) {
        const uint32_t pay = (uint32_t)((bias > 0) - (bias < 0) + 1);
        if (tr)
            for (uint32_t s = d; s < W.traj_stride; ++s) tr[3 * s] = tr[3 * s + 1] = tr[3 * s + 2] = 0;
        if (W.payoff) W.payoff[i] = (uint8_t)pay;
        if (W.traj_hash) W.traj_hash[i] = h;
        acc_w += (pay == 2);
        acc_d += (pay == 1);
        acc_l += (pay == 0);
        acc_h += h;
        acc_s += d;
    }
}

template <int LOCK>
__global__ void __launch_bounds__(W2Cfg<LOCK>::THREADS, 1) k_warp2(const Warp2Params P) {
    extern __shared__ __align__(16) unsigned char w2_smem[];
    const WarpParams &W = P.w;
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    uint32_t *rings = reinterpret_cast<uint32_t *>(w2_smem);
    double *fq_s = reinterpret_cast<double *>(w2_smem + sizeof(uint32_t) * W2_SLOTS * W2_STRIDE);
    W2Ctl *ctl = reinterpret_cast<W2Ctl *>(fq_s + (W.T > 1 ? 11u * W.fq_w : 0u));
    if (W.T > 1)
        for (uint32_t k = threadIdx.x; k < 11u * W.fq_w; k += blockDim.x) fq_s[k] = W.fq[k];
    if (threadIdx.x < W2_SLOTS) ctl->released[threadIdx.x] = 0;
    if (threadIdx.x == 0) ctl->seeded = ctl->taken = 0;
    __syncthreads();
    /* this CTA's contiguous share of the batch */
    const uint64_t per = (W.n + gridDim.x - 1) / gridDim.x;
    const uint64_t c0 = per * blockIdx.x;
    const uint32_t count = c0 >= W.n ? 0u : (uint32_t)(W.n - c0 < per ? W.n - c0 : per);
    if (wib >= W2Cfg<LOCK>::PLAYERS) {
        if (W.T > 1 && W.tseed)
            w2_seeder<W2Cfg<LOCK>::SEEDB, W2Cfg<LOCK>::NSEED, true>(P, ctl, rings, c0, count, lane, wib - W2Cfg<LOCK>::PLAYERS);
        else
            w2_seeder<W2Cfg<LOCK>::SEEDB, W2Cfg<LOCK>::NSEED, false>(P, ctl, rings, c0, count, lane, wib - W2Cfg<LOCK>::PLAYERS);
        return;
    }
    unsigned long long acc_w = 0, acc_d = 0, acc_l = 0, acc_h = 0, acc_s = 0;
    for (;;) {
        uint32_t r0 = 0;
        if (lane == 0) r0 = atomicAdd(&ctl->taken, (uint32_t)LOCK);
        r0 = __shfl_sync(PCU_FULL, r0, 0);
        if (r0 >= count) break;
        const uint32_t nr = count - r0 < LOCK ? count - r0 : LOCK;
        const volatile uint32_t *seeded = &ctl->seeded;
        while (*seeded < r0 + nr) __nanosleep(400);
        __threadfence_block();
        w2_play<LOCK>(W, rings, fq_s, c0, r0, nr, lane, acc_w, acc_d, acc_l, acc_h, acc_s);
        __syncwarp();
        __threadfence_block();
        if (lane < (int)nr) *(volatile uint32_t *)&ctl->released[(r0 + lane) % W2_SLOTS] = (r0 + lane) / W2_SLOTS + 1;
    }
    if (W.sums) {
        acc_w = warp_sum_u64(acc_w);
        acc_d = warp_sum_u64(acc_d);
        acc_l = warp_sum_u64(acc_l);
        acc_h = warp_sum_u64(acc_h);
        acc_s = warp_sum_u64(acc_s);
        if (lane == 0) {
            if (acc_w) atomicAdd(W.sums + 0, acc_w);
            if (acc_d) atomicAdd(W.sums + 1, acc_d);
            if (acc_l) atomicAdd(W.sums + 2, acc_l);
            if (acc_s) atomicAdd(W.sums + 3, acc_s);
            if (acc_h) atomicAdd(W.sums + 4, acc_h);
        }
    }
}
