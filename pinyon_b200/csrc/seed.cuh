// Seeding for k_warp (private engines whose draws leave the first refill, or T > 1).
//
// std::mt19937::seed (bits/random.tcc; reached from prng(seed), random.hh:16) is a 623-step
// serial chain.  k_warp used to run it on ONE lane per engine while the other 31 idled: 30 % of the
// kernel on config B with distinct roots (profiles/README.md r1d).  k_seed_pre runs the chains
// lane-per-rollout, 32 per instruction, and writes the seeded arrays rollout-major to a scratch slab
// in HBM through a 32 x 33 shared-memory transpose (128-byte coalesced rows); k_warp then loads its
// rollout's 624 words with 20 coalesced reads.  The throw-away prng{transition_seed} of
// random-tree.hh:118 needs only x[0..2], x[397], x[398]: its uniform() is finished here.
#pragma once

struct SeedParams {
    uint64_t i0, count;           /* slab = rollouts [i0, i0 + count) */
    const uint64_t *state_seed;   /* null: state_seed_u */
    uint64_t state_seed_u;
    const uint64_t *model_seeds;  /* mode P, else null */
    const uint64_t *tseed;        /* null: tseed_u */
    uint64_t tseed_u;
    uint32_t want_state, want_model, want_pu;
    uint32_t mulc;                /* 1812433253 as data (see lane.cuh) */
    uint32_t *xs;                 /* [count][624] seeded state engines */
    uint32_t *xm;                 /* [count][624] seeded model engines */
    double *pu;                   /* [count] prng{transition_seed}.uniform() */
};

static __global__ void __launch_bounds__(128) k_seed_pre(const SeedParams S) {
    __shared__ uint32_t tile_all[4][2][32 * 33];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    uint32_t *ts = tile_all[wib][0], *tm = tile_all[wib][1];
    const uint64_t r0 = ((uint64_t)blockIdx.x * 4 + wib) * 32; /* first rollout of this warp, slab-relative */
    if (r0 >= S.count) return;
    const uint64_t r = r0 + lane < S.count ? r0 + lane : S.count - 1;
    const uint64_t i = S.i0 + r;
    const uint32_t nrows = (uint32_t)(S.count - r0 < 32 ? S.count - r0 : 32);
    const uint32_t mulc = S.mulc;
    uint32_t vs = (uint32_t)(S.state_seed ? S.state_seed[i] : S.state_seed_u);
    uint32_t vm = S.model_seeds ? (uint32_t)S.model_seeds[i] : 0u;
    uint32_t vt = (uint32_t)(S.tseed ? S.tseed[i] : S.tseed_u);
    const uint32_t t0 = vt;
    uint32_t t1 = 0, t2 = 0, t397 = 0, t398 = 0;
#pragma unroll 1
    for (uint32_t kb = 0; kb < MT_N; kb += 32) {
        const uint32_t len = MT_N - kb < 32 ? MT_N - kb : 32;
#pragma unroll 4
        for (uint32_t j = 0; j < len; ++j) {
            const uint32_t k = kb + j;
            if (k) {
                if (S.want_state) vs = lane_seed_step(vs, mulc, k);
                if (S.want_model) vm = lane_seed_step(vm, mulc, k);
                if (S.want_pu && k <= 398) {
                    vt = lane_seed_step(vt, mulc, k);
                    if (k == 1) t1 = vt;
                    if (k == 2) t2 = vt;
                    if (k == 397) t397 = vt;
                    if (k == 398) t398 = vt;
                }
            }
            if (S.want_state) ts[j * 33 + lane] = vs;
            if (S.want_model) tm[j * 33 + lane] = vm;
        }
        __syncwarp();
        if (lane < (int)len) {
            for (uint32_t q = 0; q < nrows; ++q) {
                if (S.want_state) S.xs[(r0 + q) * MT_N + kb + lane] = ts[lane * 33 + q];
                if (S.want_model) S.xm[(r0 + q) * MT_N + kb + lane] = tm[lane * 33 + q];
            }
        }
        __syncwarp();
    }
    if (S.want_pu && r0 + lane < S.count) {
        const uint32_t o0 = mt_temper(mt_twist(t0, t1, t397));
        const uint32_t o1 = mt_temper(mt_twist(t1, t2, t398));
        S.pu[r0 + lane] = prng_uniform_from_words(o0, o1);
    }
}
