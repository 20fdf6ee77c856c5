// Leaf batches with per-rollout model words (pcu_model_rng mode PCU_MODEL_WORDS).
//
// A search that evaluates ONE leaf per iteration (TreeBandit::run_iteration, tree-bandit.hh:103,111) is batched by
// running many independent searches and collecting their parked single-leaf inferences (include/pinyon_cuda/
// batcher.hh).  Every tree keeps its own sequential model engine, so a batch has no common model stream: the host
// hands over, per rollout, the 4 * depth RAW engine words its model would draw next (the host advances that engine
// anyway to keep `model.device` in lock-step; tempering, the uniform_int arithmetic of random.hh:26,28 and the whole
// rollout run here).
//
//   k_leaf_table     leaves that share one state stream (all leaves of searches from one root: same state seed):
//                    one rollout per thread against the state-stream table, words read straight from the batch
//   k_words_to_codes everything else: words -> joint action codes, then k_lane / k_warp as in mode S
#pragma once

struct LeafParams {
    uint64_t n;
    uint32_t kind; /* KIND_TREE / KIND_MOLD */
    uint32_t cols, T, gen_words;
    FoldConst mrow, mcol;
    const uint32_t *words;    /* [4 * total_steps] raw engine words, rollout-major */
    const uint64_t *step_off; /* first step of rollout i; null: i * depth_u */
    const uint16_t *depth;
    uint32_t depth_u;
    const uint8_t *tab; /* state-stream table (host_build_table) */
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
};

static __global__ void __launch_bounds__(128) k_leaf_table(const LeafParams P) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long aw = 0, ad = 0, al = 0, ah = 0, as = 0;
    if (i < P.n) {
        const uint32_t d = P.depth ? P.depth[i] : P.depth_u;
        const uint64_t soff = P.step_off ? P.step_off[i] : i * (uint64_t)P.depth_u;
        const uint4 *w4 = reinterpret_cast<const uint4 *>(P.words) + soff;
        uint8_t *tr = P.traj ? P.traj + (size_t)i * P.traj_stride * 3 : nullptr;
        uint32_t h = PCU_HASH_INIT, pay = 0;
        if (P.kind == KIND_MOLD) {
            for (uint32_t s = 0; s < d; ++s) {
                const uint32_t code = gw_code(w4[s], P.mrow, P.mcol, P.cols);
                h = hash_step(h, code, 0);
                if (tr) put_traj<true>(tr, s, code, P.cols, 0, P.traj_stride);
            }
            ad = 1;
        } else {
            const int8_t *tb = reinterpret_cast<const int8_t *>(P.tab);
            const uint8_t *tch = P.tab + P.tab_len;
            uint32_t g = (P.state_pos ? P.state_pos[i] : P.state_pos_u) - P.tab_base - P.gen_words; /* where the current
                                                                                      chance strategies were drawn */
            int bias = P.bias ? P.bias[i] : P.bias_u;
            const uint32_t T = P.T, gen = P.gen_words;
            for (uint32_t s = 0; s < d; ++s) {
                const uint32_t code = gw_code(w4[s], P.mrow, P.mcol, P.cols);
                const uint32_t st = code * T;
                const uint32_t ch = T > 1 ? (uint32_t)tch[g + 2 * st] : 0u; /* random-tree.hh:115-135 */
                const uint32_t t = g + gen + st + ch;                       /* discard(get_transition_idx) :88-89 */
                bias += tb[t];                                               /* :97 */
                g = t + 2;
                h = hash_step(h, code, ch);
                if (tr) put_traj<true>(tr, s, code, P.cols, ch, P.traj_stride);
            }
            pay = (uint32_t)((bias > 0) - (bias < 0) + 1);
            aw = pay == 2;
            ad = pay == 1;
            al = pay == 0;
        }
        if (tr)
            for (uint32_t s = d; s < P.traj_stride; ++s) tr[3 * s] = tr[3 * s + 1] = tr[3 * s + 2] = 0;
        if (P.payoff) P.payoff[i] = (uint8_t)pay;
        if (P.traj_hash) P.traj_hash[i] = h;
        ah = h;
        as = d;
    }
    if (P.sums) {
        aw = warp_sum_u64(aw);
        ad = warp_sum_u64(ad);
        al = warp_sum_u64(al);
        ah = warp_sum_u64(ah);
        as = warp_sum_u64(as);
        if ((threadIdx.x & 31) == 0) {
            if (aw) atomicAdd(P.sums + 0, aw);
            if (ad) atomicAdd(P.sums + 1, ad);
            if (al) atomicAdd(P.sums + 2, al);
            if (as) atomicAdd(P.sums + 3, as);
            if (ah) atomicAdd(P.sums + 4, ah);
        }
    }
}

template <typename CodeT>
__global__ void __launch_bounds__(256) k_words_to_codes(const uint32_t *__restrict__ words, uint64_t total_steps, uint32_t cols,
                                                        FoldConst mrow, FoldConst mcol, CodeT *__restrict__ codes) {
    const uint64_t s = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= total_steps) return;
    codes[s] = (CodeT)gw_code(reinterpret_cast<const uint4 *>(words)[s], mrow, mcol, cols);
}
