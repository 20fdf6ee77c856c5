// k_thread — the VARIANT path: one rollout per thread, a direct transliteration of the reference's rollout loop
// (monte-carlo-model.hh:85-96 over random-tree.hh:87-135,154-199 / test-states.hh:43-49) with the pieces the fast
// kernels fix at compile time left open:
//   * the PRNG type: Pinyon's `prng` (std::mt19937, random.hh:8-64; the engine lives in the thread's local memory)
//     or `xor_shift` (random.hh:66-121; 64 bits of state in registers);
//   * the growth functions: the defaults (random-tree.hh:141-146) or GrowLib::alternating_constant_actions
//     (libpinyon/grow-lib.hh:16-24) as actions_func, under which rows and cols change from step to step.
// Batches of the default type list with default growth never come here (k_stream2 / k_lane / k_warp serve them).
// Lanes run in lock-step on their own engines; nothing is shared, so there is no synchronisation at all.
#pragma once

enum { PCU_ENG_MT19937 = 0, PCU_ENG_XORSHIFT = 1 };
enum { PCU_GROW_DEFAULT = 0, PCU_GROW_ALTERNATING_CONSTANT = 1 };
enum { TH_MODEL_SEEDS = 0, TH_MODEL_WORDS = 1, TH_MODEL_XS_STREAM = 2 };

struct ThreadParams {
    uint64_t n;
    uint32_t kind; /* KIND_TREE / KIND_MOLD */
    uint32_t rows, cols, T, growth;
    int32_t thr_p, thr_q;
    const uint64_t *state_seed; /* MT: prng seed; xor_shift: the engine's CURRENT 64-bit state */
    uint64_t state_seed_u;
    const uint32_t *state_pos; /* MT only: draws made since seeding */
    uint32_t state_pos_u;
    const uint64_t *tseed;
    uint64_t tseed_u;
    const uint16_t *depth;
    uint32_t depth_u;
    const int32_t *bias;
    int32_t bias_u;
    const double *fq;
    uint32_t fq_w;
    uint32_t model_src;
    const uint64_t *model_seeds; /* TH_MODEL_SEEDS */
    const uint32_t *words;       /* TH_MODEL_WORDS: four std::mt19937 words per step, rollout-major (16-byte aligned) */
    int32_t words_raw;           /* 1: untempered state words (mode W); 0: engine outputs (mode S, k_gen_words) */
    const uint64_t *step_off; /* first step of rollout i (null: i * depth_u) */
    uint64_t xs_m0;           /* TH_MODEL_XS_STREAM: the model engine's state before the batch */
    const uint64_t *xs_pow;   /* [64][64]: column b of A^(2^j), A = one xor_shift step */
    uint8_t *payoff;
    uint8_t *traj;
    uint32_t traj_stride;
    uint32_t *traj_hash;
    unsigned long long *sums;
};

/* ---- xor_shift (random.hh:115-120): state ^= state << 21; state ^= state >> 35; state ^= state << 4 */
__host__ __device__ __forceinline__ uint64_t xs_step(uint64_t s) {
    s ^= s << 21;
    s ^= s >> 35;
    s ^= s << 4;
    return s;
}
/* the inverse step: each xor-shift is undone by repeating it with doubled distances while they stay below 64 */
__host__ __device__ __forceinline__ uint64_t xs_unstep(uint64_t s) {
    s ^= s << 4;
    s ^= s << 8;
    s ^= s << 16;
    s ^= s << 32;
    s ^= s >> 35;
    s ^= s << 21;
    s ^= s << 42;
    return s;
}

struct EngXs {
    static constexpr uint32_t UNITS = 1; /* engine steps per 64-bit draw */
    uint64_t s;
    uint64_t at; /* draws made, relative to an arbitrary origin */
    __device__ __forceinline__ void position(uint64_t abs) {
        for (; at < abs; ++at) s = xs_step(s);
    }
    __device__ __forceinline__ uint64_t u64() {
        s = xs_step(s);
        ++at;
        return s;
    }
    /* xor_shift::uniform (random.hh:79): double(x) / double(UINT64_MAX); the divisor rounds to 2^64 */
    __device__ __forceinline__ double uniform() { return __dmul_rn((double)u64(), 5.421010862427522170037e-20); }
};

/* std::mt19937 in libstdc++'s form, in the thread's local memory; `at` counts 32-bit draws since seeding. */
struct EngMt {
    static constexpr uint32_t UNITS = 2; /* 32-bit draws per 64-bit draw (random.hh:28: high word first) */
    uint32_t x[MT_N];
    uint32_t p;
    uint64_t at;
    __device__ __forceinline__ void seed(uint64_t sd) {
        uint32_t v = (uint32_t)sd;
        x[0] = v;
        for (uint32_t i = 1; i < MT_N; ++i) {
            v = mt_init_step(v, i);
            x[i] = v;
        }
        p = MT_N;
        at = 0;
    }
    __device__ __forceinline__ void refill() {
        for (uint32_t k = 0; k < MT_N; ++k) {
            const uint32_t k1 = k + 1 == MT_N ? 0 : k + 1, km = k + MT_M >= MT_N ? k + MT_M - MT_N : k + MT_M;
            x[k] = mt_twist(x[k], x[k1], x[km]);
        }
        p = 0;
    }
    __device__ __forceinline__ void position(uint64_t abs) { /* prng::discard (random.hh:63) */
        uint64_t n = abs - at;
        at = abs;
        while (n) {
            if (p >= MT_N) refill();
            const uint64_t take = n < (uint64_t)(MT_N - p) ? n : (uint64_t)(MT_N - p);
            p += (uint32_t)take;
            n -= take;
        }
    }
    __device__ __forceinline__ uint32_t next() {
        if (p >= MT_N) refill();
        ++at;
        return mt_temper(x[p++]);
    }
    __device__ __forceinline__ uint64_t u64() {
        const uint32_t hi = next(), lo = next();
        return ((uint64_t)hi << 32) | lo;
    }
    __device__ __forceinline__ double uniform() { /* random.hh:23: the first draw is the LOW word */
        const uint32_t a = next(), b = next();
        return prng_uniform_from_words(a, b);
    }
};

/* state <- A^k state for the xor_shift step matrix A, from the columns of A^(2^j). */
__device__ __forceinline__ uint64_t xs_jump(uint64_t s, uint64_t k, const uint64_t *pw) {
    for (int j = 0; k; ++j, k >>= 1) {
        if (!(k & 1)) continue;
        const uint64_t *col = pw + 64 * j;
        uint64_t y = 0;
        for (int b = 0; b < 64; ++b)
            if ((s >> b) & 1) y ^= col[b];
        s = y;
    }
    return s;
}

/* GrowLib::alternating_constant_actions (grow-lib.hh:16-24) applied as actions_func to rows, then to cols
 * (random-tree.hh:105-106): the second call already sees the new rows. */
__device__ __forceinline__ void th_grow(uint32_t growth, uint32_t &rows, uint32_t &cols) {
    if (growth == PCU_GROW_ALTERNATING_CONSTANT) {
        rows = rows == 1 ? max(rows, cols) : 1;
        cols = cols == 1 ? max(rows, cols) : 1;
    }
}

template <typename Eng>
__device__ __forceinline__ void th_rollout(const ThreadParams &P, uint64_t i, Eng &S, Eng &M, unsigned long long *acc) {
    uint32_t rows = P.rows, cols = P.cols;
    const uint32_t T = P.T;
    uint32_t d = P.depth ? P.depth[i] : P.depth_u;
    const uint64_t soff = P.step_off ? P.step_off[i] : i * (uint64_t)P.depth_u;
    int bias = 0;
    uint64_t cur = 0; /* draw index (in engine units) where the current node's chance strategies start */
    double pu = 0.0;
    if (P.kind == KIND_TREE) {
        bias = P.bias ? P.bias[i] : P.bias_u;
        const uint64_t sd = P.state_seed ? P.state_seed[i] : P.state_seed_u;
        if constexpr (Eng::UNITS == 2) {
            S.seed(sd);
            cur = (uint64_t)(P.state_pos ? P.state_pos[i] : P.state_pos_u) - 2ull * rows * cols * T;
        } else { /* the state's engine as it stands: behind the rows cols T draws of its latest strategies */
            uint64_t s = sd;
            const uint32_t gen = rows * cols * T;
            for (uint32_t k = 0; k < gen; ++k) s = xs_unstep(s);
            S.s = s;
            S.at = 0;
            cur = 0;
        }
        if (T > 1) { /* prng temp{transition_seed}; p = temp.uniform() (random-tree.hh:118-121): once per rollout */
            const uint64_t ts = P.tseed ? P.tseed[i] : P.tseed_u;
            if constexpr (Eng::UNITS == 2) {
                uint32_t v = (uint32_t)ts, t1 = 0, t2 = 0, t397 = 0;
                const uint32_t t0 = v;
                for (uint32_t k = 1; k <= 398; ++k) {
                    v = mt_init_step(v, k);
                    if (k == 1) t1 = v;
                    if (k == 2) t2 = v;
                    if (k == 397) t397 = v;
                }
                pu = prng_uniform_from_words(mt_temper(mt_twist(t0, t1, t397)), mt_temper(mt_twist(t1, t2, v)));
            } else {
                Eng tmp;
                tmp.s = ts;
                tmp.at = 0;
                pu = tmp.uniform();
            }
        }
    }
    uint32_t h = PCU_HASH_INIT;
    uint8_t *tr = P.traj ? P.traj + (size_t)i * P.traj_stride * 3 : nullptr;
    const uint32_t steps = d;
    for (uint32_t s = 0; s < steps; ++s) {
        uint32_t r, c;
        if (P.model_src == TH_MODEL_WORDS) { /* the step's two uniform_64 draws, high word first (random.hh:26,28) */
            uint4 w = reinterpret_cast<const uint4 *>(P.words)[soff + s];
            if (P.words_raw) {
                w.x = mt_temper(w.x);
                w.y = mt_temper(w.y);
                w.z = mt_temper(w.z);
                w.w = mt_temper(w.w);
            }
            r = (uint32_t)((((uint64_t)w.x << 32) | w.y) % rows);
            c = (uint32_t)((((uint64_t)w.z << 32) | w.w) % cols);
        } else {
            r = (uint32_t)(M.u64() % rows); /* monte-carlo-model.hh:89-90 */
            c = (uint32_t)(M.u64() % cols);
        }
        uint32_t ch = 0;
        const uint32_t joint = r * cols + c;
        if (P.kind == KIND_TREE) {
            const uint32_t st = joint * T;
            if (T > 1) { /* the chosen cell's strategies (random-tree.hh:154-199), then apply_actions(row, col) :115-135 */
                S.position(cur + (uint64_t)Eng::UNITS * st);
                int nk[8], sum = 0;
                for (uint32_t k = 0; k < T; ++k) {
                    nk[k] = (int)(S.u64() % 10) + 1;
                    if (nk[k] * P.thr_q < P.thr_p * 10) nk[k] = 0;
                    sum += nk[k];
                }
                if (sum == 0) {
                    nk[0] = 1;
                    sum = 1;
                }
                double pp = pu;
                for (uint32_t k = 0; k < T; ++k) {
                    pp = __dsub_rn(pp, P.fq[(uint32_t)nk[k] * P.fq_w + (uint32_t)sum]);
                    if (pp <= 0) {
                        ch = k;
                        break;
                    }
                }
            }
            /* apply_actions(row, col, chance) :87-109: discard(idx), one bias draw, new strategies */
            const uint64_t gen = (uint64_t)Eng::UNITS * rows * cols * T;
            S.position(cur + gen + st + ch);
            bias += (int)(S.u64() % 3) - 1;
            cur = cur + gen + st + ch + Eng::UNITS;
        }
        h = hash_step(h, joint, ch);
        if (tr && s < P.traj_stride) {
            tr[3 * s + 0] = (uint8_t)r;
            tr[3 * s + 1] = (uint8_t)c;
            tr[3 * s + 2] = (uint8_t)ch;
        }
        if (P.kind == KIND_TREE && s + 1 < steps) th_grow(P.growth, rows, cols);
    }
    uint32_t pay = 0;
    if (P.kind == KIND_TREE) pay = (uint32_t)((bias > 0) - (bias < 0) + 1);
    if (tr)
        for (uint32_t s = steps; s < P.traj_stride; ++s) tr[3 * s] = tr[3 * s + 1] = tr[3 * s + 2] = 0;
    if (P.payoff) P.payoff[i] = (uint8_t)pay;
    if (P.traj_hash) P.traj_hash[i] = h;
    if (P.kind == KIND_TREE) {
        acc[0] = pay == 2;
        acc[1] = pay == 1;
        acc[2] = pay == 0;
    } else {
        acc[1] = 1;
    }
    acc[3] = steps;
    acc[4] = h;
}

template <int ENGINE>
__global__ void __launch_bounds__(128) k_thread(const ThreadParams P) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long acc[5] = {0, 0, 0, 0, 0};
    if (i < P.n) {
        if (ENGINE == PCU_ENG_XORSHIFT) {
            EngXs S, M;
            S.s = 0;
            S.at = 0;
            M.at = 0;
            if (P.model_src == TH_MODEL_SEEDS) {
                M.s = P.model_seeds[i];
            } else { /* ONE model engine consumed in index order: two draws per step (monte-carlo-model.hh:58-67) */
                const uint64_t soff = P.step_off ? P.step_off[i] : i * (uint64_t)P.depth_u;
                M.s = xs_jump(P.xs_m0, 2 * soff, P.xs_pow);
            }
            th_rollout<EngXs>(P, i, S, M, acc);
        } else {
            EngMt S, M;
            S.p = MT_N;
            S.at = 0;
            if (P.model_src == TH_MODEL_SEEDS) M.seed(P.model_seeds[i]);
            th_rollout<EngMt>(P, i, S, M, acc);
        }
    }
    if (P.sums) {
#pragma unroll
        for (int k = 0; k < 5; ++k) {
            const unsigned long long v = warp_sum_u64(acc[k]);
            if ((threadIdx.x & 31) == 0 && v) atomicAdd(P.sums + k, v);
        }
    }
}
