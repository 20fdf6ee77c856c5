// TEST INFRASTRUCTURE ONLY — the reference itself, compiled as the parity oracle.
//
// This single translation unit includes baskuit/pinyon's own headers, unmodified,
// from /root/reference/src (plus oracle/shim/gmpxx.h, a parse-only GMP stand-in)
// and wraps the reference's MonteCarloModel (src/model/monte-carlo-model.hh:25-98)
// over RandomTree (src/state/random-tree.hh:10-201) and MoldState
// (src/state/test-states.hh:9-68) behind a plain C ABI so that tests/ and
// bench.py's cpu_baseline / --impl reference legs can drive it through ctypes.
// It is built into oracle/_ref/ (git-ignored) by oracle/Makefile and is never
// linked into, loaded by or called from the product library.
//
// Single TU on purpose: src/libpinyon/math.hh:14,21,28 defines non-inline
// functions, so two TUs including it would not link.
//
// Trajectories are observed without touching the reference: TracedTree/TracedMold
// derive from the reference State classes and shadow apply_actions(row, col) with
// a recorder that forwards to the base implementation; the rollout loop that runs
// is the reference's own protected MonteCarloModel::Model::rollout
// (monte-carlo-model.hh:85-96).

// The reference's headers rely on their includer for some standard headers
// (off-policy.hh uses std::chrono, tree-bandit-flat.hh <cstring>/<unordered_map>).
#include <array>
#include <atomic>
#include <chrono>
#include <cstdint>
#include <cstring>
#include <memory>
#include <thread>
#include <unordered_map>
#include <vector>

#include <model/monte-carlo-model.hh>
#include <state/random-tree.hh>
#include <state/test-states.hh>
#include <libpinyon/grow-lib.hh>

#include <algorithm/tree-bandit/bandit/exp3.hh>
#include <algorithm/tree-bandit/tree/off-policy.hh>
#include <algorithm/tree-bandit/tree/tree-bandit.hh>
#include <tree/tree.hh>

namespace {

struct TraceSink {
    uint8_t *traj = nullptr;  // 3 bytes per step: row, col, chance
    uint32_t steps = 0;
};
thread_local TraceSink *g_sink = nullptr;

template <typename TL>
struct TracedTreeT : RandomTree<TL> {
    using Base = typename RandomTree<TL>::State;
    class State : public Base {
       public:
        using Base::Base;
        State(const Base &b) : Base{b} {}
        using Base::apply_actions;
        void apply_actions(int row_action, int col_action) {
            Base::apply_actions(row_action, col_action);
            if (g_sink) {
                if (g_sink->traj) {
                    uint8_t *t = g_sink->traj + 3 * static_cast<size_t>(g_sink->steps);
                    t[0] = static_cast<uint8_t>(row_action);
                    t[1] = static_cast<uint8_t>(col_action);
                    t[2] = static_cast<uint8_t>(this->obs);
                }
                ++g_sink->steps;
            }
        }
    };
};

template <typename TL>
struct TracedMoldT : MoldState<TL> {
    using Base = typename MoldState<TL>::State;
    class State : public Base {
       public:
        using Base::Base;
        using Base::apply_actions;
        void apply_actions(int row_action, int col_action) {
            Base::apply_actions(row_action, col_action);
            if (g_sink) {
                if (g_sink->traj) {
                    uint8_t *t = g_sink->traj + 3 * static_cast<size_t>(g_sink->steps);
                    t[0] = static_cast<uint8_t>(row_action);
                    t[1] = static_cast<uint8_t>(col_action);
                    t[2] = 0;
                }
                ++g_sink->steps;
            }
        }
    };
};

using TracedTree = TracedTreeT<RandomTreeFloatTypes>;
using TracedMold = TracedMoldT<SimpleTypes>;

// ---- the xor_shift PRNG (src/types/random.hh:66-121) as a type list.  The class as shipped fails IsPRNG
// (types.hh:165-190 wants random_seed()), so no reference type list can name it; the one missing member is added
// here, exactly as a user would have to, and nothing else is touched.
struct xor_shift_prng : xor_shift {
    using xor_shift::xor_shift;
    uint64_t random_seed() { return uniform_64(); }
};
using XsTreeTypes = DefaultTypes<double, int, int, double, ConstantSum<1, 1>::Value, std::vector, Matrix, std::mutex,
                                 uint64_t, xor_shift_prng>;
using XsMoldTypes = DefaultTypes<double, int, int, double, PairReal, std::vector, Matrix, std::mutex, uint64_t, xor_shift_prng>;
static_assert(IsTypeList<XsTreeTypes>);

// xor_shift keeps its 64-bit state private: read it through the explicit-instantiation access rule.
template <typename Tag, typename Tag::type Member>
struct PrivateAccess {
    friend typename Tag::type harness_get(Tag) { return Member; }
};
struct XsStateTag {
    using type = uint64_t xor_shift::*;
    friend type harness_get(XsStateTag);
};
template struct PrivateAccess<XsStateTag, &xor_shift::state>;
inline uint64_t xs_state_of(const xor_shift &d) { return d.*harness_get(XsStateTag{}); }

int g_growth = 0;  // 0: default growth functions; 1: GrowLib::alternating_constant_actions as actions_func

using TreeMC = MonteCarloModel<TracedTree>;
using MoldMC = MonteCarloModel<TracedMold>;
using PlainTreeMC = MonteCarloModel<RandomTree<>>;
using PlainMoldMC = MonteCarloModel<MoldState<>>;

static_assert(IsSingleModelTypes<PlainTreeMC>);
static_assert(IsBatchModelTypes<PlainTreeMC>);
static_assert(IsSingleModelTypes<PlainMoldMC>);

inline uint8_t payoff_code(double row_value) { return static_cast<uint8_t>(row_value * 2.0); }

template <typename TL>
uint64_t tree_rollouts(uint64_t n, const uint64_t *state_seeds, uint64_t state_seed_u, const uint32_t *state_pre_discards,
                       uint32_t state_pre_discard_u, const uint64_t *transition_seeds, uint64_t transition_seed_u,
                       const uint32_t *depths, uint32_t depth_u, const int32_t *payoff_biases, int32_t payoff_bias_u,
                       uint32_t rows, uint32_t cols, uint32_t transitions, int thr_p, int thr_q, const uint64_t *model_seeds,
                       uint64_t model_seed_u, uint64_t model_pre_discard, uint8_t *payoff_x2, uint8_t *traj,
                       uint32_t traj_depth_max, uint64_t *state_after_ctor) {
    using Traced = TracedTreeT<TL>;
    using MC = MonteCarloModel<Traced>;
    using PRNG = typename TL::PRNG;
    using BaseState = typename RandomTree<TL>::State;
    typename MC::Model shared_model{PRNG{model_seed_u}};
    shared_model.device.discard(model_pre_discard);
    uint64_t total_steps = 0;
    TraceSink sink;
    g_sink = &sink;

    const bool uniform_root = !state_seeds && !state_pre_discards && !depths;
    std::unique_ptr<typename Traced::State> root_cache;

    for (uint64_t i = 0; i < n; ++i) {
        const uint64_t sseed = state_seeds ? state_seeds[i] : state_seed_u;
        const uint32_t spre = state_pre_discards ? state_pre_discards[i] : state_pre_discard_u;
        const uint32_t depth = depths ? depths[i] : depth_u;
        if (!uniform_root || !root_cache) {
            PRNG d{sseed};
            d.discard(spre);
            if (g_growth == 1)  // GrowLib::alternating_constant_actions as actions_func (random-tree.hh:43-57)
                root_cache = std::make_unique<typename Traced::State>(BaseState{
                    d, depth, rows, cols, transitions, Rational<>{thr_p, thr_q}, &BaseState::depth_bound_default,
                    &GrowLib::alternating_constant_actions<BaseState>, &BaseState::payoff_bias_default});
            else
                root_cache = std::make_unique<typename Traced::State>(d, depth, rows, cols, transitions, Rational<>{thr_p, thr_q});
        }
        if constexpr (std::is_base_of_v<xor_shift, PRNG>)
            if (state_after_ctor) state_after_ctor[i] = xs_state_of(root_cache->device);
        typename Traced::State state{*root_cache};
        state.randomize_transition(transition_seeds ? transition_seeds[i] : transition_seed_u);
        state.payoff_bias = payoff_biases ? payoff_biases[i] : payoff_bias_u;

        sink.traj = traj ? traj + 3 * static_cast<size_t>(traj_depth_max) * i : nullptr;
        sink.steps = 0;
        typename MC::ModelOutput out;
        if (model_seeds) {
            typename MC::Model model{PRNG{model_seeds[i]}};
            model.inference(std::move(state), out);
        } else {
            shared_model.inference(std::move(state), out);
        }
        total_steps += sink.steps;
        if (payoff_x2) payoff_x2[i] = payoff_code(out.value.get_row_value());
    }
    g_sink = nullptr;
    return total_steps;
}

template <typename TL>
uint64_t mold_rollouts(uint64_t n, uint32_t n_actions, const uint32_t *depths, uint32_t depth_u, const uint64_t *model_seeds,
                       uint64_t model_seed_u, uint64_t model_pre_discard, double *payoff_row, double *payoff_col, uint8_t *traj,
                       uint32_t traj_depth_max) {
    using Traced = TracedMoldT<TL>;
    using MC = MonteCarloModel<Traced>;
    using PRNG = typename TL::PRNG;
    typename MC::Model shared_model{PRNG{model_seed_u}};
    shared_model.device.discard(model_pre_discard);
    uint64_t total_steps = 0;
    TraceSink sink;
    g_sink = &sink;
    for (uint64_t i = 0; i < n; ++i) {
        const uint32_t depth = depths ? depths[i] : depth_u;
        typename Traced::State state{n_actions, depth};
        sink.traj = traj ? traj + 3 * static_cast<size_t>(traj_depth_max) * i : nullptr;
        sink.steps = 0;
        typename MC::ModelOutput out;
        if (model_seeds) {
            typename MC::Model model{PRNG{model_seeds[i]}};
            model.inference(std::move(state), out);
        } else {
            shared_model.inference(std::move(state), out);
        }
        total_steps += sink.steps;
        if (payoff_row) payoff_row[i] = out.value.get_row_value();
        if (payoff_col) payoff_col[i] = out.value.get_col_value();
    }
    g_sink = nullptr;
    return total_steps;
}

}  // namespace

extern "C" {

// prng primitives (src/types/random.hh:8-64), for pinning the restatement.
void ref_prng_uniform_64(uint64_t seed, uint64_t discard, uint64_t n, uint64_t *out) {
    prng d{seed};
    d.discard(discard);
    for (uint64_t i = 0; i < n; ++i) out[i] = d.uniform_64();
}
void ref_prng_uniform(uint64_t seed, uint64_t discard, uint64_t n, double *out) {
    prng d{seed};
    d.discard(discard);
    for (uint64_t i = 0; i < n; ++i) out[i] = d.uniform();
}
void ref_prng_random_int(uint64_t seed, uint64_t discard, int modulus, uint64_t n, int32_t *out) {
    prng d{seed};
    d.discard(discard);
    for (uint64_t i = 0; i < n; ++i) out[i] = d.random_int(modulus);
}
uint64_t ref_sizeof_prng() { return sizeof(prng); }
uint64_t ref_sizeof_random_tree_state() { return sizeof(RandomTree<>::State); }

// Root chance strategies as the reference's constructor generates them
// (random-tree.hh:31-41,154-199).  out must hold rows*cols*transitions doubles.
void ref_random_tree_root_strategies(uint64_t state_seed, uint64_t state_pre_discard, uint32_t depth,
                                     uint32_t rows, uint32_t cols, uint32_t transitions, int thr_p, int thr_q,
                                     double *out) {
    prng d{state_seed};
    d.discard(state_pre_discard);
    RandomTree<>::State root{d, depth, rows, cols, transitions, Rational<>{thr_p, thr_q}};
    for (size_t i = 0; i < root.chance_strategies.size(); ++i) out[i] = root.chance_strategies[i];
}

// RandomTree rollouts through the reference MonteCarloModel.
//   per-rollout arrays may be null -> the *_u scalar is used for every rollout.
//   model_seeds == null: "mode S" — ONE Model{prng{model_seed_u}} (after
//     discard(model_pre_discard)) consumes the rollouts in index order, i.e. the
//     reference's batched inference (monte-carlo-model.hh:58-67).
//   model_seeds != null: "mode P" — rollout i runs on a fresh Model{prng{model_seeds[i]}}.
//   traj (nullable): n * depth_max * 3 bytes, row/col/chance per step, zero padded.
// Returns total steps executed.
uint64_t ref_random_tree_rollouts(uint64_t n, const uint64_t *state_seeds, uint64_t state_seed_u,
                                  const uint32_t *state_pre_discards, uint32_t state_pre_discard_u,
                                  const uint64_t *transition_seeds, uint64_t transition_seed_u,
                                  const uint32_t *depths, uint32_t depth_u, const int32_t *payoff_biases,
                                  int32_t payoff_bias_u, uint32_t rows, uint32_t cols, uint32_t transitions,
                                  int thr_p, int thr_q, const uint64_t *model_seeds, uint64_t model_seed_u,
                                  uint64_t model_pre_discard, uint8_t *payoff_x2, uint8_t *traj,
                                  uint32_t traj_depth_max) {
    return tree_rollouts<RandomTreeFloatTypes>(n, state_seeds, state_seed_u, state_pre_discards, state_pre_discard_u,
                                               transition_seeds, transition_seed_u, depths, depth_u, payoff_biases,
                                               payoff_bias_u, rows, cols, transitions, thr_p, thr_q, model_seeds,
                                               model_seed_u, model_pre_discard, payoff_x2, traj, traj_depth_max, nullptr);
}

// The same over the xor_shift type list.  Seeds are engine STATES (xor_shift(seed): state = seed, random.hh:72);
// state_after_ctor (nullable) receives each root's engine state as the constructor left it — what the CUDA library's
// batch carries for an xor_shift state.
uint64_t ref_xs_random_tree_rollouts(uint64_t n, const uint64_t *state_seeds, uint64_t state_seed_u,
                                     const uint32_t *state_pre_discards, uint32_t state_pre_discard_u,
                                     const uint64_t *transition_seeds, uint64_t transition_seed_u,
                                     const uint32_t *depths, uint32_t depth_u, const int32_t *payoff_biases,
                                     int32_t payoff_bias_u, uint32_t rows, uint32_t cols, uint32_t transitions,
                                     int thr_p, int thr_q, const uint64_t *model_seeds, uint64_t model_seed_u,
                                     uint64_t model_pre_discard, uint8_t *payoff_x2, uint8_t *traj,
                                     uint32_t traj_depth_max, uint64_t *state_after_ctor) {
    return tree_rollouts<XsTreeTypes>(n, state_seeds, state_seed_u, state_pre_discards, state_pre_discard_u,
                                      transition_seeds, transition_seed_u, depths, depth_u, payoff_biases, payoff_bias_u,
                                      rows, cols, transitions, thr_p, thr_q, model_seeds, model_seed_u, model_pre_discard,
                                      payoff_x2, traj, traj_depth_max, state_after_ctor);
}

void ref_set_growth(int growth) { g_growth = growth; }

// MoldState rollouts (test-states.hh:13-67; benchmark/mold-state.cc:3-12).
// payoff_row/payoff_col receive the (never-set, hence 0,0) PairReal payoff.
uint64_t ref_mold_rollouts(uint64_t n, uint32_t n_actions, const uint32_t *depths, uint32_t depth_u,
                           const uint64_t *model_seeds, uint64_t model_seed_u, uint64_t model_pre_discard,
                           double *payoff_row, double *payoff_col, uint8_t *traj, uint32_t traj_depth_max) {
    return mold_rollouts<SimpleTypes>(n, n_actions, depths, depth_u, model_seeds, model_seed_u, model_pre_discard, payoff_row,
                                      payoff_col, traj, traj_depth_max);
}
uint64_t ref_xs_mold_rollouts(uint64_t n, uint32_t n_actions, const uint32_t *depths, uint32_t depth_u,
                              const uint64_t *model_seeds, uint64_t model_seed_u, uint64_t model_pre_discard,
                              double *payoff_row, double *payoff_col, uint8_t *traj, uint32_t traj_depth_max) {
    return mold_rollouts<XsMoldTypes>(n, n_actions, depths, depth_u, model_seeds, model_seed_u, model_pre_discard, payoff_row,
                                      payoff_col, traj, traj_depth_max);
}

// CPU baseline timing: the reference's untraced MonteCarloModel, `threads`
// std::threads, each owning a Model copy and a contiguous slice of the batch
// (the reference has no threaded inference of its own; this mirrors the
// per-thread model copies of multithreaded.hh:112-113).  Every rollout starts
// from a copy of one root.  kind 0 = RandomTree, 1 = MoldState.
// Returns seconds; wdl[3] receives win/draw/loss counts (RandomTree).
double ref_time_rollouts(int kind, uint64_t n, uint32_t threads, uint64_t state_seed, uint32_t depth,
                         uint32_t rows, uint32_t cols, uint32_t transitions, uint64_t model_seed, uint64_t *wdl,
                         uint64_t *steps_out) {
    if (threads == 0) threads = 1;
    std::vector<std::thread> pool;
    std::vector<std::array<uint64_t, 4>> partial(threads, std::array<uint64_t, 4>{0, 0, 0, 0});
    std::atomic<int> ready{0};
    std::atomic<bool> go{false};
    const auto worker = [&](uint32_t tid) {
        const uint64_t lo = n * tid / threads, hi = n * (tid + 1) / threads;
        std::array<uint64_t, 4> acc{0, 0, 0, 0};
        if (kind == 0) {
            const RandomTree<>::State root{prng{state_seed}, depth, rows, cols, transitions};
            PlainTreeMC::Model model{prng{model_seed + tid}};
            ++ready;
            while (!go.load()) std::this_thread::yield();
            for (uint64_t i = lo; i < hi; ++i) {
                PlainTreeMC::ModelOutput out;
                model.inference(RandomTree<>::State{root}, out);
                acc[2 - payoff_code(out.value.get_row_value())] += 1;
            }
            acc[3] = (hi - lo) * depth;
        } else {
            PlainMoldMC::Model model{prng{model_seed + tid}};
            ++ready;
            while (!go.load()) std::this_thread::yield();
            double sink_v = 0;
            for (uint64_t i = lo; i < hi; ++i) {
                PlainMoldMC::ModelOutput out;
                model.inference(MoldState<>::State{rows, depth}, out);
                sink_v += out.value.get_row_value();
            }
            acc[1] = static_cast<uint64_t>(sink_v);
            acc[3] = (hi - lo) * depth;
        }
        partial[tid] = acc;
    };
    for (uint32_t t = 0; t < threads; ++t) pool.emplace_back(worker, t);
    while (ready.load() < static_cast<int>(threads)) std::this_thread::yield();
    const auto t0 = std::chrono::steady_clock::now();
    go.store(true);
    for (auto &th : pool) th.join();
    const auto t1 = std::chrono::steady_clock::now();
    uint64_t tot[4] = {0, 0, 0, 0};
    for (auto &p : partial)
        for (int k = 0; k < 4; ++k) tot[k] += p[k];
    if (wdl) {
        wdl[0] = tot[0];
        wdl[1] = tot[1];
        wdl[2] = tot[2];
    }
    if (steps_out) *steps_out = tot[3];
    return std::chrono::duration<double>(t1 - t0).count();
}

// TreeBandit<Exp3<MonteCarloModel<RandomTree<>>>> root statistics (SURVEY.md
// Appendix B4; tree-bandit.hh:58-76, exp3.hh:9-16).  stats_out:
// [0]=matrix nodes, [1]=root visits, [2..2+rows)=row_visits, then col_visits;
// value_total_row receives root value_total row value.
void ref_tree_bandit_exp3(uint64_t state_seed, uint32_t depth, uint32_t rows, uint32_t cols, uint32_t transitions,
                          uint64_t device_seed, uint64_t model_seed, uint64_t iterations, int64_t *stats_out,
                          double *value_total_row) {
    using Types = TreeBandit<Exp3<PlainTreeMC>>;
    Types::PRNG device{device_seed};
    Types::State state{prng{state_seed}, depth, rows, cols, transitions};
    Types::Model model{prng{model_seed}};
    Types::MatrixNode root{};
    Types::Search search{};
    search.run_for_iterations(iterations, device, state, model, root);
    stats_out[0] = static_cast<int64_t>(root.count_matrix_nodes());
    stats_out[1] = root.stats.visits;
    for (uint32_t r = 0; r < rows; ++r) stats_out[2 + r] = root.stats.row_visits[r];
    for (uint32_t c = 0; c < cols; ++c) stats_out[2 + rows + c] = root.stats.col_visits[c];
    *value_total_row = static_cast<double>(root.stats.value_total.get_row_value());
}

}  // extern "C"
