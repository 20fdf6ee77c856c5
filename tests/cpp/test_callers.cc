// The remaining callers north_star names — MatrixUCB, FullTraversal — and the reference's tests/flat.cc equivalence,
// over PinyonCuda::MonteCarloModel, each against the same run over the reference's own MonteCarloModel.
//
// Compiled with g++ -std=c++23 against the UNMODIFIED reference headers.  Two things the reference checkout lacks are
// stood in for by test infrastructure:
//   * extern/lrslib (absent submodule): tests/cpp/lrs_shim/ — lib.h's entry points over a support-enumeration solver;
//   * MatrixUCB has no get_refined_value, which IsBanditAlgorithmTypes demands (algorithm.hh:28-30 vs
//     matrix-ucb.hh:41-157): MatrixUCBFixed below derives the one missing member.
// Both runs of every comparison use the same stand-ins, so equality of their results is a statement about the models.
//
//   test_callers --cpu   reference model only (the stand-ins work; tree == flat for the reference)
//   test_callers --gpu   reference vs GPU model, on every visible GPU (Device::shared_all)
#include <cstdio>
#include <cstring>
#include <memory>
#include <unordered_map>
#include <vector>

#include <model/monte-carlo-model.hh>
#include <state/random-tree.hh>
#include <state/test-states.hh>

#include <algorithm/solver/full-traversal.hh>
#include <algorithm/tree-bandit/bandit/exp3.hh>
#include <algorithm/tree-bandit/bandit/matrix-ucb.hh>
#include <algorithm/tree-bandit/tree/tree-bandit.hh>
#include <algorithm/tree-bandit/tree/tree-bandit-flat.hh>
#include <tree/tree.hh>

#include <pinyon_cuda/monte_carlo_model.hh>

using RefTree = MonteCarloModel<RandomTree<>>;
using GpuTree = PinyonCuda::MonteCarloModel<RandomTree<>>;

// MatrixUCB + the member IsBanditAlgorithmTypes requires (algorithm.hh:28-30): the value of the refined (Nash)
// strategies on the empirical value matrix, the counterpart of Exp3::get_refined_value (exp3.hh).
template <IsValueModelTypes Types>
struct MatrixUCBFixed : MatrixUCB<Types> {
    using Base = MatrixUCB<Types>;
    class BanditAlgorithm : public Base::BanditAlgorithm {
       public:
        using Base::BanditAlgorithm::BanditAlgorithm;
        void get_refined_value(const typename Base::MatrixStats &stats, typename Types::Value &value) const {
            typename Types::VectorReal row_strategy, col_strategy;
            const auto matrix = this->get_ev_matrix(stats);
            value = LRSNash::solve(matrix, row_strategy, col_strategy);
        }
    };
};

static_assert(IsBanditAlgorithmTypes<MatrixUCBFixed<GpuTree>>);
static_assert(IsSearchTypes<TreeBandit<MatrixUCBFixed<GpuTree>>>);
static_assert(IsSingleModelTypes<GpuTree> && IsChanceStateTypes<GpuTree>);  // FullTraversal's constraints (:20-23)
static_assert(IsSearchTypes<TreeBandit<Exp3<GpuTree>>>);

static int g_failures = 0;
#define CHECK(cond)                                                          \
    do {                                                                     \
        if (!(cond)) {                                                       \
            std::printf("FAIL %s:%d: %s\n", __FILE__, __LINE__, #cond);      \
            ++g_failures;                                                    \
        }                                                                    \
    } while (0)

// ---- FullTraversal: exhaustive solve down to max_depth, model.inference at the cut-off (full-traversal.hh:112-119),
// LRSNash::solve at every internal node (:187)
template <typename Types>
static std::vector<double> full_traversal(uint64_t seed, size_t depth, size_t actions, size_t T, size_t cutoff) {
    using FT = FullTraversal<Types>;
    typename Types::State state{prng{seed}, depth, actions, actions, T};
    typename Types::Model model{prng{1}};
    typename FT::MatrixNode root{};
    typename FT::Search search{};
    search.run(cutoff, state, model, root);
    std::vector<double> out{double(root.stats.payoff.get_row_value()), double(root.stats.matrix_node_count)};
    for (auto v : root.stats.row_solution) out.push_back(double(v));
    for (auto v : root.stats.col_solution) out.push_back(double(v));
    out.push_back(double(model.device.uniform_64() >> 11));  // the model engine ends in the same place
    return out;
}

// ---- TreeBandit<MatrixUCB>: select solves the UCB matrix (matrix-ucb.hh:95-108), leaves go to model.inference
template <typename Types>
static std::vector<double> matrix_ucb_root(size_t iterations) {
    using S = TreeBandit<MatrixUCBFixed<Types>>;
    typename S::PRNG device{0};
    typename S::State state{prng{0}, 6, 3, 3, 1};
    typename S::Model model{prng{0}};
    typename S::MatrixNode root{};
    typename S::Search search{};
    search.run_for_iterations(iterations, device, state, model, root);
    std::vector<double> out{double(root.count_matrix_nodes()), double(root.stats.total_visits)};
    for (int e = 0; e < 9; ++e) {
        out.push_back(double(root.stats.data_matrix[e].visits));
        out.push_back(double(root.stats.data_matrix[e].value.get_row_value()));
    }
    typename S::Value refined;
    search.get_refined_value(root.stats, refined);
    out.push_back(double(refined.get_row_value()));
    return out;
}

// ---- tests/flat.cc:3-29: TreeBandit and TreeBanditFlat give identical root statistics
using FlatOptions = SearchOptions<void, void, void, void, 1 << 12, 1 << 5>;
template <typename Types>
static void flat_equivalence(uint64_t seed, size_t T, Rational<> thr, size_t iterations, typename Exp3<Types>::MatrixStats &tree_out,
                             typename Exp3<Types>::MatrixStats &flat_out) {
    using B = Exp3<Types>;
    const typename B::State state{prng{seed}, 20, 3, 3, T, thr};
    const typename B::PRNG device{0};
    const typename B::Model model{prng{0}};
    {
        auto device_ = device;
        auto model_ = model;
        typename TreeBandit<B>::MatrixNode node{};
        typename TreeBandit<B>::Search search{};
        search.run_for_iterations(iterations, device_, state, model_, node);
        tree_out = node.stats;
    }
    {
        auto device_ = device;
        auto model_ = model;
        auto search = std::make_unique<typename TreeBanditFlat<B, FlatOptions>::Search>();
        search->run_for_iterations(iterations, device_, state, model_);
        flat_out = search->matrix_data[0].stats;
    }
}

int main(int argc, char **argv) {
    const bool gpu = argc > 1 && !std::strcmp(argv[1], "--gpu");
    try {
        {   // the stand-ins do their job on the reference model alone
            const auto a = full_traversal<RefTree>(3, 4, 2, 1, 2);
            CHECK(a[0] >= 0 && a[0] <= 1 && a[1] > 1);
            const auto b = matrix_ucb_root<RefTree>(300);
            CHECK(b[1] == 299 || b[1] == 300);
            Exp3<RefTree>::MatrixStats t, f;
            flat_equivalence<RefTree>(7, 1, Rational<>{0, 1}, 2000, t, f);
            std::printf("reference: FullTraversal value %.6f over %.0f nodes; MatrixUCB visits %.0f; tree == flat: %d\n", a[0],
                        a[1], b[1], int(t == f));
        }
        if (gpu) {
            for (const auto &[seed, depth, actions, T, cutoff] :
                 {std::tuple<uint64_t, size_t, size_t, size_t, size_t>{3, 5, 2, 1, 3}, {0, 4, 3, 1, 2}, {11, 4, 2, 2, 2}}) {
                const auto a = full_traversal<RefTree>(seed, depth, actions, T, cutoff);
                const auto b = full_traversal<GpuTree>(seed, depth, actions, T, cutoff);
                CHECK(a == b);
                std::printf("FullTraversal seed %llu depth %zu %zux%zux%zu cut-off %zu: value %.6f, %.0f nodes (ref) vs %.6f, %.0f (gpu)\n",
                            (unsigned long long)seed, depth, actions, actions, T, cutoff, a[0], a[1], b[0], b[1]);
            }
            {
                const auto a = matrix_ucb_root<RefTree>(1500);
                const auto b = matrix_ucb_root<GpuTree>(1500);
                CHECK(a == b);
                std::printf("TreeBandit<MatrixUCB>: nodes %.0f visits %.0f refined value %.6f (ref) vs %.0f %.0f %.6f (gpu)\n", a[0],
                            a[1], a.back(), b[0], b[1], b.back());
            }
            int cases = 0, tree_is_flat = 0;
            for (uint64_t seed : {1ull, 2ull, 3ull})
                for (size_t T : {size_t(1), size_t(2)})
                    for (Rational<> thr : {Rational<>{0, 1}, Rational<>{1, 2}}) {
                        Exp3<RefTree>::MatrixStats rt, rf;
                        Exp3<GpuTree>::MatrixStats gt, gf;
                        flat_equivalence<RefTree>(seed, T, thr, 3000, rt, rf);
                        flat_equivalence<GpuTree>(seed, T, thr, 3000, gt, gf);
                        auto same = [](const auto &x, const auto &y) {
                            return x.row_gains == y.row_gains && x.col_gains == y.col_gains && x.row_visits == y.row_visits &&
                                   x.col_visits == y.col_visits && x.visits == y.visits &&
                                   x.value_total.row_value == y.value_total.row_value &&
                                   x.value_total.col_value == y.value_total.col_value;
                        };
                        CHECK(same(rt, gt));       // the tree search is model-independent
                        CHECK(same(rf, gf));       // so is the flat search
                        CHECK(gt == gf);           // tests/flat.cc:28 over the GPU model
                        ++cases;
                        tree_is_flat += int(rt == rf);
                    }
            std::printf("flat equivalence: %d cases, tree == flat over the GPU model in all of them (reference: %d of %d)\n", cases,
                        tree_is_flat, cases);
        }
    } catch (const std::exception &e) {
        std::printf("FAIL exception: %s\n", e.what());
        ++g_failures;
    }
    std::printf(g_failures ? "FAILED (%d)\n" : "ALL OK\n", g_failures);
    return g_failures ? 1 : 0;
}
