// bench_search — BASELINE.json configs[4] / SURVEY.md 8(d) M4: the reference's own OffPolicy<Exp3<...>> search
// (src/algorithm/tree-bandit/tree/off-policy.hh:83-109, unmodified, compiled from the reference's headers)
// over (a) the reference MonteCarloModel and (b) the drop-in PinyonCuda::MonteCarloModel, with `leaves`
// actor iterations per learner iteration (= leaves per batched inference, 4096 in the config) on
// M1's root RandomTree{prng{0}, 10, 3, 3, 1}.
//
//   bench_search <ref|gpu|ref_tb|gpu_tb> <gpu_id> <device_seed> <leaves_per_batch> <learner_iterations> <warmup_iterations>
//   (ref_tb / gpu_tb: TreeBandit<Exp3>, one leaf per iteration; leaves_per_batch is ignored)
//
// Prints one JSON object: leaves/s over the timed learner iterations and the root statistics
// (exp3.hh:9-16) so that the caller can check ref == gpu for equal seeds and iteration counts, and
// sum them over root-parallel ranks.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include <model/monte-carlo-model.hh>
#include <state/random-tree.hh>

#include <algorithm/tree-bandit/bandit/exp3.hh>
#include <algorithm/tree-bandit/tree/off-policy.hh>
#include <algorithm/tree-bandit/tree/tree-bandit.hh>
#include <tree/tree.hh>

#include <pinyon_cuda/monte_carlo_model.hh>

using RefTree = MonteCarloModel<RandomTree<>>;
using GpuTree = PinyonCuda::MonteCarloModel<RandomTree<>>;

template <typename Types, typename MakeModel>
static int run(const char *impl, uint64_t device_seed, size_t leaves, size_t iters, size_t warmup, MakeModel make) {
    typename Types::PRNG device{device_seed};
    std::vector<typename Types::State> states;
    states.emplace_back(prng{0}, 10, 3, 3, 1);
    typename Types::Model model = make();
    std::vector<typename Types::MatrixNode> roots(1);
    typename Types::Search search{};
    if (warmup) search.run_for_iterations(warmup, leaves, device, states, model, roots);
    const auto t0 = std::chrono::steady_clock::now();
    search.run_for_iterations(iters, leaves, device, states, model, roots);
    const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    const auto &st = roots[0].stats;
    std::string rv, cv;
    for (auto v : st.row_visits) rv += (rv.empty() ? "" : ",") + std::to_string(v);
    for (auto v : st.col_visits) cv += (cv.empty() ? "" : ",") + std::to_string(v);
    std::printf("{\"impl\": \"%s\", \"leaves_per_batch\": %zu, \"learner_iterations\": %zu, \"warmup_iterations\": %zu, "
                "\"seconds\": %.6f, \"leaves_per_s\": %.3f, \"nodes\": %zu, \"visits\": %lld, \"value_total_row\": %.6f, "
                "\"row_visits\": [%s], \"col_visits\": [%s]}\n",
                impl, leaves, iters, warmup, sec, double(leaves) * double(iters) / sec, roots[0].count_matrix_nodes(),
                (long long)st.visits, double(st.value_total.get_row_value()), rv.c_str(), cv.c_str());
    return 0;
}

// Unmodified TreeBandit<Exp3<...>> (tree-bandit.hh:79-164): one single-state inference per iteration
// (monte-carlo-model.hh:41-56) -> one k_small launch per leaf on the GPU model.
template <typename Types, typename MakeModel>
static int run_tb(const char *impl, uint64_t device_seed, size_t iters, size_t warmup, MakeModel make) {
    typename Types::PRNG device{device_seed};
    typename Types::State state{prng{0}, 10, 3, 3, 1};
    typename Types::Model model = make();
    typename Types::MatrixNode root{};
    typename Types::Search search{};
    if (warmup) search.run_for_iterations(warmup, device, state, model, root);
    const auto t0 = std::chrono::steady_clock::now();
    search.run_for_iterations(iters, device, state, model, root);
    const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    const auto &st = root.stats;
    std::string rv, cv;
    for (auto v : st.row_visits) rv += (rv.empty() ? "" : ",") + std::to_string(v);
    for (auto v : st.col_visits) cv += (cv.empty() ? "" : ",") + std::to_string(v);
    std::printf("{\"impl\": \"%s\", \"leaves_per_batch\": 1, \"learner_iterations\": %zu, \"warmup_iterations\": %zu, "
                "\"seconds\": %.6f, \"leaves_per_s\": %.3f, \"nodes\": %zu, \"visits\": %lld, \"value_total_row\": %.6f, "
                "\"row_visits\": [%s], \"col_visits\": [%s]}\n",
                impl, iters, warmup, sec, double(iters) / sec, root.count_matrix_nodes(), (long long)st.visits,
                double(st.value_total.get_row_value()), rv.c_str(), cv.c_str());
    return 0;
}

int main(int argc, char **argv) {
    if (argc < 7) {
        std::fprintf(stderr, "usage: %s <ref|gpu> <gpu_id> <device_seed> <leaves> <iterations> <warmup>\n", argv[0]);
        return 2;
    }
    const bool gpu = !std::strcmp(argv[1], "gpu");
    const int gpu_id = std::atoi(argv[2]);
    const uint64_t seed = std::strtoull(argv[3], nullptr, 10);
    const size_t leaves = std::strtoull(argv[4], nullptr, 10), iters = std::strtoull(argv[5], nullptr, 10),
                 warmup = std::strtoull(argv[6], nullptr, 10);
    try {
        if (!std::strcmp(argv[1], "gpu_tb"))
            return run_tb<TreeBandit<Exp3<GpuTree>>>("gpu_tb", seed, iters, warmup, [&] {
                return GpuTree::Model{prng{0}, PinyonCuda::Device::shared(gpu_id)};
            });
        if (!std::strcmp(argv[1], "ref_tb"))
            return run_tb<TreeBandit<Exp3<RefTree>>>("ref_tb", seed, iters, warmup, [] { return RefTree::Model{prng{0}}; });
        if (gpu)
            return run<OffPolicy<Exp3<GpuTree>>>("gpu", seed, leaves, iters, warmup, [&] {
                return GpuTree::Model{prng{0}, PinyonCuda::Device::shared(gpu_id)};
            });
        return run<OffPolicy<Exp3<RefTree>>>("ref", seed, leaves, iters, warmup, [] { return RefTree::Model{prng{0}}; });
    } catch (const std::exception &e) {
        std::fprintf(stderr, "bench_search: %s\n", e.what());
        return 1;
    }
}
