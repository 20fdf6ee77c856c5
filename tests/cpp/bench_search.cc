// bench_search — BASELINE.json configs[4] / SURVEY.md 8(d) M4: the reference's own OffPolicy<Exp3<...>> search
// (src/algorithm/tree-bandit/tree/off-policy.hh:83-109, unmodified, compiled from the reference's headers)
// over (a) the reference MonteCarloModel and (b) the drop-in PinyonCuda::MonteCarloModel, with `leaves`
// actor iterations per learner iteration (= leaves per batched inference, 4096 in the config) on
// M1's root RandomTree{prng{0}, 10, 3, 3, 1}.
//
//   bench_search <ref|gpu|ref_tb|gpu_tb> <gpu_id> <device_seed> <leaves_per_batch> <learner_iterations> <warmup_iterations>
//   (ref_tb / gpu_tb: TreeBandit<Exp3>, one leaf per iteration; leaves_per_batch is ignored)
//   bench_search gpu_tb_batched <gpu_id|-1> <seed> <leaves_per_batch> <iterations_per_tree> 0 <threads> <fibers_per_thread>
//   bench_search ref_tb_mt 0 <seed> 0 <iterations_per_tree> 0 <threads> <trees_per_thread>
//   bench_search <gpu_op_mt|ref_op_mt> <gpu_id> <seed> <leaves_per_batch> <learner_iterations> <warmup> <threads>
//
// Prints one JSON object: leaves/s over the timed learner iterations and the root statistics
// (exp3.hh:9-16) so that the caller can check ref == gpu for equal seeds and iteration counts, and
// sum them over root-parallel ranks.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

#include <model/monte-carlo-model.hh>
#include <state/random-tree.hh>

#include <algorithm/tree-bandit/bandit/exp3.hh>
#include <algorithm/tree-bandit/tree/off-policy.hh>
#include <algorithm/tree-bandit/tree/tree-bandit.hh>
#include <tree/tree.hh>

#include <pinyon_cuda/batcher.hh>
#include <pinyon_cuda/monte_carlo_model.hh>

using RefTree = MonteCarloModel<RandomTree<>>;
using GpuTree = PinyonCuda::MonteCarloModel<RandomTree<>>;
using BatchedTree = PinyonCuda::BatchedMonteCarloModel<RandomTree<>>;

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

// Root-parallel TreeBandit<Exp3> over the batching adapter (include/pinyon_cuda/batcher.hh): threads x fibers independent
// UNMODIFIED searches (tree k: device prng{seed + k}, model prng{0}), their single-leaf inferences evaluated `leaves` at a
// time in model-rng mode W; gpu_id < 0: one batcher per visible GPU, worker threads spread over them.
static int run_tb_batched(int gpu_id, uint64_t seed, size_t leaves, size_t iters, size_t threads, size_t fibers) {
    using B = TreeBandit<Exp3<BatchedTree>>;
    std::vector<std::shared_ptr<PinyonCuda::LeafBatcher>> batchers;
    if (gpu_id >= 0) {
        batchers.push_back(std::make_shared<PinyonCuda::LeafBatcher>(PinyonCuda::Device::shared(gpu_id), leaves));
    } else {
        for (int g = 0; g < std::max(1, pcu_visible_devices()); ++g)
            batchers.push_back(std::make_shared<PinyonCuda::LeafBatcher>(PinyonCuda::Device::shared(g), leaves));
    }
    PinyonCuda::FiberPool pool(threads, fibers, batchers);
    const size_t trees = threads * fibers;
    const RandomTree<>::State root_state{prng{0}, 10, 3, 3, 1};
    std::vector<std::unique_ptr<B::MatrixNode>> roots(trees);
    const auto t0 = std::chrono::steady_clock::now();
    pool.run(trees, [&](size_t k, size_t w) {
        B::PRNG device{seed + k};
        B::Model model{prng{0}, pool.batcher(w)};
        roots[k] = std::make_unique<B::MatrixNode>();
        B::Search search{};
        search.run_for_iterations(iters, device, root_state, model, *roots[k]);
    });
    const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    long long visits = 0;
    double value = 0;
    size_t nodes = 0;
    for (auto &r : roots) {
        visits += r->stats.visits;
        value += double(r->stats.value_total.get_row_value());
        nodes += r->count_matrix_nodes();
    }
    unsigned long long calls = 0, evaluated = 0, full = 0;
    for (auto &b : batchers) {
        const auto st = b->stats();
        calls += st.calls;
        evaluated += st.leaves;
        full += st.full;
    }
    std::printf("{\"impl\": \"gpu_tb_batched\", \"gpus\": %zu, \"threads\": %zu, \"fibers_per_thread\": %zu, \"trees\": %zu, "
                "\"iterations_per_tree\": %zu, \"leaves_per_batch\": %zu, \"seconds\": %.6f, \"leaves_per_s\": %.3f, "
                "\"gpu_calls\": %llu, \"gpu_leaves\": %llu, \"full_batches\": %llu, \"nodes\": %zu, \"visits\": %lld, "
                "\"value_total_row\": %.1f}\n",
                batchers.size(), threads, fibers, trees, iters, leaves, sec, double(trees) * double(iters) / sec, calls, evaluated,
                full, nodes, visits, value);
    return 0;
}

// The reference's TreeBandit<Exp3> over the reference model, one independent tree per host thread: what the box's
// CPU cores do root-parallel without a GPU.
static int run_tb_ref_mt(uint64_t seed, size_t iters, size_t threads, size_t trees_per_thread) {
    using R = TreeBandit<Exp3<RefTree>>;
    const RandomTree<>::State root_state{prng{0}, 10, 3, 3, 1};
    std::vector<long long> visits(threads, 0);
    const auto t0 = std::chrono::steady_clock::now();
    std::vector<std::thread> pool;
    for (size_t t = 0; t < threads; ++t)
        pool.emplace_back([&, t] {
            for (size_t j = 0; j < trees_per_thread; ++j) {
                R::PRNG device{seed + t * trees_per_thread + j};
                R::Model model{prng{0}};
                R::MatrixNode root{};
                R::Search search{};
                search.run_for_iterations(iters, device, root_state, model, root);
                visits[t] += root.stats.visits;
            }
        });
    for (auto &t : pool) t.join();
    const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    long long v = 0;
    for (auto x : visits) v += x;
    std::printf("{\"impl\": \"ref_tb_mt\", \"threads\": %zu, \"trees\": %zu, \"iterations_per_tree\": %zu, \"seconds\": %.6f, "
                "\"leaves_per_s\": %.3f, \"visits\": %lld}\n",
                threads, threads * trees_per_thread, iters, sec, double(threads * trees_per_thread) * double(iters) / sec, v);
    return 0;
}

// Root-parallel OffPolicy: `threads` host threads, each running the reference's unmodified
// OffPolicy<Exp3<...>> on its own tree (device prng{seed + t}, model prng{0}) with `leaves` actor iterations per batched
// inference; the GPU models share one context (calls serialise on its mutex: 0.06 us of GPU per leaf against ~10 us of
// host tree walk).  The same loop over the reference model is the CPU side of the comparison.
template <typename Types, typename MakeModel>
static int run_mt(const char *impl, uint64_t device_seed, size_t leaves, size_t iters, size_t warmup, size_t threads,
                  MakeModel make) {
    std::vector<long long> visits(threads, 0);
    std::vector<double> values(threads, 0.0);
    std::vector<std::string> errors(threads);
    std::vector<std::unique_ptr<typename Types::Model>> models;
    for (size_t t = 0; t < threads; ++t) models.push_back(std::make_unique<typename Types::Model>(make()));
    std::vector<std::vector<typename Types::MatrixNode>> roots; /* MatrixNode is neither copyable nor movable */
    roots.reserve(threads);
    for (size_t t = 0; t < threads; ++t) roots.emplace_back(1);
    auto body = [&](size_t t, size_t n_iter) {
        try {
            typename Types::PRNG device{device_seed + t};
            std::vector<typename Types::State> states;
            states.emplace_back(prng{0}, 10, 3, 3, 1);
            typename Types::Search search{};
            search.run_for_iterations(n_iter, leaves, device, states, *models[t], roots[t]);
        } catch (const std::exception &e) {
            errors[t] = e.what();
        }
    };
    auto sweep = [&](size_t n_iter) {
        std::vector<std::thread> pool;
        for (size_t t = 0; t < threads; ++t) pool.emplace_back(body, t, n_iter);
        for (auto &th : pool) th.join();
    };
    if (warmup) sweep(warmup);
    const auto t0 = std::chrono::steady_clock::now();
    sweep(iters);
    const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    for (auto &e : errors)
        if (!e.empty()) throw std::runtime_error(e);
    long long v = 0;
    double val = 0;
    size_t nodes = 0;
    for (size_t t = 0; t < threads; ++t) {
        v += roots[t][0].stats.visits;
        val += double(roots[t][0].stats.value_total.get_row_value());
        nodes += roots[t][0].count_matrix_nodes();
    }
    std::printf("{\"impl\": \"%s\", \"threads\": %zu, \"leaves_per_batch\": %zu, \"learner_iterations\": %zu, "
                "\"warmup_iterations\": %zu, \"seconds\": %.6f, \"leaves_per_s\": %.3f, \"nodes\": %zu, \"visits\": %lld, "
                "\"value_total_row\": %.6f}\n",
                impl, threads, leaves, iters, warmup, sec, double(threads) * double(leaves) * double(iters) / sec, nodes, v, val);
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
    const size_t threads = argc > 7 ? std::strtoull(argv[7], nullptr, 10) : std::max(1u, std::thread::hardware_concurrency());
    const size_t fibers = argc > 8 ? std::strtoull(argv[8], nullptr, 10) : 256;
    try {
        if (!std::strcmp(argv[1], "gpu_tb_batched")) return run_tb_batched(gpu_id, seed, leaves, iters, threads, fibers);
        if (!std::strcmp(argv[1], "ref_tb_mt")) return run_tb_ref_mt(seed, iters, threads, fibers);
        if (!std::strcmp(argv[1], "gpu_op_mt"))
            return run_mt<OffPolicy<Exp3<GpuTree>>>("gpu_op_mt", seed, leaves, iters, warmup, threads, [&] {
                return GpuTree::Model{prng{0}, PinyonCuda::Device::shared(gpu_id)};
            });
        if (!std::strcmp(argv[1], "ref_op_mt"))
            return run_mt<OffPolicy<Exp3<RefTree>>>("ref_op_mt", seed, leaves, iters, warmup, threads,
                                                     [] { return RefTree::Model{prng{0}}; });
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
