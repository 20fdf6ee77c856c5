// Drop-in test of PinyonCuda::MonteCarloModel against the reference's own MonteCarloModel.
//
// Compiled (tests/cpp/build.py) with g++ -std=c++23 against the UNMODIFIED reference headers
// under /root/reference/src and linked to libpinyon_cuda.so.  The reference classes here are
// the checker; the GPU model is the thing checked.  Reads like the reference's own tests
// (tests/flat.cc, tests/cart/node.cc, benchmark/mold-state.cc): same type lists, same calls.
//
//   test_dropin --cpu   host logic only (engine access, stream locator, concepts, loud failure)
//   test_dropin --gpu   everything, on cuda:0
#include <array>
#include <atomic>
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <memory>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include <model/monte-carlo-model.hh>
#include <state/random-tree.hh>
#include <state/test-states.hh>

#include <algorithm/tree-bandit/bandit/exp3.hh>
#include <algorithm/tree-bandit/tree/off-policy.hh>
#include <algorithm/tree-bandit/tree/tree-bandit.hh>
#include <tree/tree.hh>

#include <pinyon_cuda/batcher.hh>
#include <pinyon_cuda/monte_carlo_model.hh>

using RefTree = MonteCarloModel<RandomTree<>>;
using GpuTree = PinyonCuda::MonteCarloModel<RandomTree<>>;
using RefMold = MonteCarloModel<MoldState<>>;
using GpuMold = PinyonCuda::MonteCarloModel<MoldState<>>;
using GpuTreePolicy = PinyonCuda::MonteCarloModel<RandomTree<>, true>;
using BatchedTree = PinyonCuda::BatchedMonteCarloModel<RandomTree<>>;
using BatchedMold = PinyonCuda::BatchedMonteCarloModel<MoldState<>>;
// ---- variants: a type list over Pinyon's xor_shift PRNG (with the member IsPRNG misses added)
using XsTypes = DefaultTypes<double, int, int, double, ConstantSum<1, 1>::Value, std::vector, Matrix, std::mutex, uint64_t,
                             PinyonCuda::xor_shift_prng>;
using RefXsTree = MonteCarloModel<RandomTree<XsTypes>>;
using GpuXsTree = PinyonCuda::MonteCarloModel<RandomTree<XsTypes>>;
static_assert(IsTypeList<XsTypes>);
static_assert(IsValueModelTypes<GpuXsTree> && IsSingleModelTypes<GpuXsTree> && IsBatchModelTypes<GpuXsTree>);

// ---- the concepts every caller constrains its model with (src/model/model.hh:8-79)
static_assert(IsValueModelTypes<GpuTree>);
static_assert(IsSingleModelTypes<GpuTree>);
static_assert(IsBatchModelTypes<GpuTree>);
static_assert(IsValueModelTypes<GpuMold>);
static_assert(IsSingleModelTypes<GpuMold>);
static_assert(IsBatchModelTypes<GpuMold>);
static_assert(IsPolicyModelTypes<GpuTreePolicy>);
static_assert(IsSingleModelTypes<GpuTreePolicy>);
static_assert(std::is_copy_constructible_v<GpuTree::Model>);
// ---- and the callers named in north_star instantiate over it
static_assert(IsBanditAlgorithmTypes<Exp3<GpuTree>>);
static_assert(IsSearchTypes<TreeBandit<Exp3<GpuTree>>>);
static_assert(IsSearchTypes<TreeBandit<Exp3<GpuMold>>>);
// ---- the batching adapter is a model like any other
static_assert(IsValueModelTypes<BatchedTree> && IsSingleModelTypes<BatchedTree> && IsBatchModelTypes<BatchedTree>);
static_assert(IsSearchTypes<TreeBandit<Exp3<BatchedTree>>> && IsSearchTypes<TreeBandit<Exp3<BatchedMold>>>);
static_assert(std::is_copy_constructible_v<BatchedTree::Model>);

static int g_failures = 0;
#define CHECK(cond)                                                          \
    do {                                                                     \
        if (!(cond)) {                                                       \
            std::printf("FAIL %s:%d: %s\n", __FILE__, __LINE__, #cond);      \
            ++g_failures;                                                    \
        }                                                                    \
    } while (0)

// ------------------------------------------------------------------ host-only checks

static void test_engine_access() {
    for (uint64_t seed : {0ull, 7ull, 12345678901234ull}) {
        for (size_t adv : {size_t(0), size_t(1), size_t(623), size_t(624), size_t(625), size_t(5000)}) {
            prng a{seed};
            a.discard(adv);
            PinyonCuda::EngineImage fast, slow;
            PinyonCuda::EngineAccess<prng>::read(a, fast);
            PinyonCuda::detail::image_from_stream(PinyonCuda::detail::engine_of(a), slow);
            CHECK(fast.x == slow.x && fast.p == slow.p);
            // the C ABI's host engine continues the image exactly as the prng continues
            PinyonCuda::EngineImage img = fast;
            for (int i = 0; i < 1500; ++i) {
                const uint64_t hi = pcu_mt_next(img.x.data(), &img.p), lo = pcu_mt_next(img.x.data(), &img.p);
                CHECK(((hi << 32) | lo) == a.uniform_64());
            }
            // write-back puts a fresh prng in the same place
            prng b{seed + 99};
            PinyonCuda::EngineAccess<prng>::write(b, img);
            for (int i = 0; i < 1500; ++i) CHECK(a.uniform_64() == b.uniform_64());
        }
    }
    std::printf("engine access ok (libstdc++ peek: %d)\n", int(PinyonCuda::detail::GLIBCXX_PEEK));
}

static void test_stream_locator() {
    PinyonCuda::StreamLocator loc;
    for (uint64_t seed : {0ull, 3ull, (1ull << 32) + 5}) {
        prng d{seed};
        uint64_t drawn = 0;
        prng walker{seed * 31 + 1};
        for (int i = 0; i < 400; ++i) {
            PinyonCuda::EngineImage img;
            PinyonCuda::EngineAccess<prng>::read(d, img);
            CHECK(loc.locate(seed, img) == drawn);
            const size_t adv = walker.random_int(1500);
            d.discard(adv);
            drawn += adv;
        }
        PinyonCuda::EngineImage img;
        PinyonCuda::EngineAccess<prng>::read(prng{seed + 1}, img);
        prng other{seed + 1};
        other.discard(10);
        PinyonCuda::EngineAccess<prng>::read(other, img);
        CHECK(loc.locate(seed, img) != 10 || seed + 1 == seed);
    }
    // a RandomTree root and its descendants: position == draws its device made
    RandomTree<>::State s{prng{5}, 12, 4, 3, 2};
    PinyonCuda::TreeImage t = PinyonCuda::serialise_tree(s, loc);
    CHECK(t.state_pos == 2 * 4 * 3 * 2 && t.depth == 12 && t.payoff_bias == 0 && t.state_seed == 5);
    uint64_t expect = t.state_pos;
    for (int step = 0; step < 6; ++step) {
        const int r = step % 4, c = (step * 2) % 3, ch = step % 2;
        s.apply_actions(r, c, ch);
        expect += (r * 3 * 2 + c * 2 + ch) + 2 + 2 * 4 * 3 * 2;
        t = PinyonCuda::serialise_tree(s, loc);
        CHECK(t.state_pos == expect && t.depth == 12 - step - 1);
    }
    std::printf("stream locator ok\n");
}

// The fiber scheduler of the batching adapter, without a GPU: fibers park on tickets, a pass resumes exactly those
// whose ticket is ready, an exception inside a fiber surfaces in rethrow().
static void test_fibers_host_only() {
    PinyonCuda::FiberScheduler sched;
    PinyonCuda::detail::tl_scheduler = &sched;
    std::vector<PinyonCuda::Ticket> tickets(100);
    long sum = 0;
    for (int i = 0; i < 100; ++i)
        sched.spawn(
            [&, i] {
                const double x = i * 1.5; /* a floating-point local must survive the switches */
                for (int k = 0; k < 3; ++k) {
                    PinyonCuda::FiberScheduler::current()->park(&tickets[i]);
                    tickets[i].ready = 0;
                    sum += long(x) + k;
                }
                if (i == 7) throw PinyonCuda::Error(PCU_ERR_INVALID, "fiber 7");
            },
            64 << 10);
    CHECK(PinyonCuda::FiberScheduler::current() == nullptr); /* the main context is not a fiber */
    auto r = sched.pass();
    CHECK(r.first == 100 && r.second && sum == 0);
    r = sched.pass();
    CHECK(r.first == 100 && !r.second); /* every fiber is parked on an unready ticket */
    for (int i = 0; i < 50; ++i) tickets[i].ready = 1;
    r = sched.pass();
    long expect = 0;
    for (int i = 0; i < 50; ++i) expect += long(i * 1.5);
    CHECK(r.second && sum == expect);
    for (int round = 0; round < 8 && r.first; ++round) {
        for (auto &t : tickets) t.ready = 1;
        r = sched.pass();
    }
    CHECK(r.first == 0);
    bool threw = false;
    try {
        sched.rethrow();
    } catch (const PinyonCuda::Error &e) {
        threw = std::string(e.what()) == "fiber 7";
    }
    CHECK(threw);
    PinyonCuda::detail::tl_scheduler = nullptr;
    std::printf("fibers ok\n");
}

static void test_loud_failure_without_gpu() {
    pcu_ctx *probe = nullptr;
    if (pcu_create(0, &probe) == PCU_OK) { /* a GPU is present: nothing to fail about */
        pcu_destroy(probe);
        return;
    }
    bool threw = false;
    try {
        GpuTree::Model model{prng{0}};
    } catch (const PinyonCuda::Error &e) {
        threw = true;
        std::printf("no GPU -> %s\n", e.what());
    }
    CHECK(threw);
}

// ------------------------------------------------------------------ GPU parity checks

template <typename A, typename B>
static bool same_value(const A &a, const B &b) {
    return a.get_row_value() == b.get_row_value() && a.get_col_value() == b.get_col_value();
}

// Single-state inference in sequence on one model: values AND the model engine stay in lock-step.
static void test_single_inference() {
    struct Cfg { uint64_t seed; size_t depth, rows, cols, T; int thr_p, thr_q; uint64_t tseed; };
    const Cfg cfgs[] = {{0, 10, 3, 3, 1, 0, 1, 0}, {0, 20, 9, 9, 3, 0, 1, 0}, {7, 6, 2, 5, 4, 3, 10, 0xdeadbeefcafeull},
                        {(1ull << 32) + 5, 8, 4, 4, 2, 1, 2, 1}};
    for (const Cfg &c : cfgs) {
        RandomTree<>::State root{prng{c.seed}, c.depth, c.rows, c.cols, c.T, Rational<>{c.thr_p, c.thr_q}};
        root.randomize_transition(c.tseed);
        RefTree::Model ref{prng{11}};
        GpuTree::Model gpu{prng{11}};
        for (int i = 0; i < 40; ++i) {
            RefTree::ModelOutput a;
            GpuTree::ModelOutput b;
            ref.inference(RandomTree<>::State{root}, a);
            gpu.inference(RandomTree<>::State{root}, b);
            CHECK(same_value(a.value, b.value));
        }
        CHECK(ref.device.uniform_64() == gpu.device.uniform_64());
    }
    {
        RefMold::Model ref{prng{3}};
        GpuMold::Model gpu{prng{3}};
        for (int i = 0; i < 20; ++i) {
            RefMold::ModelOutput a;
            GpuMold::ModelOutput b;
            ref.inference(MoldState<>::State{2, 10}, a);
            gpu.inference(MoldState<>::State{2, 10}, b);
            CHECK(same_value(a.value, b.value));
        }
        CHECK(ref.device.uniform_64() == gpu.device.uniform_64());
    }
    {   // has_policy = true: uniform policies (monte-carlo-model.hh:45-53)
        GpuTreePolicy::Model gpu{prng{0}};
        MonteCarloModel<RandomTree<>, true>::Model ref{prng{0}};
        GpuTreePolicy::ModelOutput b;
        MonteCarloModel<RandomTree<>, true>::ModelOutput a;
        RandomTree<>::State root{prng{0}, 5, 3, 4, 1};
        ref.inference(RandomTree<>::State{root}, a);
        gpu.inference(RandomTree<>::State{root}, b);
        CHECK(same_value(a.value, b.value));
        CHECK(a.row_policy == b.row_policy && a.col_policy == b.col_policy && b.row_policy.size() == 3);
    }
    std::printf("single inference ok\n");
}

// Batched inference over mid-tree leaves (different positions, depths, biases; some terminal),
// exactly the batches a search produces (off-policy.hh:131-147).
static void test_batch_inference() {
    struct Cfg { uint64_t seed; size_t depth, rows, cols, T; int thr_p, thr_q; };
    const Cfg cfgs[] = {{0, 10, 3, 3, 1, 0, 1}, {0, 20, 9, 9, 3, 0, 1}, {99, 5, 4, 2, 3, 3, 10}};
    for (const Cfg &c : cfgs) {
        const RandomTree<>::State root{prng{c.seed}, c.depth, c.rows, c.cols, c.T, Rational<>{c.thr_p, c.thr_q}};
        prng walker{c.seed + 17};
        RefTree::Model ref{prng{5}};
        GpuTree::Model gpu{prng{5}};
        RefTree::ModelBatchInput ref_in{};
        GpuTree::ModelBatchInput gpu_in{};
        const int n = 1500;
        for (int i = 0; i < n; ++i) {
            RandomTree<>::State s{root};
            if (i % 7 == 3) s.randomize_transition(walker);
            const int down = walker.random_int(static_cast<int>(c.depth) + 1); /* == depth -> terminal */
            for (int k = 0; k < down; ++k) {
                s.apply_actions(walker.random_int(c.rows), walker.random_int(c.cols));
                s.get_actions();
            }
            RandomTree<>::State s2{s};
            ref.add_to_batch_input(std::move(s), ref_in);
            gpu.add_to_batch_input(std::move(s2), gpu_in);
        }
        CHECK(gpu_in.size() == size_t(n));
        RefTree::ModelBatchOutput ref_out{};
        GpuTree::ModelBatchOutput gpu_out{};
        ref.inference(ref_in, ref_out);
        gpu.inference(gpu_in, gpu_out);
        CHECK(gpu_out.size() == ref_out.size());
        int bad = 0;
        double mean = 0;
        size_t live = 0;
        for (int i = 0; i < n; ++i) {
            RefTree::ModelOutput a;
            GpuTree::ModelOutput b;
            ref.get_output(a, ref_out, i);
            gpu.get_output(b, gpu_out, i);
            bad += !same_value(a.value, b.value);
            if (gpu_out.slot[i] >= 0) {
                mean += a.value.get_row_value();
                ++live;
            }
        }
        CHECK(bad == 0);
        CHECK(live && gpu_out.mean_row_value() == mean / double(live)); /* exact: payoffs are 0, 1/2, 1 */
        CHECK(ref.device.uniform_64() == gpu.device.uniform_64());
    }
    {   // MoldState batch, ragged depths
        RefMold::Model ref{prng{8}};
        GpuMold::Model gpu{prng{8}};
        RefMold::ModelBatchInput ref_in{};
        GpuMold::ModelBatchInput gpu_in{};
        for (int i = 0; i < 700; ++i) {
            MoldState<>::State s{3, size_t(i % 13)}; /* depth 0 is terminal at construction */
            MoldState<>::State s2{s};
            ref.add_to_batch_input(std::move(s), ref_in);
            gpu.add_to_batch_input(std::move(s2), gpu_in);
        }
        RefMold::ModelBatchOutput ref_out{};
        GpuMold::ModelBatchOutput gpu_out{};
        ref.inference(ref_in, ref_out);
        gpu.inference(gpu_in, gpu_out);
        for (int i = 0; i < 700; ++i) {
            RefMold::ModelOutput a;
            GpuMold::ModelOutput b;
            ref.get_output(a, ref_out, i);
            gpu.get_output(b, gpu_out, i);
            CHECK(same_value(a.value, b.value));
        }
        CHECK(ref.device.uniform_64() == gpu.device.uniform_64());
    }
    std::printf("batch inference ok\n");
}

// Unmodified callers over the GPU model: TreeBandit<Exp3> (tree-bandit.hh:79-164, single-leaf
// inference) and OffPolicy<Exp3> (off-policy.hh:83-109, batched leaves) must build the same tree
// as over the reference model — same seeds, same root statistics.
template <typename Types>
static std::vector<double> tree_bandit_root(size_t iterations) {
    typename Types::PRNG device{0};
    typename Types::State state{prng{0}, 10, 3, 3, 1};
    typename Types::Model model{prng{0}};
    typename Types::MatrixNode root{};
    typename Types::Search search{};
    search.run_for_iterations(iterations, device, state, model, root);
    std::vector<double> s{double(root.count_matrix_nodes()), double(root.stats.visits),
                          double(root.stats.value_total.get_row_value())};
    for (auto v : root.stats.row_visits) s.push_back(v);
    for (auto v : root.stats.col_visits) s.push_back(v);
    return s;
}

template <typename Types>
static std::vector<double> off_policy_root(size_t learner, size_t actor) {
    typename Types::PRNG device{0};
    std::vector<typename Types::State> states;
    states.emplace_back(prng{0}, 10, 3, 3, 1);
    typename Types::Model model{prng{0}};
    std::vector<typename Types::MatrixNode> roots(1);
    typename Types::Search search{};
    search.run_for_iterations(learner, actor, device, states, model, roots);
    std::vector<double> s{double(roots[0].count_matrix_nodes()), double(roots[0].stats.visits),
                          double(roots[0].stats.value_total.get_row_value())};
    for (auto v : roots[0].stats.row_visits) s.push_back(v);
    for (auto v : roots[0].stats.col_visits) s.push_back(v);
    return s;
}

static void test_search_callers() {
    const auto a = tree_bandit_root<TreeBandit<Exp3<RefTree>>>(1024);
    const auto b = tree_bandit_root<TreeBandit<Exp3<GpuTree>>>(1024);
    CHECK(a == b);
    std::printf("TreeBandit<Exp3>: nodes %.0f visits %.0f value_total %.1f  (ref)  vs  %.0f %.0f %.1f  (gpu)\n", a[0],
                a[1], a[2], b[0], b[1], b[2]);
    const auto c = off_policy_root<OffPolicy<Exp3<RefTree>>>(16, 256);
    const auto d = off_policy_root<OffPolicy<Exp3<GpuTree>>>(16, 256);
    CHECK(c == d);
    std::printf("OffPolicy<Exp3>:  nodes %.0f visits %.0f value_total %.1f  (ref)  vs  %.0f %.0f %.1f  (gpu)\n", c[0],
                c[1], c[2], d[0], d[1], d[2]);
}

// Copies of a model share the GPU context and evaluate concurrently from threads
// (multithreaded.hh:112-113 gives every thread its own Model copy).
static void test_threaded_copies() {
    GpuTree::Model proto{prng{2}};
    const RandomTree<>::State root{prng{4}, 8, 3, 3, 2};
    std::vector<std::vector<double>> got(4);
    std::vector<std::thread> pool;
    for (int t = 0; t < 4; ++t)
        pool.emplace_back([&, t] {
            GpuTree::Model mine{proto};
            for (int i = 0; i < 50; ++i) {
                GpuTree::ModelOutput o;
                mine.inference(RandomTree<>::State{root}, o);
                got[t].push_back(o.value.get_row_value());
            }
        });
    for (auto &th : pool) th.join();
    RefTree::Model ref{prng{2}};
    std::vector<double> want;
    for (int i = 0; i < 50; ++i) {
        RefTree::ModelOutput o;
        ref.inference(RandomTree<>::State{root}, o);
        want.push_back(o.value.get_row_value());
    }
    for (int t = 0; t < 4; ++t) CHECK(got[t] == want); /* same start state in every copy (Appendix C8) */
    std::printf("threaded copies ok\n");
}

// Root-parallel batching of UNMODIFIED TreeBandit (tree-bandit.hh:79-164): many independent searches in fibers, their
// single-leaf inferences evaluated as batches in model-rng mode W.  Every tree must be identical to the same tree
// searched alone over the reference model: same device seed, same model seed -> same root statistics, and the same
// model engine afterwards.
template <typename Types>
static std::vector<double> root_stats(const typename Types::MatrixNode &root) {
    std::vector<double> s{double(root.count_matrix_nodes()), double(root.stats.visits), double(root.stats.value_total.get_row_value())};
    for (auto v : root.stats.row_visits) s.push_back(v);
    for (auto v : root.stats.col_visits) s.push_back(v);
    for (auto v : root.stats.row_gains) s.push_back(double(v));
    return s;
}

static void test_batched_tree_bandit(size_t threads, size_t fibers, size_t iterations) {
    using B = TreeBandit<Exp3<BatchedTree>>;
    using R = TreeBandit<Exp3<RefTree>>;
    const size_t trees = threads * fibers;
    struct Cfg { uint64_t seed; size_t depth, rows, cols, T; };
    const Cfg cfgs[] = {{0, 10, 3, 3, 1}, {5, 6, 3, 2, 2}};
    for (const Cfg &c : cfgs) {
        auto batcher = std::make_shared<PinyonCuda::LeafBatcher>(PinyonCuda::Device::shared(0), 512);
        PinyonCuda::FiberPool pool(threads, fibers, {batcher});
        std::vector<std::unique_ptr<B::MatrixNode>> roots(trees);
        std::vector<uint64_t> engine_after(trees);
        const RandomTree<>::State root_state{prng{c.seed}, c.depth, c.rows, c.cols, c.T};
        pool.run(trees, [&](size_t k, size_t w) {
            B::PRNG device{k};
            B::Model model{prng{1000 + k}, pool.batcher(w)};
            roots[k] = std::make_unique<B::MatrixNode>();
            B::Search search{};
            search.run_for_iterations(iterations, device, root_state, model, *roots[k]);
            engine_after[k] = model.device.uniform_64();
        });
        size_t bad = 0;
        for (size_t k = 0; k < trees; ++k) {
            R::PRNG device{k};
            R::Model model{prng{1000 + k}};
            R::MatrixNode ref_root{};
            R::Search search{};
            search.run_for_iterations(iterations, device, root_state, model, ref_root);
            bad += root_stats<R>(ref_root) != root_stats<B>(*roots[k]);
            bad += model.device.uniform_64() != engine_after[k];
        }
        CHECK(bad == 0);
        const auto st = batcher->stats();
        CHECK(st.leaves > 0 && st.calls > 0 && st.leaves / st.calls > 1);
        std::printf("batched TreeBandit<Exp3>: %zu trees x %zu iterations on %zu threads, %llu leaves in %llu GPU calls (%.0f per call), "
                    "every tree == reference\n", trees, iterations, threads, (unsigned long long)st.leaves,
                    (unsigned long long)st.calls, double(st.leaves) / double(st.calls));
    }
    {   // from a plain thread the same model is a synchronous drop-in (single leaves, flushed at once)
        RefTree::Model ref{prng{4}};
        BatchedTree::Model gpu{prng{4}};
        const RandomTree<>::State root{prng{2}, 7, 4, 3, 2};
        for (int i = 0; i < 30; ++i) {
            RefTree::ModelOutput a;
            BatchedTree::ModelOutput b;
            ref.inference(RandomTree<>::State{root}, a);
            gpu.inference(RandomTree<>::State{root}, b);
            CHECK(same_value(a.value, b.value));
        }
        CHECK(ref.device.uniform_64() == gpu.device.uniform_64());
        RefMold::Model mref{prng{6}};
        BatchedMold::Model mgpu{prng{6}};
        for (int i = 0; i < 10; ++i) {
            RefMold::ModelOutput a;
            BatchedMold::ModelOutput b;
            mref.inference(MoldState<>::State{3, 9}, a);
            mgpu.inference(MoldState<>::State{3, 9}, b);
            CHECK(same_value(a.value, b.value));
        }
        CHECK(mref.device.uniform_64() == mgpu.device.uniform_64());
    }
    std::printf("batching adapter ok\n");
}

// Variants: the xor_shift type list, and RandomTree states grown with GrowLib::alternating_constant_actions, through the
// drop-in model against the reference's own MonteCarloModel over the same type list (single and batched inference,
// mid-tree leaves, model engine in lock-step).
template <typename Ref, typename Gpu, typename MakeState>
static void variant_case(const char *name, MakeState make_root, uint64_t model_seed) {
    using State = typename Ref::State;
    using PRNG = typename Ref::PRNG;
    const State root = make_root();
    typename Ref::Model ref{PRNG{model_seed}};
    typename Gpu::Model gpu{PRNG{model_seed}};
    for (int i = 0; i < 25; ++i) { /* single leaves */
        typename Ref::ModelOutput a;
        typename Gpu::ModelOutput b;
        ref.inference(State{root}, a);
        gpu.inference(State{root}, b);
        CHECK(same_value(a.value, b.value));
    }
    CHECK(ref.device.uniform_64() == gpu.device.uniform_64());
    PRNG walker{model_seed + 17};
    typename Ref::ModelBatchInput ref_in{};
    typename Gpu::ModelBatchInput gpu_in{};
    const int n = 1200;
    for (int i = 0; i < n; ++i) { /* mid-tree leaves, some terminal */
        State s{root};
        if (i % 5 == 2) s.randomize_transition(walker);
        const int down = walker.random_int(static_cast<int>(root.depth_bound) + 1);
        for (int k = 0; k < down && !s.is_terminal(); ++k) {
            s.apply_actions(walker.random_int(static_cast<int>(s.rows)), walker.random_int(static_cast<int>(s.cols)));
            s.get_actions();
        }
        State s2{s};
        ref.add_to_batch_input(std::move(s), ref_in);
        gpu.add_to_batch_input(std::move(s2), gpu_in);
    }
    typename Ref::ModelBatchOutput ref_out{};
    typename Gpu::ModelBatchOutput gpu_out{};
    ref.inference(ref_in, ref_out);
    gpu.inference(gpu_in, gpu_out);
    int bad = 0;
    for (int i = 0; i < n; ++i) {
        typename Ref::ModelOutput a;
        typename Gpu::ModelOutput b;
        ref.get_output(a, ref_out, i);
        gpu.get_output(b, gpu_out, i);
        bad += !same_value(a.value, b.value);
    }
    CHECK(bad == 0);
    CHECK(ref.device.uniform_64() == gpu.device.uniform_64());
    std::printf("variant %s ok\n", name);
}

static void test_variants() {
    using XsState = RandomTree<XsTypes>::State;
    variant_case<RefXsTree, GpuXsTree>("xor_shift 3x3x1", [] { return XsState{PinyonCuda::xor_shift_prng{5}, 8, 3, 3, 1}; }, 77);
    variant_case<RefXsTree, GpuXsTree>("xor_shift 4x3x3 thr 3/10",
                                       [] { return XsState{PinyonCuda::xor_shift_prng{99}, 6, 4, 3, 3, Rational<>{3, 10}}; }, 1234567);
    using St = RandomTree<>::State;
    auto grown = [](uint64_t seed, size_t rows, size_t cols, size_t T) {
        return St{prng{seed}, 7, rows, cols, T, Rational<>{0, 1}, &St::depth_bound_default,
                  &GrowLib::alternating_constant_actions<St>, &St::payoff_bias_default};
    };
    variant_case<RefTree, GpuTree>("alternating_constant_actions 1x3x2", [&] { return grown(3, 1, 3, 2); }, 5);
    variant_case<RefTree, GpuTree>("alternating_constant_actions 3x3x1", [&] { return grown(8, 3, 3, 1); }, 6);
    using XsGrown = RandomTree<XsTypes>::State;
    variant_case<RefXsTree, GpuXsTree>("xor_shift + alternating_constant_actions 3x1x2", [] {
        return XsGrown{PinyonCuda::xor_shift_prng{41}, 6, 3, 1, 2, Rational<>{0, 1}, &XsGrown::depth_bound_default,
                       &GrowLib::alternating_constant_actions<XsGrown>, &XsGrown::payoff_bias_default};
    }, 9);
    std::printf("variants ok\n");
}

// ONE model.inference(batch_in, batch_out) on a context that spans several GPUs (pcu_create_multi: contiguous
// rollout-index shards, per-shard slices of the one model stream, NCCL all-reduce of the counters) must equal the same
// call on one GPU byte for byte, including model.device afterwards (monte-carlo-model.hh:58-67).  With one GPU in the box
// the shards share it (device 0 listed three times), which runs the same plan, offsets and engine hand-back.
static void test_multi_gpu_batch(size_t n) {
    std::vector<int> ids;
    const int visible = pcu_visible_devices();
    if (visible >= 2)
        for (int i = 0; i < visible; ++i) ids.push_back(i);
    else
        ids = {0, 0, 0};
    auto multi = std::make_shared<PinyonCuda::Device>(ids);
    GpuTree::Model one{prng{3}, PinyonCuda::Device::shared(0)}, many{prng{3}, multi};
    const RandomTree<>::State root{prng{0}, 10, 3, 3, 1};
    GpuTree::ModelBatchInput in_one{};
    for (size_t i = 0; i < n; ++i) one.add_to_batch_input(RandomTree<>::State{root}, in_one);
    GpuTree::ModelBatchInput in_many = in_one;
    GpuTree::ModelBatchOutput out_one{}, out_many{};
    one.inference(in_one, out_one);
    many.inference(in_many, out_many);
    /* batches below the context's multi_min_rollouts (default 2^16 per GPU) stay on its first GPU: same answers, no sharding */
    if (n >= (size_t(1) << 20)) CHECK(std::string(pcu_last_path(multi->ctx)).find(" x") != std::string::npos);
    CHECK(pcu_device_count(multi->ctx) == int(ids.size()));
    CHECK(out_one.payoff_x2.size() == n && out_one.payoff_x2 == out_many.payoff_x2);
    for (int k = 0; k < 8; ++k) CHECK(out_one.sums[k] == out_many.sums[k]);
    CHECK(out_one.sums[0] + out_one.sums[1] + out_one.sums[2] == n);
    CHECK(one.device.uniform_64() == many.device.uniform_64());
    // and the head of the batch against the reference model itself
    RefTree::Model ref{prng{3}};
    for (size_t i = 0; i < std::min<size_t>(3000, n); ++i) {
        RefTree::ModelOutput a;
        GpuTree::ModelOutput b;
        ref.inference(RandomTree<>::State{root}, a);
        many.get_output(b, out_many, (long)i);
        CHECK(same_value(a.value, b.value));
    }
    std::printf("multi-GPU batch ok: %zu leaves over %zu shards (%d visible GPUs), path %s\n", n, ids.size(), visible,
                pcu_last_path(multi->ctx));
}

int main(int argc, char **argv) {
    if (argc > 2 && !std::strcmp(argv[1], "--multi")) {
        try {
            test_multi_gpu_batch(std::strtoull(argv[2], nullptr, 10));
        } catch (const std::exception &e) {
            std::printf("FAIL exception: %s\n", e.what());
            ++g_failures;
        }
        std::printf(g_failures ? "FAILED (%d)\n" : "ALL OK\n", g_failures);
        return g_failures ? 1 : 0;
    }
    const bool gpu = argc > 1 && !std::strcmp(argv[1], "--gpu");
    test_engine_access();
    test_stream_locator();
    test_fibers_host_only();
    if (!gpu) {
        test_loud_failure_without_gpu();
    } else {
        try {
            test_single_inference();
            test_batch_inference();
            test_search_callers();
            test_threaded_copies();
            test_batched_tree_bandit(4, 64, 200);
            test_variants();
        } catch (const std::exception &e) {
            std::printf("FAIL exception: %s\n", e.what());
            ++g_failures;
        }
    }
    std::printf(g_failures ? "FAILED (%d)\n" : "ALL OK\n", g_failures);
    return g_failures ? 1 : 0;
}
