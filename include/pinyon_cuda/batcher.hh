// Root-parallel batching of single-leaf inference (BASELINE configs[4]).
//
// TreeBandit::run_iteration evaluates ONE leaf per iteration through model.inference(State&&, ModelOutput&)
// (src/algorithm/tree-bandit/tree/tree-bandit.hh:103,111) and cannot be suspended: served one leaf at a time a GPU
// loses to a CPU core (a launch and a synchronisation per 10-step rollout).  The way to feed the GPU from UNMODIFIED
// search code is to run many independent searches (root-parallel: one tree each, as TreeBanditThreaded gives every
// thread its own Model copy, multithreaded.hh:113) and let their single-leaf inferences rendezvous:
//
//   BatchedMonteCarloModel<Types>::Model::inference(State&&, ...)   serialises the leaf, takes the 4 * depth raw words
//        its OWN model engine would draw next (so `model.device` advances exactly as the reference's would and every
//        tree stays bit-identical to the same tree searched alone with the reference model), hands both to the
//        LeafBatcher and PARKS the calling fiber;
//   FiberPool     T host threads x F fibers (ucontext), each fiber running user code that calls an unmodified
//        Search::run_for_iterations on its own tree; a thread switches to its next runnable fiber whenever one parks;
//   LeafBatcher   collects parked leaves from all threads; a flusher thread issues ONE pcu_rollout_* call in model-rng
//        mode W per >= batch_size leaves (or as soon as every worker thread has run out of runnable fibers) and wakes
//        the fibers whose leaves came back.  While a batch is on the GPU the threads keep descending the other trees.
//
// Called from a plain thread (no fiber), inference() submits its leaf, asks for an immediate flush and blocks: the
// model stays usable everywhere a MonteCarloModel is.  There is no CPU evaluation anywhere in this file.
#pragma once

#include <sys/mman.h>
#include <ucontext.h>

#include <atomic>
#include <condition_variable>
#include <exception>
#include <functional>
#include <memory>
#include <mutex>
#include <thread>
#include <vector>

#include "monte_carlo_model.hh"

namespace PinyonCuda {

/* Where a parked inference finds its answer.  Lives in the frame of the parked call. */
struct Ticket {
    std::atomic<int> ready{0};
    uint8_t payoff_x2 = 0;
};

/* ------------------------------------------------------------------------------------------ fibers */

/* Switching fibers: glibc's swapcontext saves and restores the signal mask with a system call per switch (two per
 * parked leaf, ~1 us); on x86-64 the switch is six pushes, a stack-pointer swap and six pops instead.  Other
 * architectures (or -DPINYON_CUDA_FIBERS_UCONTEXT) use ucontext. */
#if defined(__x86_64__) && !defined(PINYON_CUDA_FIBERS_UCONTEXT)
#define PINYON_CUDA_FIBER_ASM 1
extern "C" void pinyon_cuda_fiber_switch(void **save_sp, void *load_sp);
asm(R"(
.text
.weak pinyon_cuda_fiber_switch
.type pinyon_cuda_fiber_switch,@function
pinyon_cuda_fiber_switch:
    pushq %rbp
    pushq %rbx
    pushq %r12
    pushq %r13
    pushq %r14
    pushq %r15
    movq %rsp, (%rdi)
    movq %rsi, %rsp
    popq %r15
    popq %r14
    popq %r13
    popq %r12
    popq %rbx
    popq %rbp
    ret
.size pinyon_cuda_fiber_switch,.-pinyon_cuda_fiber_switch
)");
#endif

class FiberScheduler;
namespace detail {
inline thread_local FiberScheduler *tl_scheduler = nullptr; /* non-null while this thread runs a FiberPool worker */
inline thread_local void *tl_starting_fiber = nullptr;      /* the fiber being entered for the first time */
}

class FiberScheduler {
    struct Fiber {
#ifdef PINYON_CUDA_FIBER_ASM
        void *sp = nullptr;
#else
        ucontext_t context;
#endif
        void *stack = nullptr;
        size_t stack_bytes = 0;
        std::function<void()> body;
        Ticket *waiting = nullptr;
        bool done = false;
        FiberScheduler *owner = nullptr;
    };
#ifdef PINYON_CUDA_FIBER_ASM
    void *main_sp_ = nullptr;
#else
    ucontext_t main_;
#endif
    std::vector<std::unique_ptr<Fiber>> fibers_;
    Fiber *running_ = nullptr;
    std::exception_ptr error_;

    static void run_body(Fiber *f) {
        try {
            f->body();
        } catch (...) {
            if (!f->owner->error_) f->owner->error_ = std::current_exception();
        }
        f->done = true;
        f->owner->to_main(f);
    }
#ifdef PINYON_CUDA_FIBER_ASM
    static void entry() { /* reached by the `ret` of the first switch into a fiber */
        run_body(static_cast<Fiber *>(detail::tl_starting_fiber));
        __builtin_trap(); /* a finished fiber is never resumed */
    }
    void to_main(Fiber *f) { pinyon_cuda_fiber_switch(&f->sp, main_sp_); }
    void to_fiber(Fiber *f) {
        detail::tl_starting_fiber = f;
        pinyon_cuda_fiber_switch(&main_sp_, f->sp);
    }
#else
    static void trampoline(unsigned hi, unsigned lo) {
        run_body(reinterpret_cast<Fiber *>((uintptr_t(hi) << 32) | uintptr_t(lo)));
    }
    void to_main(Fiber *f) { swapcontext(&f->context, &main_); }
    void to_fiber(Fiber *f) { swapcontext(&main_, &f->context); }
#endif

   public:
    FiberScheduler() = default;
    FiberScheduler(const FiberScheduler &) = delete;
    ~FiberScheduler() {
        for (auto &f : fibers_)
            if (f->stack) munmap(f->stack, f->stack_bytes);
    }

    /* The scheduler of the fiber the caller runs in, or null on a plain thread. */
    static FiberScheduler *current() {
        FiberScheduler *s = detail::tl_scheduler;
        return (s && s->running_) ? s : nullptr;
    }

    void spawn(std::function<void()> body, size_t stack_bytes) {
        auto f = std::make_unique<Fiber>();
        /* mmap: pages are touched (and zero-filled) only where a fiber's stack actually reaches */
        f->stack = mmap(nullptr, stack_bytes, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS | MAP_NORESERVE | MAP_STACK, -1, 0);
        if (f->stack == MAP_FAILED) throw Error(PCU_ERR_NOMEM, "mmap of a fiber stack failed");
        f->stack_bytes = stack_bytes;
        f->body = std::move(body);
        f->owner = this;
#ifdef PINYON_CUDA_FIBER_ASM
        /* [A] = &entry with A 16-byte aligned, so that entry() starts with the stack alignment of a called function;
         * six zeroed callee-saved registers below it for the first switch to pop */
        uintptr_t top = (reinterpret_cast<uintptr_t>(f->stack) + stack_bytes) & ~uintptr_t(15);
        void **a = reinterpret_cast<void **>(top - 32);
        a[0] = reinterpret_cast<void *>(&FiberScheduler::entry);
        a[1] = nullptr;
        for (int i = 1; i <= 6; ++i) a[-i] = nullptr;
        f->sp = a - 6;
#else
        getcontext(&f->context);
        f->context.uc_stack.ss_sp = f->stack;
        f->context.uc_stack.ss_size = stack_bytes;
        f->context.uc_link = nullptr;
        const uintptr_t p = reinterpret_cast<uintptr_t>(f.get());
        makecontext(&f->context, reinterpret_cast<void (*)()>(&FiberScheduler::trampoline), 2, unsigned(p >> 32), unsigned(p));
#endif
        fibers_.push_back(std::move(f));
    }

    /* From inside a fiber: give the thread away until `ticket` is ready. */
    void park(Ticket *ticket) {
        Fiber *f = running_;
        f->waiting = ticket;
        to_main(f);
    }

    /* One pass over the fibers: resume every one that can run.  Returns {alive, progressed}. */
    std::pair<size_t, bool> pass() {
        size_t alive = 0;
        bool progressed = false;
        for (auto &f : fibers_) {
            if (f->done) continue;
            ++alive;
            if (f->waiting) {
                if (!f->waiting->ready.load(std::memory_order_acquire)) continue;
                f->waiting = nullptr;
            }
            running_ = f.get();
            to_fiber(f.get());
            running_ = nullptr;
            progressed = true;
            if (f->done) --alive;
        }
        return {alive, progressed};
    }

    void rethrow() {
        if (error_) std::rethrow_exception(error_);
    }
};

/* ------------------------------------------------------------------------------------------ batcher */

class LeafBatcher {
   public:
    struct Leaf {
        Shape shape;
        bool mold = false;
        uint64_t state_seed = 0, transition_seed = 0;
        uint32_t state_pos = 0;
        uint16_t depth = 0;
        int32_t payoff_bias = 0;
    };
    struct Stats {
        uint64_t calls = 0, leaves = 0, full = 0; /* GPU calls, leaves evaluated, calls that reached batch_size */
    };

    std::shared_ptr<Device> gpu;
    const size_t batch_size;

    explicit LeafBatcher(std::shared_ptr<Device> gpu, size_t batch_size = 4096)
        : gpu{std::move(gpu)}, batch_size{batch_size ? batch_size : 1}, flusher_{[this] { flush_loop(); }} {}
    LeafBatcher(const LeafBatcher &) = delete;
    ~LeafBatcher() {
        {
            std::lock_guard<std::mutex> lock(mu_);
            stop_ = true;
        }
        cv_work_.notify_all();
        flusher_.join();
    }

    /* Queue one leaf with the 4 * depth raw model words of its own engine.  `ticket` is signalled by the flusher. */
    void submit(const Leaf &leaf, const uint32_t *words, Ticket *ticket) {
        std::lock_guard<std::mutex> lock(mu_);
        Batch *b = nullptr;
        for (auto &cand : pending_)
            if (cand.shape == leaf.shape && cand.mold == leaf.mold) b = &cand;
        if (!b) {
            pending_.emplace_back();
            b = &pending_.back();
            b->shape = leaf.shape;
            b->mold = leaf.mold;
        }
        b->state_seed.push_back(leaf.state_seed);
        b->transition_seed.push_back(leaf.transition_seed);
        b->state_pos.push_back(leaf.state_pos);
        b->depth.push_back(leaf.depth);
        b->payoff_bias.push_back(leaf.payoff_bias);
        b->words.insert(b->words.end(), words, words + 4 * size_t(leaf.depth));
        b->tickets.push_back(ticket);
        if (++pending_leaves_ >= batch_size) cv_work_.notify_one();
    }

    /* Completed flushes so far.  A worker reads it BEFORE polling its fibers and passes it to wait_for_flush: a flush
     * that completes in between is then never slept through. */
    uint64_t generation() const { return generation_.load(std::memory_order_acquire); }

    /* A worker thread with no runnable fiber: sleep until a flush completes after `seen`. */
    void wait_for_flush(uint64_t seen) {
        std::unique_lock<std::mutex> lock(mu_);
        ++idle_;
        cv_work_.notify_one(); /* possibly the last worker to go idle: flush what is there */
        cv_done_.wait(lock, [&] { return generation_.load(std::memory_order_relaxed) != seen || stop_; });
        --idle_;
    }

    /* A plain thread (no fiber) that submitted a leaf: flush now, block until it is back. */
    void wait_ticket(Ticket *ticket) {
        std::unique_lock<std::mutex> lock(mu_);
        ++urgent_;
        cv_work_.notify_one();
        cv_done_.wait(lock, [&] { return ticket->ready.load(std::memory_order_acquire) != 0 || stop_; });
        --urgent_;
    }

    void add_worker() {
        std::lock_guard<std::mutex> lock(mu_);
        ++workers_;
    }
    void remove_worker() {
        {
            std::lock_guard<std::mutex> lock(mu_);
            --workers_;
        }
        cv_work_.notify_one();
    }

    Stats stats() const {
        std::lock_guard<std::mutex> lock(mu_);
        return stats_;
    }
    std::exception_ptr error() const {
        std::lock_guard<std::mutex> lock(mu_);
        return error_;
    }

   private:
    struct Batch {
        Shape shape;
        bool mold = false;
        std::vector<uint64_t> state_seed, transition_seed;
        std::vector<uint32_t> state_pos;
        std::vector<uint16_t> depth;
        std::vector<int32_t> payoff_bias;
        std::vector<uint32_t> words;
        std::vector<Ticket *> tickets;
    };

    mutable std::mutex mu_;
    std::condition_variable cv_work_, cv_done_;
    std::vector<Batch> pending_;
    size_t pending_leaves_ = 0, workers_ = 0, idle_ = 0, urgent_ = 0;
    std::atomic<uint64_t> generation_{0};
    bool stop_ = false;
    Stats stats_;
    std::exception_ptr error_;
    std::vector<uint8_t> payoff_;
    std::thread flusher_; /* last member: starts when everything above exists */

    template <typename T>
    static bool all_equal(const std::vector<T> &v) {
        for (size_t i = 1; i < v.size(); ++i)
            if (v[i] != v[0]) return false;
        return true;
    }

    void evaluate(Batch &b) {
        const size_t n = b.tickets.size();
        payoff_.assign(n, 0);
        pcu_model_rng rng{};
        rng.mode = PCU_MODEL_WORDS;
        rng.words = b.words.data();
        pcu_out out{};
        out.mem = PCU_MEM_HOST;
        out.payoff_x2 = payoff_.data();
        std::lock_guard<std::mutex> lock(gpu->mutex);
        if (b.mold) {
            pcu_mold_batch mb{};
            mb.n = n;
            mb.mem = PCU_MEM_HOST;
            mb.n_actions = b.shape.rows;
            if (all_equal(b.depth))
                mb.depth_u = b.depth[0];
            else
                mb.depth = b.depth.data();
            gpu->check(pcu_rollout_mold(gpu->ctx, &mb, &rng, &out));
        } else {
            pcu_rt_batch tb{};
            tb.n = n;
            tb.mem = PCU_MEM_HOST;
            tb.rows = b.shape.rows;
            tb.cols = b.shape.cols;
            tb.transitions = b.shape.transitions;
            tb.thr_p = b.shape.thr_p;
            tb.thr_q = b.shape.thr_q;
            tb.growth = b.shape.growth;
            /* batch-uniform fields travel as scalars: leaves of searches from one root share the state seed and
             * take the library's state-stream table kernel */
            if (all_equal(b.state_seed)) tb.state_seed_u = b.state_seed[0]; else tb.state_seed = b.state_seed.data();
            if (all_equal(b.state_pos)) tb.state_pos_u = b.state_pos[0]; else tb.state_pos = b.state_pos.data();
            if (b.shape.transitions == 1 || all_equal(b.transition_seed)) tb.transition_seed_u = b.transition_seed[0];
            else tb.transition_seed = b.transition_seed.data();
            if (all_equal(b.depth)) tb.depth_u = b.depth[0]; else tb.depth = b.depth.data();
            if (all_equal(b.payoff_bias)) tb.payoff_bias_u = b.payoff_bias[0]; else tb.payoff_bias = b.payoff_bias.data();
            gpu->check(pcu_rollout_random_tree(gpu->ctx, &tb, &rng, &out));
        }
    }

    void flush_loop() {
        std::unique_lock<std::mutex> lock(mu_);
        for (;;) {
            cv_work_.wait(lock, [&] {
                return stop_ || pending_leaves_ >= batch_size ||
                       (pending_leaves_ > 0 && (urgent_ > 0 || (workers_ > 0 && idle_ >= workers_)));
            });
            if (stop_) return;
            std::vector<Batch> work;
            work.swap(pending_);
            const bool full = pending_leaves_ >= batch_size;
            pending_leaves_ = 0;
            lock.unlock();
            std::exception_ptr err;
            size_t leaves = 0;
            for (Batch &b : work) {
                try {
                    evaluate(b);
                    for (size_t i = 0; i < b.tickets.size(); ++i) b.tickets[i]->payoff_x2 = payoff_[i];
                } catch (...) {
                    err = std::current_exception(); /* the waiters are released and find the error */
                }
                leaves += b.tickets.size();
                for (Ticket *t : b.tickets) t->ready.store(err ? -1 : 1, std::memory_order_release);
            }
            lock.lock();
            if (err && !error_) error_ = err;
            stats_.calls += work.size();
            stats_.leaves += leaves;
            stats_.full += full;
            generation_.fetch_add(1, std::memory_order_release);
            cv_done_.notify_all();
        }
    }
};

/* ------------------------------------------------------------------------------------------ the model */

/* MonteCarloModel whose single-state inference goes through a LeafBatcher.  Batched inference (mode S) is inherited. */
template <IsPerfectInfoStateTypes Types, bool has_policy = false>
struct BatchedMonteCarloModel : MonteCarloModel<Types, has_policy> {
    using Base = MonteCarloModel<Types, has_policy>;
    using ModelOutput = typename Base::ModelOutput;

    class Model : public Base::Model {
       public:
        std::shared_ptr<LeafBatcher> batcher;

        Model(const typename Types::PRNG &device, std::shared_ptr<LeafBatcher> batcher)
            : Base::Model{device, batcher->gpu}, batcher{std::move(batcher)} {}
        /* same construction as the reference (monte-carlo-model.hh:39): a batcher of its own on the shared context */
        Model(const typename Types::PRNG &device)
            : Base::Model{device}, batcher{std::make_shared<LeafBatcher>(this->gpu)} {}

        using Base::Model::inference; /* the batched overload */

        void inference(typename Types::State &&state, ModelOutput &output) {
            Base::Model::fill_policy(state, output);
            if (state.is_terminal()) { /* rollout() does nothing (monte-carlo-model.hh:87) */
                output.value = state.get_payoff();
                return;
            }
            LeafBatcher::Leaf leaf;
            if constexpr (Base::IS_TREE) {
                static thread_local StreamLocator locator; /* per thread: no lock on the search's hot path */
                const TreeImage t = serialise_tree(state, locator);
                leaf.shape = t.shape;
                leaf.state_seed = t.state_seed;
                leaf.transition_seed = t.transition_seed;
                leaf.state_pos = t.state_pos;
                leaf.depth = t.depth;
                leaf.payoff_bias = t.payoff_bias;
            } else {
                Base::Model::check_square(state);
                leaf.mold = true;
                leaf.shape.rows = leaf.shape.cols = static_cast<uint32_t>(state.row_actions.size());
                leaf.depth = Base::Model::checked_depth(state.max_depth);
                output.value = state.get_payoff(); /* MoldState never writes its payoff (test-states.hh:43-49) */
            }
            /* the words this model's engine draws for the rollout: 2 random_int per step, 2 words each
             * (monte-carlo-model.hh:89-90; random.hh:26,28); `device` ends where the reference model's would */
            using Access = EngineAccess<typename Types::PRNG>;
            static_assert(Access::ENGINE == PCU_ENGINE_MT19937, "the batching adapter (model-rng mode W) serves prng type lists");
            EngineImage img;
            Access::read(this->device, img);
            uint32_t small[4 * 64];
            std::vector<uint32_t> large;
            uint32_t *words = small;
            if (leaf.depth > 64) {
                large.resize(4 * size_t(leaf.depth));
                words = large.data();
            }
            pcu_mt_take_raw(img.x.data(), &img.p, 4 * uint64_t(leaf.depth), words);
            Access::write(this->device, img);
            Ticket ticket;
            batcher->submit(leaf, words, &ticket);
            if (FiberScheduler *sched = FiberScheduler::current())
                sched->park(&ticket);
            else
                batcher->wait_ticket(&ticket);
            if (ticket.ready.load(std::memory_order_acquire) != 1) {
                if (auto err = batcher->error()) std::rethrow_exception(err);
                throw Error(PCU_ERR_CUDA, "leaf batch failed");
            }
            if constexpr (Base::IS_TREE) output.value = Base::Model::value_from_code(ticket.payoff_x2);
        }
    };
};

/* ------------------------------------------------------------------------------------------ the pool */

/* `threads` host threads x `fibers_per_thread` fibers.  run(n, fn) calls fn(task, worker) for task = 0..n-1, each inside
 * a fiber; worker w feeds batchers[w % batchers.size()] (one batcher per GPU), which fn should hand to the models it
 * builds: pool.batcher(worker). */
class FiberPool {
    size_t threads_, fibers_, stack_bytes_;
    std::vector<std::shared_ptr<LeafBatcher>> batchers_;

   public:
    FiberPool(size_t threads, size_t fibers_per_thread, std::vector<std::shared_ptr<LeafBatcher>> batchers,
              size_t stack_bytes = size_t(256) << 10)
        : threads_{threads ? threads : 1}, fibers_{fibers_per_thread ? fibers_per_thread : 1}, stack_bytes_{stack_bytes},
          batchers_{std::move(batchers)} {
        if (batchers_.empty()) throw Error(PCU_ERR_INVALID, "FiberPool needs at least one LeafBatcher");
    }

    size_t threads() const { return threads_; }
    size_t capacity() const { return threads_ * fibers_; } /* searches in flight */
    const std::shared_ptr<LeafBatcher> &batcher(size_t worker) const { return batchers_[worker % batchers_.size()]; }

    template <typename Fn>
    void run(size_t n_tasks, Fn fn) {
        std::atomic<size_t> next{0};
        std::vector<std::exception_ptr> errors(threads_);
        auto worker = [&](size_t w) {
            LeafBatcher &mine = *batcher(w);
            FiberScheduler sched;
            detail::tl_scheduler = &sched;
            mine.add_worker();
            try {
                for (size_t f = 0; f < fibers_; ++f)
                    sched.spawn(
                        [&, w] {
                            for (size_t k = next.fetch_add(1); k < n_tasks; k = next.fetch_add(1)) fn(k, w);
                        },
                        stack_bytes_);
                for (;;) {
                    const uint64_t seen = mine.generation();
                    const auto [alive, progressed] = sched.pass();
                    if (!alive) break;
                    if (!progressed) mine.wait_for_flush(seen);
                }
                sched.rethrow();
            } catch (...) {
                errors[w] = std::current_exception();
            }
            mine.remove_worker();
            detail::tl_scheduler = nullptr;
        };
        std::vector<std::thread> pool;
        for (size_t w = 1; w < threads_; ++w) pool.emplace_back(worker, w);
        worker(0);
        for (auto &t : pool) t.join();
        for (auto &e : errors)
            if (e) std::rethrow_exception(e);
    }
};

}  // namespace PinyonCuda
