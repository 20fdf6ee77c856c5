// PinyonCuda::MonteCarloModel — drop-in replacement of Pinyon's MonteCarloModel type list
// (src/model/monte-carlo-model.hh:25-98) whose rollouts run on a B200 through the C ABI of
// libpinyon_cuda (include/pinyon_cuda.h).  Header-only C++23, no CUDA types.
//
//   using Types = TreeBandit<Exp3<PinyonCuda::MonteCarloModel<RandomTree<>>>>;
//   Types::Model model{prng{0}};           // same construction as the reference (:39)
//
// Same names, argument meaning and consumption semantics as the reference:
//   Model(const PRNG &)                      :39   public member `device`      :37
//   inference(State &&, ModelOutput &)       :41-56   (a batch of one)
//   inference(ModelBatchInput &, ModelBatchOutput &)  :58-67  one model engine consumed in
//                                            index order; `device` ends where the CPU model's would
//   add_to_batch_input(State &&, ModelBatchInput &) const   :69-74
//   get_output(ModelOutput &, ModelBatchOutput &, long) const :76-82
// The reference reports no errors; this model throws PinyonCuda::Error when the library
// cannot evaluate a state (no GPU, non-default growth functions, depth_bound 0) — there is
// no CPU fallback to hide behind.
#pragma once

#include <algorithm>
#include <memory>
#include <utility>

#include <model/model.hh>
#include <state/state.hh>

#include "batch.hh"

namespace PinyonCuda {

namespace detail {
template <typename Types, bool has_policy>
struct ModelOutputImpl;
template <typename Types>
struct ModelOutputImpl<Types, false> {
    typename Types::Value value;
};
template <typename Types>
struct ModelOutputImpl<Types, true> {
    typename Types::Value value;
    typename Types::VectorReal row_policy, col_policy;
};
}  // namespace detail

template <IsPerfectInfoStateTypes Types, bool has_policy = false>
struct MonteCarloModel : Types {
    using ModelOutput = detail::ModelOutputImpl<Types, has_policy>;
    using ModelBatchInput = BatchInput<typename Types::Value>;
    using ModelBatchOutput = BatchOutput<typename Types::Value>;

    static constexpr bool IS_TREE = IsRandomTreeState<typename Types::State>;
    static constexpr bool IS_MOLD = IsMoldState<typename Types::State>;
    static_assert(IS_TREE || IS_MOLD, "PinyonCuda::MonteCarloModel evaluates RandomTree and MoldState states");

    class Model {
       public:
        typename Types::PRNG device;
        std::shared_ptr<Device> gpu; /* shared by copies: models are copied per thread */

        /* every visible GPU: batched inference shards over them inside the library (pcu_create_multi) */
        Model(const typename Types::PRNG &device) : device{device}, gpu{Device::shared_all()} {}
        Model(const typename Types::PRNG &device, std::shared_ptr<Device> gpu) : device{device}, gpu{std::move(gpu)} {}

        /* 2 * row payoff -> Types::Value exactly as random-tree.hh:99-103 builds it */
        static typename Types::Value value_from_code(uint8_t payoff_x2)
            requires IS_TREE
        {
            typename Types::Q row_payoff{static_cast<int>(payoff_x2), 2};
            row_payoff.canonicalize();
            return typename Types::Value{typename Types::Real{row_payoff}};
        }

        void inference(typename Types::State &&state, ModelOutput &output) {
            fill_policy(state, output);
            if (state.is_terminal()) { /* rollout() does nothing (:87) */
                output.value = state.get_payoff();
                return;
            }
            uint8_t payoff = 0;
            pcu_out out{};
            out.mem = PCU_MEM_HOST;
            out.payoff_x2 = &payoff;
            std::lock_guard<std::mutex> lock(gpu->mutex);
            if constexpr (IS_TREE) {
                const TreeImage t = serialise_tree(state, gpu->locator);
                pcu_rt_batch b{};
                b.n = 1;
                b.mem = PCU_MEM_HOST;
                b.state_seed_u = t.state_seed;
                b.state_pos_u = t.state_pos;
                b.transition_seed_u = t.transition_seed;
                b.depth_u = t.depth;
                b.payoff_bias_u = t.payoff_bias;
                b.rows = t.shape.rows;
                b.cols = t.shape.cols;
                b.transitions = t.shape.transitions;
                b.thr_p = t.shape.thr_p;
                b.thr_q = t.shape.thr_q;
                b.growth = t.shape.growth;
                with_engine([&](pcu_model_rng &rng) { gpu->check(pcu_rollout_random_tree(gpu->ctx, &b, &rng, &out)); });
                output.value = value_from_code(payoff);
            } else {
                pcu_mold_batch b{};
                b.n = 1;
                b.mem = PCU_MEM_HOST;
                b.depth_u = checked_depth(state.max_depth);
                b.n_actions = static_cast<uint32_t>(state.row_actions.size());
                check_square(state);
                with_engine([&](pcu_model_rng &rng) { gpu->check(pcu_rollout_mold(gpu->ctx, &b, &rng, &out)); });
                output.value = state.get_payoff();
            }
        }

        void add_to_batch_input(typename Types::State &&state, ModelBatchInput &input) const {
            const int64_t index = static_cast<int64_t>(input.slot.size());
            if (state.is_terminal()) {
                input.slot.push_back(-1);
                input.terminal_index.push_back(index);
                input.terminal_value.push_back(state.get_payoff());
                return;
            }
            input.slot.push_back(static_cast<int64_t>(input.depth.size()));
            if constexpr (IS_TREE) {
                TreeImage t;
                {
                    std::lock_guard<std::mutex> lock(gpu->mutex);
                    t = serialise_tree(state, gpu->locator);
                }
                input.shape.push_back(t.shape);
                input.state_seed.push_back(t.state_seed);
                input.state_pos.push_back(t.state_pos);
                input.transition_seed.push_back(t.transition_seed);
                input.depth.push_back(t.depth);
                input.payoff_bias.push_back(t.payoff_bias);
            } else {
                check_square(state);
                Shape sh;
                sh.rows = sh.cols = static_cast<uint32_t>(state.row_actions.size());
                input.is_mold = true;
                input.mold_value.push_back(state.get_payoff());
                input.shape.push_back(sh);
                input.depth.push_back(checked_depth(state.max_depth));
            }
        }

        /* Batched inference (monte-carlo-model.hh:58-67): the batch is consumed in index order
         * by this model's one engine; consecutive runs of equal shape are one library call each. */
        void inference(ModelBatchInput &input, ModelBatchOutput &output) {
            const size_t n = input.device_size();
            output.slot = input.slot;
            output.terminal_index = input.terminal_index;
            output.terminal_value = input.terminal_value;
            output.is_mold = input.is_mold;
            output.mold_value = input.mold_value;
            output.payoff_x2.assign(n, 0);
            std::fill(std::begin(output.sums), std::end(output.sums), 0);
            std::lock_guard<std::mutex> lock(gpu->mutex);
            size_t first = 0;
            while (first < n) {
                size_t last = first + 1;
                while (last < n && input.shape[last] == input.shape[first]) ++last;
                run_range(input, output, first, last);
                first = last;
            }
        }

        void get_output(ModelOutput &model_output, ModelBatchOutput &batch_output, const long int index) const {
            const int64_t slot = batch_output.slot[index];
            if (slot < 0) {
                const auto it = std::lower_bound(batch_output.terminal_index.begin(), batch_output.terminal_index.end(),
                                                 static_cast<int64_t>(index));
                model_output.value = batch_output.terminal_value[it - batch_output.terminal_index.begin()];
            } else if constexpr (IS_MOLD) {
                model_output.value = batch_output.mold_value[slot];
            } else {
                model_output.value = value_from_code(batch_output.payoff_x2[slot]);
            }
        }

       protected: /* shared with BatchedMonteCarloModel (batcher.hh) */
        template <typename F>
        void with_engine(F &&call) {
            using Access = EngineAccess<typename Types::PRNG>;
            pcu_model_rng rng{};
            rng.mode = PCU_MODEL_STREAM;
            rng.engine = Access::ENGINE;
            if constexpr (Access::ENGINE == PCU_ENGINE_XOR_SHIFT) {
                uint64_t st = Access::state(device);
                rng.xs_state = &st;
                call(rng);
                Access::set_state(device, st);
            } else {
                EngineImage img;
                Access::read(device, img);
                rng.mt_x = img.x.data();
                rng.mt_p = &img.p;
                call(rng);
                Access::write(device, img); /* `device` continues where the CPU model's would */
            }
        }

        static uint16_t checked_depth(size_t max_depth) {
            if (max_depth == 0 || max_depth > 65535) throw Error(PCU_ERR_INVALID, "MoldState max_depth must be 1..65535");
            return static_cast<uint16_t>(max_depth);
        }
        static void check_square(const typename Types::State &state) {
            if (state.row_actions.size() != state.col_actions.size())
                throw Error(PCU_ERR_UNSUPPORTED, "MoldState with different row and column action counts");
        }

        static void fill_policy(const typename Types::State &state, ModelOutput &output) {
            if constexpr (has_policy) { /* uniform policies, monte-carlo-model.hh:45-53 */
                const size_t rows = state.row_actions.size();
                const size_t cols = state.col_actions.size();
                const typename Types::Real row_uniform{Rational{1, static_cast<int>(rows)}};
                output.row_policy.resize(rows, row_uniform);
                const typename Types::Real col_uniform{Rational{1, static_cast<int>(cols)}};
                output.col_policy.resize(cols, col_uniform);
            }
        }

        template <typename T>
        static bool all_equal(const std::vector<T> &v, size_t first, size_t last) {
            for (size_t i = first + 1; i < last; ++i)
                if (v[i] != v[first]) return false;
            return true;
        }

        void run_range(ModelBatchInput &in, ModelBatchOutput &outp, size_t first, size_t last) {
            const size_t n = last - first;
            uint64_t sums[8] = {0, 0, 0, 0, 0, 0, 0, 0};
            pcu_out out{};
            out.mem = PCU_MEM_HOST;
            out.payoff_x2 = outp.payoff_x2.data() + first;
            out.sums = sums;
            const Shape &sh = in.shape[first];
            if (in.is_mold) {
                pcu_mold_batch b{};
                b.n = n;
                b.mem = PCU_MEM_HOST;
                b.n_actions = sh.rows;
                if (all_equal(in.depth, first, last))
                    b.depth_u = in.depth[first];
                else
                    b.depth = in.depth.data() + first;
                with_engine([&](pcu_model_rng &rng) { gpu->check(pcu_rollout_mold(gpu->ctx, &b, &rng, &out)); });
            } else {
                pcu_rt_batch b{};
                b.n = n;
                b.mem = PCU_MEM_HOST;
                b.rows = sh.rows;
                b.cols = sh.cols;
                b.transitions = sh.transitions;
                b.thr_p = sh.thr_p;
                b.thr_q = sh.thr_q;
                b.growth = sh.growth;
                /* batch-uniform fields travel as scalars: a batch of leaves of one root then takes
                 * the library's shared-stream kernel */
                if (all_equal(in.state_seed, first, last))
                    b.state_seed_u = in.state_seed[first];
                else
                    b.state_seed = in.state_seed.data() + first;
                if (all_equal(in.state_pos, first, last))
                    b.state_pos_u = in.state_pos[first];
                else
                    b.state_pos = in.state_pos.data() + first;
                if (all_equal(in.transition_seed, first, last))
                    b.transition_seed_u = in.transition_seed[first];
                else
                    b.transition_seed = in.transition_seed.data() + first;
                if (all_equal(in.depth, first, last))
                    b.depth_u = in.depth[first];
                else
                    b.depth = in.depth.data() + first;
                if (all_equal(in.payoff_bias, first, last))
                    b.payoff_bias_u = in.payoff_bias[first];
                else
                    b.payoff_bias = in.payoff_bias.data() + first;
                with_engine([&](pcu_model_rng &rng) { gpu->check(pcu_rollout_random_tree(gpu->ctx, &b, &rng, &out)); });
            }
            for (int k = 0; k < 8; ++k) outp.sums[k] += sums[k];
        }
    };
};

}  // namespace PinyonCuda
