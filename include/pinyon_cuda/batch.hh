// Structure-of-arrays batch containers of the drop-in model and the host serialisers
// RandomTree::State / MoldState::State -> SoA (north_star (a)).
//
// IsBatchModelTypes (src/model/model.hh:45-79) constrains the batch containers only to be
// constructible from `{1}`; their element types are free, so the reference's
// std::vector<State> (5232 B per RandomTree state, src/model/monte-carlo-model.hh:31) is
// replaced by ~30 bytes per state.
#pragma once

#include <cstdint>
#include <type_traits>
#include <vector>

#include <libpinyon/grow-lib.hh>

#include "device.hh"

namespace PinyonCuda {

template <typename S>
concept IsRandomTreeState = requires(S &s) {
    s.device;
    s.transition_seed;
    s.depth_bound;
    s.rows;
    s.cols;
    s.transitions;
    s.payoff_bias;
    s.chance_threshold;
    s.depth_bound_func;
};

template <typename S>
concept IsMoldState = requires(S &s) { s.max_depth; } && !IsRandomTreeState<S>;

/* rows, cols, transitions and chance_threshold are batch-uniform in the C ABI; a batch
 * mixing shapes is evaluated as consecutive runs of equal shape (order is preserved, so the
 * shared model stream is consumed exactly as monte-carlo-model.hh:58-67 does). */
struct Shape {
    uint32_t rows = 0, cols = 0, transitions = 1;
    int32_t thr_p = 0, thr_q = 1;
    uint32_t growth = 0; /* PCU_GROWTH_* */
    bool operator==(const Shape &) const = default;
};

template <typename Value>
struct BatchInput {
    /* one entry per add_to_batch_input, in order */
    std::vector<int64_t> slot;        /* index into the SoA arrays, or -1 for a terminal state */
    std::vector<Value> terminal_value; /* payoff of terminal entries (rollout is a no-op, :87) */
    std::vector<int64_t> terminal_index;
    /* SoA over non-terminal states */
    std::vector<Shape> shape;
    std::vector<uint64_t> state_seed;
    std::vector<uint32_t> state_pos;
    std::vector<uint64_t> transition_seed;
    std::vector<uint16_t> depth;
    std::vector<int32_t> payoff_bias;
    bool is_mold = false;
    std::vector<Value> mold_value; /* per slot: MoldState never writes its payoff (test-states.hh:43-49) */

    BatchInput() = default;
    explicit BatchInput(size_t reserve_hint) {
        slot.reserve(reserve_hint);
        shape.reserve(reserve_hint);
        depth.reserve(reserve_hint);
    }
    size_t size() const { return slot.size(); }
    size_t device_size() const { return depth.size(); }
    void clear() { *this = BatchInput{}; }
};

template <typename Value>
struct BatchOutput {
    std::vector<uint8_t> payoff_x2; /* per non-terminal state: 2 * row payoff (pinyon_cuda.h pcu_out) */
    std::vector<int64_t> slot;
    std::vector<Value> terminal_value;
    std::vector<int64_t> terminal_index;
    uint64_t sums[8] = {0, 0, 0, 0, 0, 0, 0, 0}; /* wins, draws, losses, steps, hash sum */
    bool is_mold = false;
    std::vector<Value> mold_value;

    BatchOutput() = default;
    explicit BatchOutput(size_t n) { payoff_x2.reserve(n); }
    size_t size() const { return slot.size(); }
    /* exact batch mean of the row value: integer counts, one division */
    double mean_row_value() const {
        const uint64_t n = sums[0] + sums[1] + sums[2];
        return n ? double(2 * sums[0] + sums[1]) / (2.0 * double(n)) : 0.0;
    }
};

/* ---- serialisers ----------------------------------------------------------------------- */

struct TreeImage {
    Shape shape;
    uint64_t state_seed = 0, transition_seed = 0;
    uint32_t state_pos = 0;
    uint16_t depth = 0;
    int32_t payoff_bias = 0;
};

/* RandomTree::State (src/state/random-tree.hh:12-41) -> TreeImage.  Throws Error for
 * states the device path does not reproduce. */
template <IsRandomTreeState State>
TreeImage serialise_tree(const State &s, StreamLocator &locator) {
    using PRNG = std::remove_cvref_t<decltype(s.device)>;
    uint32_t growth = PCU_GROWTH_DEFAULT;
    if (s.actions_func == &GrowLib::alternating_constant_actions<State>)
        growth = PCU_GROWTH_ALTERNATING_CONSTANT; /* the one instantiable grow-lib function (libpinyon/grow-lib.hh:16-24) */
    else if (s.actions_func != &State::actions_default)
        throw Error(PCU_ERR_UNSUPPORTED, "RandomTree actions_func other than the default / GrowLib::alternating_constant_actions");
    if (s.depth_bound_func != &State::depth_bound_default || s.payoff_bias_func != &State::payoff_bias_default)
        throw Error(PCU_ERR_UNSUPPORTED, "RandomTree depth/bias growth functions other than the defaults are host code");

    if (s.chance_denominator != 10) throw Error(PCU_ERR_UNSUPPORTED, "chance_denominator != 10");
    if (s.depth_bound == 0 || s.depth_bound > 65535)
        throw Error(PCU_ERR_INVALID, "depth_bound must be 1..65535 (0 never terminates, random-tree.hh:95-99)");
    TreeImage t;
    t.shape.rows = static_cast<uint32_t>(s.rows);
    t.shape.cols = static_cast<uint32_t>(s.cols);
    t.shape.transitions = static_cast<uint32_t>(s.transitions);
    t.shape.thr_p = static_cast<int32_t>(s.chance_threshold.p);
    t.shape.thr_q = static_cast<int32_t>(s.chance_threshold.q);
    if (t.shape.thr_q < 0) { /* Rational keeps no sign convention */
        t.shape.thr_q = -t.shape.thr_q;
        t.shape.thr_p = -t.shape.thr_p;
    }
    if (t.shape.thr_p < 0) t.shape.thr_p = 0; /* x >= 1/10 > any negative threshold */
    t.shape.growth = growth;
    t.transition_seed = static_cast<uint64_t>(s.transition_seed);
    t.depth = static_cast<uint16_t>(s.depth_bound);
    t.payoff_bias = s.payoff_bias;
    if constexpr (EngineAccess<PRNG>::ENGINE == PCU_ENGINE_XOR_SHIFT) {
        t.state_seed = EngineAccess<PRNG>::state(s.device); /* the whole engine: its current 64-bit state */
        t.state_pos = 0;
        (void)locator;
    } else {
        t.state_seed = EngineAccess<PRNG>::seed(s.device);
        EngineImage img;
        EngineAccess<PRNG>::read(s.device, img);
        const uint64_t pos = locator.locate(t.state_seed, img);
        if (pos == UINT64_MAX || pos > 0xffffffffull)
            throw Error(PCU_ERR_UNSUPPORTED, "state.device is not within 2^32 draws of prng{device.get_seed()}");
        t.state_pos = static_cast<uint32_t>(pos);
    }
    return t;
}

}  // namespace PinyonCuda
