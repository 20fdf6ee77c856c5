// Host-side access to the std::mt19937 inside Pinyon's `prng` (src/types/random.hh:8-13).
//
// Batched inference in the reference shares ONE model engine sequentially
// (src/model/monte-carlo-model.hh:58-67), so the device path needs the engine's exact
// state (624 words + position) and must hand back the post-batch state.  `prng` keeps
// its engine private and has no accessor.  This header reads and
// writes it without modifying the reference:
//   * the private member `prng::engine` is reached through the explicit-instantiation
//     access rule ([temp.spec]/6: access checking is not applied to the arguments of an
//     explicit instantiation), which is standard C++ and layout independent;
//   * the std::mt19937 state itself is read/written through its standard stream
//     operators (<random>: 624 words followed by the position), or, under libstdc++,
//     by a checked peek at _M_x / _M_p (same values, ~50x faster).
// Other PRNG types plug in by specialising PinyonCuda::EngineAccess.
#pragma once

#include <array>
#include <cstdint>
#include <cstring>
#include <random>
#include <sstream>
#include <stdexcept>
#include <type_traits>

#include <types/random.hh>

namespace PinyonCuda {

constexpr int MT_WORDS = 624;

/* libstdc++-style image of a std::mt19937: x[624] and the index p (0..624) of the next word. */
struct EngineImage {
    std::array<uint32_t, MT_WORDS> x;
    uint32_t p;
};

namespace detail {

template <typename Tag, typename Tag::type Member>
struct PrivateAccess {
    friend typename Tag::type pinyon_cuda_get(Tag) { return Member; }
};
struct PrngEngineTag {
    using type = std::mt19937 prng::*;
    friend type pinyon_cuda_get(PrngEngineTag);
};
template struct PrivateAccess<PrngEngineTag, &prng::engine>;

inline std::mt19937 &engine_of(prng &p) { return p.*pinyon_cuda_get(PrngEngineTag{}); }
inline const std::mt19937 &engine_of(const prng &p) { return p.*pinyon_cuda_get(PrngEngineTag{}); }

/* Portable form: operator<< writes the 624 state words then the position. */
inline void image_from_stream(const std::mt19937 &e, EngineImage &img) {
    std::stringstream ss;
    ss << e;
    for (int i = 0; i < MT_WORDS; ++i) {
        unsigned long long v;
        ss >> v;
        img.x[i] = static_cast<uint32_t>(v);
    }
    unsigned long long p;
    ss >> p;
    img.p = static_cast<uint32_t>(p);
}
inline void image_to_stream(std::mt19937 &e, const EngineImage &img) {
    std::stringstream ss;
    for (int i = 0; i < MT_WORDS; ++i) ss << img.x[i] << ' ';
    ss << img.p;
    ss >> e;
}

#if defined(__GLIBCXX__)
/* libstdc++: `_UIntType _M_x[624]; size_t _M_p;` (bits/random.h), uint_fast32_t words. */
using glibcxx_word = std::mt19937::result_type;
constexpr bool GLIBCXX_PEEK = sizeof(std::mt19937) == MT_WORDS * sizeof(glibcxx_word) + sizeof(size_t);
#else
constexpr bool GLIBCXX_PEEK = false;
#endif

inline void read_engine(const std::mt19937 &e, EngineImage &img) {
    if constexpr (GLIBCXX_PEEK) {
        const unsigned char *raw = reinterpret_cast<const unsigned char *>(&e);
        for (int i = 0; i < MT_WORDS; ++i) {
            std::mt19937::result_type w;
            std::memcpy(&w, raw + sizeof(w) * i, sizeof(w));
            img.x[i] = static_cast<uint32_t>(w);
        }
        size_t p;
        std::memcpy(&p, raw + sizeof(std::mt19937::result_type) * MT_WORDS, sizeof(p));
        img.p = static_cast<uint32_t>(p);
    } else {
        image_from_stream(e, img);
    }
}
inline void write_engine(std::mt19937 &e, const EngineImage &img) {
    if constexpr (GLIBCXX_PEEK) {
        unsigned char *raw = reinterpret_cast<unsigned char *>(&e);
        for (int i = 0; i < MT_WORDS; ++i) {
            const std::mt19937::result_type w = img.x[i];
            std::memcpy(raw + sizeof(w) * i, &w, sizeof(w));
        }
        const size_t p = img.p;
        std::memcpy(raw + sizeof(std::mt19937::result_type) * MT_WORDS, &p, sizeof(p));
    } else {
        image_to_stream(e, img);
    }
}

}  // namespace detail

/* Pinyon's xor_shift (random.hh:66-121) lacks the random_seed() member IsPRNG demands (types.hh:165-190), so no type
 * list can name it as shipped.  This is the class with that one member added — what a user has to write anyway:
 *     using XsTypes = DefaultTypes<double, int, int, double, ConstantSum<1, 1>::Value, std::vector, Matrix, std::mutex,
 *                                  uint64_t, PinyonCuda::xor_shift_prng>;                                          */
struct xor_shift_prng : xor_shift {
    using xor_shift::xor_shift;
    uint64_t random_seed() { return uniform_64(); }
};

namespace detail {
struct XsStateTag {
    using type = uint64_t xor_shift::*;
    friend type pinyon_cuda_get(XsStateTag);
};
template struct PrivateAccess<XsStateTag, &xor_shift::state>;
inline uint64_t &xs_state_of(xor_shift &p) { return p.*pinyon_cuda_get(XsStateTag{}); }
inline const uint64_t &xs_state_of(const xor_shift &p) { return p.*pinyon_cuda_get(XsStateTag{}); }
}  // namespace detail

/* Customisation point: how to get at the engine of a Types::PRNG.  ENGINE is the C ABI's PCU_ENGINE_* value. */
template <typename PRNG>
struct EngineAccess;

template <>
struct EngineAccess<prng> {
    static constexpr uint32_t ENGINE = 0; /* PCU_ENGINE_MT19937 */
    static void read(const prng &device, EngineImage &img) { detail::read_engine(detail::engine_of(device), img); }
    static void write(prng &device, const EngineImage &img) { detail::write_engine(detail::engine_of(device), img); }
    static uint64_t seed(const prng &device) { return device.get_seed(); }
};

/* xor_shift and anything derived from it: the whole engine is one 64-bit word. */
template <typename PRNG>
    requires std::is_base_of_v<xor_shift, PRNG>
struct EngineAccess<PRNG> {
    static constexpr uint32_t ENGINE = 1; /* PCU_ENGINE_XOR_SHIFT */
    static uint64_t state(const PRNG &device) { return detail::xs_state_of(device); }
    static void set_state(PRNG &device, uint64_t s) { detail::xs_state_of(device) = s; }
};

}  // namespace PinyonCuda
