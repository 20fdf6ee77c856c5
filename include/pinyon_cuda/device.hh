// Shared device context of the drop-in model, and the host-side stream locator.
//
// Pinyon models must be cheaply copy-constructible (src/model/readme.md "Resource
// Management"; per-thread copies at src/algorithm/tree-bandit/tree/multithreaded.hh:113,139)
// so the GPU context lives behind a std::shared_ptr, as src/readme.md:202-222 prescribes
// for libtorch models.  One libpinyon_cuda context is single-threaded; the mutex here
// serialises model copies that share it.
#pragma once

#include <array>
#include <cstdint>
#include <cstdlib>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <vector>

#include "../pinyon_cuda.h"
#include "engine_access.hh"

namespace PinyonCuda {

struct Error : std::runtime_error {
    int code;
    Error(int code, const std::string &what) : std::runtime_error(what), code(code) {}
};

/* Finds how many draws an engine has made since `prng{seed}`: RandomTree states carry a
 * COPY of a prng (src/state/random-tree.hh:31-41) advanced by data-dependent discards
 * (:88-89), and std::mt19937 does not record its absolute position.  The locator keeps,
 * per seed, three words of every refilled block of that seed's stream and matches the
 * engine's current block against them. */
class StreamLocator {
    struct PerSeed {
        std::vector<uint32_t> x;                    /* current block of the generator */
        std::vector<std::array<uint32_t, 3>> heads; /* heads[k] = (x[0], x[1], x[623]) after k refills */
        size_t last = 0;                            /* block of the previous hit: searches revisit the same few blocks */
    };
    std::unordered_map<uint32_t, PerSeed> seeds_;
    static constexpr size_t MAX_BLOCKS = size_t(1) << 16; /* 40.9 M draws */
    static constexpr size_t MAX_SEEDS = 256; /* cache bound: a batch of distinct roots must not keep 2.5 KB per root
                                              * forever (a seed seen again after eviction is re-seeded: 623 steps) */

    static void refill(std::vector<uint32_t> &x) {
        for (uint32_t k = 0; k < MT_WORDS; ++k) {
            const uint32_t y = (x[k] & 0x80000000u) | (x[(k + 1) % MT_WORDS] & 0x7fffffffu);
            x[k] = x[(k + 397) % MT_WORDS] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
        }
    }

   public:
    /* Draws made since seeding, or UINT64_MAX if the engine is not on prng{seed}'s stream. */
    uint64_t locate(uint64_t seed, const EngineImage &img) {
        if (seeds_.size() >= MAX_SEEDS && !seeds_.count(static_cast<uint32_t>(seed))) seeds_.clear();
        PerSeed &s = seeds_[static_cast<uint32_t>(seed)];
        if (s.x.empty()) {
            s.x.resize(MT_WORDS);
            s.x[0] = static_cast<uint32_t>(seed); /* mt19937(seed) uses seed mod 2^32 (random.hh:16) */
            for (uint32_t i = 1; i < MT_WORDS; ++i) s.x[i] = 1812433253u * (s.x[i - 1] ^ (s.x[i - 1] >> 30)) + i;
            s.heads.push_back({s.x[0], s.x[1], s.x[MT_WORDS - 1]});
        }
        const std::array<uint32_t, 3> want{img.x[0], img.x[1], img.x[MT_WORDS - 1]};
        auto hit = [&](size_t k) -> uint64_t {
            s.last = k;
            if (k == 0) return img.p == MT_WORDS ? 0 : UINT64_MAX; /* a seeded engine has drawn nothing */
            return uint64_t(MT_WORDS) * (k - 1) + img.p;
        };
        /* the blocks already generated, starting where the last query ended and wrapping once */
        const size_t have = s.heads.size();
        for (size_t i = 0; i < have; ++i) {
            const size_t k = (s.last + i) % have;
            if (s.heads[k] == want) return hit(k);
        }
        for (size_t k = have;; ++k) { /* further along the stream than ever before: generate on */
            if (k >= MAX_BLOCKS) return UINT64_MAX;
            refill(s.x);
            s.heads.push_back({s.x[0], s.x[1], s.x[MT_WORDS - 1]});
            if (s.heads[k] == want) return hit(k);
        }
    }
};

/* One GPU context shared by all copies of a model. */
class Device {
   public:
    pcu_ctx *ctx = nullptr;
    std::mutex mutex;
    StreamLocator locator;

    explicit Device(int device_id = 0) {
        const int rc = pcu_create(device_id, &ctx);
        if (rc != PCU_OK)
            throw Error(rc, "pcu_create(" + std::to_string(device_id) +
                                ") failed: libpinyon_cuda needs an sm_100 GPU and has no CPU fallback");
    }
    /* One context over several GPUs (pcu_create_multi): batched inference is sharded by rollout index inside the
     * library and its counters are reduced with NCCL; single leaves and small batches stay on device_ids[0]. */
    explicit Device(const std::vector<int> &device_ids) {
        const int rc = pcu_create_multi(device_ids.data(), static_cast<int>(device_ids.size()), &ctx);
        if (rc != PCU_OK)
            throw Error(rc, "pcu_create_multi over " + std::to_string(device_ids.size()) +
                                " GPUs failed: libpinyon_cuda needs sm_100 GPUs (and NCCL for more than one); "
                                "there is no CPU fallback");
    }
    Device(const Device &) = delete;
    Device &operator=(const Device &) = delete;
    ~Device() { pcu_destroy(ctx); }

    void check(int rc) const {
        if (rc != PCU_OK) throw Error(rc, std::string("libpinyon_cuda: ") + pcu_last_error(ctx));
    }

    /* Process-wide context over EVERY visible sm_100 GPU (what `Types::Model model{prng{0}}` gets), created on first
     * use.  PINYON_CUDA_DEVICES="0,2,3" restricts the set. */
    static std::shared_ptr<Device> shared_all() {
        static std::mutex mu;
        static std::weak_ptr<Device> cache;
        std::lock_guard<std::mutex> lock(mu);
        std::shared_ptr<Device> d = cache.lock();
        if (!d) {
            std::vector<int> ids;
            if (const char *env = std::getenv("PINYON_CUDA_DEVICES")) {
                for (const char *p = env; *p;) {
                    char *end = nullptr;
                    const long v = std::strtol(p, &end, 10);
                    if (end == p) break;
                    ids.push_back(static_cast<int>(v));
                    p = (*end == ',') ? end + 1 : end;
                }
            } else {
                const int n = pcu_visible_devices();
                for (int i = 0; i < n; ++i) ids.push_back(i);
            }
            if (ids.empty()) ids.push_back(0); /* no usable GPU: let pcu_create report it */
            d = ids.size() == 1 ? std::make_shared<Device>(ids[0]) : std::make_shared<Device>(ids);
            cache = d;
        }
        return d;
    }

    /* Process-wide context per GPU, created on first use. */
    static std::shared_ptr<Device> shared(int device_id = 0) {
        static std::mutex mu;
        static std::unordered_map<int, std::weak_ptr<Device>> cache;
        std::lock_guard<std::mutex> lock(mu);
        std::shared_ptr<Device> d = cache[device_id].lock();
        if (!d) {
            d = std::make_shared<Device>(device_id);
            cache[device_id] = d;
        }
        return d;
    }
};

}  // namespace PinyonCuda
