// MT19937 jump-ahead for the model stream ("mode S at scale").
//
// The reference's batched inference consumes ONE std::mt19937 sequentially
// (monte-carlo-model.hh:58-67).  To let hundreds of CTAs each start in the middle
// of that stream, the engine state must be advanced by D draws in O(1) kernels:
// with phi the characteristic polynomial of the MT19937 transition (degree 19937)
// and g = x^D mod phi,   X[D + j] = XOR_{i : g_i = 1} X[i + j]   for the raw state
// word sequence X.
//   Host side (mt_poly.cpp): phi by Berlekamp–Massey on the engine's own output; the
//     family g_c = x^(c * stride) mod phi, c = 0..count-1, cached per stride
//     (g_2k = g_k^2, g_2k+1 = g_2k * g_1; PCLMULQDQ multiplies).
//   Device side (mt_jump.cu): ONE extension of the base window to 20.6 K words and ONE
//     launch in which CTA c applies g_c to it out of shared memory.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <vector>

constexpr int MTJ_DEG = 19937;
constexpr int MTJ_POLY_WORDS = 624;                  /* 19968 bits >= 19937 coefficients, as u32 */
constexpr int MTJ_GROUPS = 90;                       /* 90 * 224 = 20160 >= 19937 extension words */
constexpr int MTJ_EXT_LEN = 624 + MTJ_GROUPS * 224;  /* words of sequence generated from the base window */

/* True once phi has been recovered (always, unless the host arithmetic is broken). */
bool mt_jump_ready();

/* x^n mod phi applied to a 624-word window on the host (any n).  false on failure. */
bool mt_jump_host(uint32_t *window, uint64_t n);

/* The same through the cached family of `stride`: window advanced by c * stride draws. */
bool mt_jump_host_family(uint32_t *window, uint64_t stride, uint64_t c);

/* Bitmaps (MTJ_POLY_WORDS u32 each, bit i = coefficient of x^i) of x^(c * stride) mod phi
 * for c = first..first+count-1; computed once per stride and cached process-wide (least recently
 * used families are dropped once the cache exceeds its byte cap, 256 MiB by default). */
bool mt_jump_polys(uint64_t stride, uint64_t first, uint64_t count, std::vector<uint32_t> &bitmaps);
void mt_jump_set_host_cache_bytes(size_t bytes);

struct MtJump {
    /* device copies of polynomial families, keyed by stride */
    struct DevFamily {
        uint64_t stride;
        uint64_t first; /* bitmaps of c = first .. first + count - 1 */
        uint64_t count;
        uint32_t *bits;
        uint64_t last_use;
    };
    std::vector<DevFamily> families;
    size_t max_families = 8; /* device copies kept per context (option "jump_families"); evicting one costs a
                              * stream synchronisation now and a 1.5 MB upload when its stride comes back */
    uint64_t tick = 0;
    uint32_t *d_ext = nullptr; /* MTJ_EXT_LEN extended sequence, then the 624-word base window */
    double last_build_ms = 0;  /* host time spent building/uploading a family in the last run (0: cache hit) */

    /* Fill d_ck[j] (j = 0..count-1) with the window preceding draw (first + j) * stride counted from
     * `base_window` (for first == 0, d_ck[0] is base_window itself).  `first` > 0 is how one GPU of a
     * sharded batch takes its slice of the global checkpoint family without any host-side jump.
     * base_window must stay valid until the stream has consumed it (pinned staging memory).
     * Asynchronous on `stream` unless a family has to be built. */
    /* pstep / extra: the fused kernel pairs a forward and a backward stream around one window, so it wants windows at
     * every SECOND stream position only — window j at (first + pstep * j) * stride — plus, when the batch's last
     * stream has an odd index, one more (stored behind the regular ones) at position extra * stride. */
    int run(cudaStream_t stream, int sms, const uint32_t *base_window, uint64_t first, uint64_t count, uint64_t stride,
            uint32_t *d_ck, uint64_t *launches, std::string *err, uint64_t pstep = 1, int64_t extra = -1);
    /* Build and upload the family now (pcu_warmup): the ~80 ms of GF(2) arithmetic a new stride costs. */
    int prepare(cudaStream_t stream, uint64_t first, uint64_t count, uint64_t stride, std::string *err, uint64_t pstep = 1,
                int64_t extra = -1);
    void release();
};
