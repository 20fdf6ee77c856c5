// libpinyon_cuda — C ABI implementation (include/pinyon_cuda.h) over the sm_100a
// kernels of kernels.cuh.  Host logic only: validation, staging, path selection,
// model-stream checkpoints, launches.  No CPU fallback exists anywhere in this file:
// every rollout is computed by a CUDA kernel or the call fails.
#include "../../include/pinyon_cuda.h"

#include <cuda_runtime.h>

#include <dlfcn.h>
#include <nccl.h> /* types and enums only: the library is dlopen'ed by pcu_create_multi, single-GPU use does not need it */

#include <algorithm>
#include <chrono>
#include <condition_variable>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <functional>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "kernels.cuh"
#include "mt_jump.h"

namespace {

struct DevBuf {
    void *p = nullptr;
    size_t cap = 0;
    cudaError_t ensure(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
        size_t want = bytes + bytes / 4 + 256;
        cudaError_t e = cudaMalloc(&p, want);
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
};

constexpr size_t PIN_OUT_BYTES = 32768, PIN_OUT_HASH = 64, PIN_OUT_TRAJ = 256;
constexpr int PIN_BASE_WINDOW = 1024; /* word offset in pin_small of the staged base window */

enum BufId {
    B_STATE_SEED, B_STATE_POS, B_TSEED, B_DEPTH, B_BIAS, B_MSEED, B_PAYOFF, B_TRAJ, B_HASH, B_SUMS,
    B_CKPT, B_CODES, B_STEPOFF, B_TAB, B_FQ, B_ENDWIN, B_SCRATCH, B_BOUND, B_L397, B_SEEDED, B_PACKED, B_PAYTMP, B_WORDS, B_XSPOW,
    B_COUNT
};

}  // namespace

struct pcu_ctx {
    int device = 0;
    int sms = 0;
    cudaStream_t own_stream = nullptr;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr; /* bracket the rollout kernel of the last call */
    cudaStream_t copy_stream = nullptr;       /* H2D staging of batch inputs, overlapped with the model-stream kernels */
    cudaEvent_t ev_in = nullptr;
    struct PendingCopy {
        void *dst;
        const void *src;
        size_t bytes;
    };
    std::vector<PendingCopy> pending_in; /* stage_in copies not yet enqueued (see inputs_ready) */
    std::vector<uint64_t> host_step_off;  /* mode W, ragged host batches: first step of every rollout */
    bool ev_valid = false;
    std::string err;
    std::string path;
    uint64_t launches = 0;
    int streams_per_sm = 0; /* model streams (generator warps) resident per SM; 0 = by batch size (run_stream) */
    double dense_streams_from = 5e7; /* rollout-steps per call from which 8 streams per SM pay for their checkpoints */
    int force_general = 0;
    int rb = 0;         /* rollouts per consumer batch of the fused kernels (32, 64, 128; 0 = by output, see run_stream) */
    int small_path = 1; /* <= 32 rollouts in mode S from host memory: one k_small launch */
    bool small_attr[2] = {false, false};
    uint8_t *pin_out = nullptr; /* pinned staging for k_small's outputs (PIN_OUT_BYTES) */
    int zero_copy = 1; /* payoffs of large batches go straight to pinned host buffers */
    int lane_full_columns = 0; /* test hook: k_lane keeps every row (no ring) */
    int force_warp = 0; /* test hook: keep distinct-root batches on k_warp instead of k_lane */
    int warp2_lock = 1; /* rollouts a k_warp2 player warp steps in lock-step: 1, 2 or 4 */
    int warp2 = 1;      /* test hook: 0 keeps modes S/W of the general path on k_seed_pre + k_warp instead of k_warp2 */
    int host_checkpoints = 0; /* test hook: brute-force checkpoints on the host instead of the jump kernels */
    DevBuf buf[B_COUNT];
    uint32_t *pin_small = nullptr; /* pinned staging: engine window out */
    bool engine_pending = false;    /* pin_small holds the post-batch engine window */
    /* cached state-stream table */
    struct {
        bool valid = false;
        uint64_t seed = 0;
        uint32_t base = 0, len = 0, T = 0;
        double pu = 0;
        int32_t thr_p = 0, thr_q = 1;
    } tab_key;
    /* cached fq table */
    uint32_t fq_w = 0;
    bool xs_pow_ready = false; /* buf[B_XSPOW] holds the powers of the xor_shift step matrix */
    MtJump jump; /* host polynomials + device images for the model-stream jump-ahead */
    /* statistics of the last rollout call (pcu_get_stat) */
    uint64_t last_chunk = 0, last_streams = 0;
    double last_family_ms = 0, last_table_ms = 0;
    /* set by a multi-GPU parent around the call it hands to this (child) context: the child plays streams
     * [first_stream, first_stream + its share) of ONE global family of model streams */
    struct ShardHint {
        bool active = false;
        uint64_t chunk = 0;        /* rollouts per model stream, chosen for the whole batch */
        uint64_t first_stream = 0; /* global index of this shard's first stream */
        bool last = true;          /* owns the end of the batch: dumps the post-batch engine window */
        bool keep_dev_sums = false; /* accumulate the counters in buf[B_SUMS] for the parent's reduction */
        bool no_sync = false;      /* return after enqueueing; the parent synchronises */
    } hint;
    struct Multi *multi = nullptr; /* non-null: this context is a multi-GPU parent (pcu_create_multi) */
};

namespace {

/* A thread that stays bound to one child context: launches for the GPUs of a sharded call are issued
 * concurrently (a serial loop over eight devices would skew their start by ~0.3 ms). */
class Worker {
    std::thread th;
    std::mutex mu;
    std::condition_variable cv;
    std::function<void()> job;
    bool has = false, quit = false, done = true;

    void loop() {
        std::unique_lock<std::mutex> lk(mu);
        for (;;) {
            cv.wait(lk, [&] { return has || quit; });
            if (quit) return;
            std::function<void()> f = std::move(job);
            has = false;
            lk.unlock();
            f();
            lk.lock();
            done = true;
            cv.notify_all();
        }
    }

   public:
    Worker() : th([this] { loop(); }) {}
    ~Worker() {
        {
            std::lock_guard<std::mutex> lk(mu);
            quit = true;
        }
        cv.notify_all();
        th.join();
    }
    void start(std::function<void()> f) {
        std::lock_guard<std::mutex> lk(mu);
        job = std::move(f);
        has = true;
        done = false;
        cv.notify_all();
    }
    void wait() {
        std::unique_lock<std::mutex> lk(mu);
        cv.wait(lk, [&] { return done; });
    }
};

struct NcclApi {
    void *lib = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t *, int, const int *) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
    bool load() {
        /* by SONAME: a process that already carries an NCCL (torch bundles one) shares it */
        lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!lib) lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (!lib) return false;
        CommInitAll = reinterpret_cast<decltype(CommInitAll)>(dlsym(lib, "ncclCommInitAll"));
        CommDestroy = reinterpret_cast<decltype(CommDestroy)>(dlsym(lib, "ncclCommDestroy"));
        AllReduce = reinterpret_cast<decltype(AllReduce)>(dlsym(lib, "ncclAllReduce"));
        GroupStart = reinterpret_cast<decltype(GroupStart)>(dlsym(lib, "ncclGroupStart"));
        GroupEnd = reinterpret_cast<decltype(GroupEnd)>(dlsym(lib, "ncclGroupEnd"));
        GetErrorString = reinterpret_cast<decltype(GetErrorString)>(dlsym(lib, "ncclGetErrorString"));
        return CommInitAll && CommDestroy && AllReduce && GroupStart && GroupEnd && GetErrorString;
    }
};

}  // namespace

struct Multi {
    std::vector<pcu_ctx *> kids;
    std::vector<std::unique_ptr<Worker>> workers;
    bool same_device = false; /* every child on one GPU (a test configuration): counters summed by k_sum_shards */
    NcclApi nccl;
    std::vector<ncclComm_t> comms;
    std::vector<cudaEvent_t> ev_done; /* same_device: child streams -> child 0's stream */
    uint64_t *pin_sums = nullptr;
    uint64_t min_rollouts = 0; /* batches below this many rollouts go to the first GPU alone */
    uint64_t nccl_calls = 0;
    double t_dispatch_ms = 0, t_reduce_ms = 0, t_sync_ms = 0; /* host wall clock of the last call's phases */
};

namespace {

int fail(pcu_ctx *ctx, int code, const char *fmt, ...) {
    if (ctx) {
        char tmp[512];
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(tmp, sizeof(tmp), fmt, ap);
        va_end(ap);
        ctx->err = tmp;
    }
    return code;
}

#define CU(call)                                                                                      \
    do {                                                                                              \
        cudaError_t e__ = (call);                                                                     \
        if (e__ != cudaSuccess)                                                                       \
            return fail(ctx, PCU_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__),   \
                        __FILE__, __LINE__);                                                          \
    } while (0)

/* ---------------------------------------------------------------- host engine */

void host_mt_seed(uint64_t seed, uint32_t *x, uint32_t *p) {
    x[0] = (uint32_t)seed;
    for (uint32_t i = 1; i < MT_N; ++i) x[i] = mt_init_step(x[i - 1], i);
    *p = MT_N;
}

void host_mt_refill(uint32_t *x) {
    for (uint32_t k = 0; k < MT_N; ++k)
        x[k] = mt_twist(x[k], x[(k + 1) % MT_N], x[(k + MT_M) % MT_N]);
}

/* The 624 raw words preceding the engine's next draw ("window").  libstdc++ keeps a
 * whole refilled block x[0..623] and the position p of the next word, so for p < 624 the
 * window starts 624 - p words BEFORE the block; those words were overwritten by the
 * in-place refill and are recovered by running the recurrence backwards: from
 * v = X[n] ^ X[n-227] = (y >> 1) ^ (lsb(y) ? 0x9908b0df : 0), y = msb(X[n-624]) | low31(X[n-623]),
 * the top bit of v is lsb(y) (the constant has its top bit set, y >> 1 has not). */
void host_window_from_engine(const uint32_t *x, uint32_t p, uint32_t *win) {
    if (p >= MT_N) {
        memcpy(win, x, sizeof(uint32_t) * MT_N);
        return;
    }
    const uint32_t r = MT_N - p; /* words to recover in front of the block */
    std::vector<uint32_t> y(r + MT_N, 0);
    memcpy(y.data() + r, x, sizeof(uint32_t) * MT_N);
    /* y[i] = X[a - r + i]; step n (index into y) yields low31(y[n-623]) and msb(y[n-624]) */
    for (uint32_t n = r + MT_N - 1; n + 1 >= MT_N; --n) { /* n >= 623 */
        uint32_t v = y[n] ^ y[n - MT_LAG];
        uint32_t yy;
        if (v & 0x80000000u)
            yy = ((v ^ 0x9908b0dfu) << 1) | 1u;
        else
            yy = v << 1;
        if (n >= MT_N - 1 && n - (MT_N - 1) < r) y[n - (MT_N - 1)] = (y[n - (MT_N - 1)] & 0x80000000u) | (yy & 0x7fffffffu);
        if (n >= MT_N && n - MT_N < r) y[n - MT_N] = (y[n - MT_N] & 0x7fffffffu) | (yy & 0x80000000u);
        if (n == MT_N - 1) break;
    }
    memcpy(win, y.data(), sizeof(uint32_t) * MT_N);
}

/* Advance a window by `words` draws by plain stepping (O(words)). */
void host_window_advance(uint32_t *win, uint64_t words) {
    std::vector<uint32_t> ring(2 * MT_N);
    memcpy(ring.data(), win, sizeof(uint32_t) * MT_N);
    /* circular buffer of 624: position h is the oldest word */
    uint32_t *x = ring.data();
    uint32_t h = 0;
    for (uint64_t i = 0; i < words; ++i) {
        const uint32_t a = x[h], b = x[(h + 1) % MT_N], c = x[(h + MT_M) % MT_N];
        x[h] = mt_twist(a, b, c);
        h = (h + 1) % MT_N;
    }
    std::vector<uint32_t> out(MT_N);
    for (uint32_t i = 0; i < MT_N; ++i) out[i] = x[(h + i) % MT_N];
    memcpy(win, out.data(), sizeof(uint32_t) * MT_N);
}

/* ------------------------------------------------------- state-stream table */

void host_build_fq(uint32_t T, std::vector<double> &fq, uint32_t *w_out);

/* Two byte tables over draw indices base .. base + len - 1 of prng{seed}, with
 * u64_j = (draw[j] << 32) | draw[j+1] (random.hh:26,28):
 *   tab[j]         = u64_j % 3 - 1 as int8: the payoff-bias step drawn at j (random-tree.hh:97)
 *   tab[len + j]   (T > 1) the chance action apply_actions(row, col) picks (random-tree.hh:115-135)
 *                  for a cell whose T strategy draws start at j: n_k = u64_{j+2k} % 10 + 1,
 *                  thresholded and normalised as get_chance_strategies does (:154-199), then the
 *                  reference's own double arithmetic p -= (double)(float)(n_k / S) from
 *                  p = prng{transition_seed}.uniform().  One entry per POSITION, so a rollout
 *                  step is two table reads whatever T is. */
void host_build_table(uint64_t seed, uint32_t base, uint32_t len, uint32_t T, double pu, int thr_p, int thr_q,
                      std::vector<uint8_t> &tab) {
    std::vector<uint32_t> x(MT_N);
    uint32_t p;
    host_mt_seed(seed, x.data(), &p);
    const uint32_t extra = 2 * T + 2;
    const uint64_t need = (uint64_t)base + len + extra;
    std::vector<uint32_t> w(len + extra);
    uint64_t idx = 0;
    while (idx < need) {
        host_mt_refill(x.data());
        for (uint32_t k = 0; k < MT_N && idx < need; ++k, ++idx)
            if (idx >= base) w[idx - base] = mt_temper(x[k]);
    }
    tab.assign((size_t)len * (T > 1 ? 2 : 1), 0);
    std::vector<uint8_t> n10(len + extra - 1);
    for (uint32_t j = 0; j + 1 < len + extra; ++j) {
        const uint64_t u = ((uint64_t)w[j] << 32) | w[j + 1];
        n10[j] = (uint8_t)(u % 10);
        if (j < len) tab[j] = (uint8_t)(int8_t)((int)(u % 3) - 1);
    }
    if (T > 1) {
        std::vector<double> fq;
        uint32_t fw;
        host_build_fq(T, fq, &fw);
        for (uint32_t j = 0; j < len; ++j) {
            int nk[9], S = 0;
            for (uint32_t k = 0; k < T; ++k) {
                nk[k] = (int)n10[j + 2 * k] + 1;
                if (nk[k] * thr_q < thr_p * 10) nk[k] = 0;
                S += nk[k];
            }
            if (S == 0) {
                nk[0] = 1;
                S = 1;
            }
            volatile double pp = pu;
            uint32_t ch = 0;
            for (uint32_t k = 0; k < T; ++k) {
                pp = pp - fq[(size_t)nk[k] * fw + S];
                if (pp <= 0) {
                    ch = k;
                    break;
                }
            }
            tab[(size_t)len + j] = (uint8_t)ch;
        }
    }
}

double host_uniform_from_seed(uint64_t seed) {
    uint32_t x0 = (uint32_t)seed, v = x0, x1 = 0, x2 = 0, x397 = 0, x398 = 0;
    for (uint32_t i = 1; i <= 398; ++i) {
        v = mt_init_step(v, i);
        if (i == 1) x1 = v;
        if (i == 2) x2 = v;
        if (i == 397) x397 = v;
        if (i == 398) x398 = v;
    }
    const uint32_t o0 = mt_temper(mt_twist(x0, x1, x397));
    const uint32_t o1 = mt_temper(mt_twist(x1, x2, x398));
    return prng_uniform_from_words_host(o0, o1);
}

/* fq[n * w + S] = (double)(float)((double)n / (double)S): the value random-tree.hh:123
 * subtracts — math::to_double returns float (math.hh:30-33) — for a strategy entry
 * n/S (equal to the reference's reduced fraction p/(double)q: one correctly rounded
 * divide of the same rational). */
void host_build_fq(uint32_t T, std::vector<double> &fq, uint32_t *w_out) {
    const uint32_t w = 10 * T + 1;
    fq.assign((size_t)11 * w, 0.0);
    for (uint32_t n = 0; n <= 10; ++n)
        for (uint32_t S = 1; S < w; ++S) {
            volatile double q = (double)n / (double)S;
            volatile float f = (float)q;
            fq[(size_t)n * w + S] = (double)f;
        }
    *w_out = w;
}

template <typename T>
int stage_in(pcu_ctx *ctx, int mem, const T *src, uint64_t n, BufId id, const T **dev_out) {
    if (!src) {
        *dev_out = nullptr;
        return PCU_OK;
    }
    if (mem == PCU_MEM_DEVICE) {
        *dev_out = src;
        return PCU_OK;
    }
    CU(ctx->buf[id].ensure(sizeof(T) * n));
    ctx->pending_in.push_back({ctx->buf[id].p, src, sizeof(T) * n});
    *dev_out = reinterpret_cast<const T *>(ctx->buf[id].p);
    return PCU_OK;
}

/* Enqueue the copies stage_in collected on the copy stream and make the work queued on ctx->stream from here on wait
 * for them.  Called late on purpose where the first kernels need no batch arrays: H2D copies of all streams drain
 * through one hardware queue in issue order, so a 134 MB seed array enqueued first would hold back the 2.5 KB engine
 * window the model-stream kernels start from (measured: no overlap at all). */
int inputs_ready(pcu_ctx *ctx) {
    for (const auto &c : ctx->pending_in)
        CU(cudaMemcpyAsync(c.dst, c.src, c.bytes, cudaMemcpyHostToDevice, ctx->copy_stream));
    ctx->pending_in.clear();
    CU(cudaEventRecord(ctx->ev_in, ctx->copy_stream));
    CU(cudaStreamWaitEvent(ctx->stream, ctx->ev_in, 0));
    return PCU_OK;
}

template <typename T>
int stage_out(pcu_ctx *ctx, int mem, T *dst, uint64_t count, BufId id, T **dev_out) {
    if (!dst) {
        *dev_out = nullptr;
        return PCU_OK;
    }
    if (mem == PCU_MEM_DEVICE) {
        *dev_out = dst;
        return PCU_OK;
    }
    CU(ctx->buf[id].ensure(sizeof(T) * count));
    *dev_out = reinterpret_cast<T *>(ctx->buf[id].p);
    return PCU_OK;
}

template <typename T>
int copy_back(pcu_ctx *ctx, int mem, T *dst, const T *dev, uint64_t count) {
    if (!dst || mem == PCU_MEM_DEVICE) return PCU_OK;
    CU(cudaMemcpyAsync(dst, dev, sizeof(T) * count, cudaMemcpyDeviceToHost, ctx->stream));
    return PCU_OK;
}

struct OutDev {
    bool payoff_direct = false; /* payoff points at the caller's pinned host buffer (zero-copy stores) */
    bool packed_direct = false;
    uint8_t *payoff = nullptr;
    uint8_t *packed = nullptr; /* 2-bit payoffs, four rollouts per byte */
    uint8_t *traj = nullptr;
    uint32_t *hash = nullptr;
    unsigned long long *sums = nullptr;
};

/* Device-visible address of a pinned host buffer, for kernels that store their output straight over PCIe. */
uint8_t *zero_copy_pointer(void *host) {
    cudaPointerAttributes at;
    uint8_t *dev = nullptr;
    if (cudaPointerGetAttributes(&at, host) == cudaSuccess && at.type == cudaMemoryTypeHost && at.devicePointer)
        dev = static_cast<uint8_t *>(at.devicePointer);
    cudaGetLastError(); /* pageable memory reports an error on some drivers: not ours */
    return dev;
}

int prepare_out(pcu_ctx *ctx, const pcu_out *out, uint64_t n, OutDev &od) {
    int rc;
    /* Large batches into PINNED host memory: the kernels store the payoff bytes straight into the caller's
     * buffer over PCIe while they run, instead of a device buffer + a D2H copy after the last rollout. */
    od.payoff_direct = od.packed_direct = false;
    const bool direct_ok = ctx->zero_copy && out->mem == PCU_MEM_HOST && n >= (1ull << 20);
    /* with a packed output the byte payoffs stay in HBM: the packing kernel of the general paths reads them */
    if (direct_ok && out->payoff_x2 && !out->payoff_packed2 && (od.payoff = zero_copy_pointer(out->payoff_x2)))
        od.payoff_direct = true;
    if (direct_ok && out->payoff_packed2 && (od.packed = zero_copy_pointer(out->payoff_packed2))) od.packed_direct = true;
    if (!od.payoff_direct && (rc = stage_out(ctx, out->mem, out->payoff_x2, n, B_PAYOFF, &od.payoff))) return rc;
    if (!od.packed_direct && (rc = stage_out(ctx, out->mem, out->payoff_packed2, (n + 3) / 4, B_PACKED, &od.packed)))
        return rc;
    if ((rc = stage_out(ctx, out->mem, out->traj, n * (uint64_t)out->traj_stride * 3, B_TRAJ, &od.traj)))
        return rc;
    if ((rc = stage_out(ctx, out->mem, out->traj_hash, n, B_HASH, &od.hash))) return rc;
    uint64_t *s = nullptr;
    if (ctx->hint.keep_dev_sums) {
        CU(ctx->buf[B_SUMS].ensure(8 * sizeof(uint64_t)));
        s = reinterpret_cast<uint64_t *>(ctx->buf[B_SUMS].p);
    } else if ((rc = stage_out(ctx, out->mem, out->sums, (uint64_t)8, B_SUMS, &s))) {
        return rc;
    }
    od.sums = reinterpret_cast<unsigned long long *>(s);
    if (od.sums) CU(cudaMemsetAsync(od.sums, 0, 8 * sizeof(uint64_t), ctx->stream));
    return PCU_OK;
}

__global__ void k_pack2(const uint8_t *__restrict__ pay, uint8_t *__restrict__ packed, uint64_t n) {
    const uint64_t q = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; /* output byte = rollouts 4q .. 4q + 3 */
    if (4 * q >= n) return;
    uint32_t v = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j)
        if (4 * q + j < n) v |= (uint32_t)(pay[4 * q + j] & 3u) << (2 * j);
    packed[q] = (uint8_t)v;
}

/* General paths write payoff bytes; a packed output is made from them by one more pass over HBM. */
int pack_source(pcu_ctx *ctx, OutDev &od, uint64_t n, uint8_t **pay) {
    *pay = od.payoff;
    if (od.packed && !od.payoff) {
        CU(ctx->buf[B_PAYTMP].ensure(n));
        *pay = reinterpret_cast<uint8_t *>(ctx->buf[B_PAYTMP].p);
    }
    return PCU_OK;
}
int pack_after(pcu_ctx *ctx, const OutDev &od, uint64_t n, const uint8_t *pay) {
    if (!od.packed || !n) return PCU_OK;
    k_pack2<<<(unsigned)(((n + 3) / 4 + 255) / 256), 256, 0, ctx->stream>>>(pay, od.packed, n);
    ctx->launches++;
    CU(cudaGetLastError());
    return PCU_OK;
}

int finish_out(pcu_ctx *ctx, const pcu_out *out, uint64_t n, const OutDev &od) {
    int rc;
    if (!od.payoff_direct && (rc = copy_back(ctx, out->mem, out->payoff_x2, od.payoff, n))) return rc;
    if (!od.packed_direct && (rc = copy_back(ctx, out->mem, out->payoff_packed2, od.packed, (n + 3) / 4))) return rc;
    if ((rc = copy_back(ctx, out->mem, out->traj, od.traj, n * (uint64_t)out->traj_stride * 3))) return rc;
    if ((rc = copy_back(ctx, out->mem, out->traj_hash, od.hash, n))) return rc;
    if (!ctx->hint.keep_dev_sums &&
        (rc = copy_back(ctx, out->mem, out->sums, reinterpret_cast<const uint64_t *>(od.sums), (uint64_t)8)))
        return rc;
    if (ctx->hint.no_sync) return PCU_OK; /* the multi-GPU parent synchronises every shard after its reduction */
    CU(cudaStreamSynchronize(ctx->stream));
    return PCU_OK;
}

void engine_from_window(const uint32_t *win, uint32_t *mt_x, uint32_t *mt_p);

int finish_all(pcu_ctx *ctx, const pcu_out *out, uint64_t n, const OutDev &od, pcu_model_rng *rng) {
    int rc = finish_out(ctx, out, n, od);
    if (rc) {
        ctx->engine_pending = false;
        return rc;
    }
    if (ctx->hint.no_sync) return PCU_OK; /* engine_pending stays set: the parent reads pin_small after its sync */
    if (ctx->engine_pending) {
        engine_from_window(ctx->pin_small, rng->mt_x, rng->mt_p);
        ctx->engine_pending = false;
    }
    return PCU_OK;
}

/* ------------------------------------------------ model-stream checkpoints */

/* Fill B_CKPT with windows: window j (j < count) holds the 624 raw words preceding model draw number
 * (first + pstep * j) * stride, counted from the engine's next draw; with extra >= 0 one more window, behind them,
 * at draw extra * stride (see MtJump::run). */
int make_checkpoints(pcu_ctx *ctx, const uint32_t *mt_x, uint32_t mt_p, uint64_t first, uint64_t count, uint64_t stride,
                     uint64_t pstep = 1, int64_t extra = -1) {
    uint32_t *win = ctx->pin_small + PIN_BASE_WINDOW; /* pinned: the upload below needs no host sync */
    host_window_from_engine(mt_x, mt_p, win);
    const uint64_t windows = count + (extra >= 0 ? 1 : 0);
    CU(ctx->buf[B_CKPT].ensure(sizeof(uint32_t) * MT_N * (windows + 1)));
    CU(ctx->buf[B_ENDWIN].ensure(sizeof(uint32_t) * MT_N));
    uint32_t *d_ck = reinterpret_cast<uint32_t *>(ctx->buf[B_CKPT].p);
    if (ctx->host_checkpoints && (windows > 1 || first > 0)) {
        /* test hook: plain stepping on the host, the definition the jump kernels must match */
        std::vector<uint32_t> all((size_t)MT_N * windows);
        std::vector<uint32_t> w(win, win + MT_N);
        uint64_t at = 0; /* position of w in units of stride */
        auto advance_to = [&](uint64_t pos) {
            for (; at < pos; ++at) host_window_advance(w.data(), stride);
        };
        for (uint64_t c = 0; c < count; ++c) {
            advance_to(first + pstep * c);
            memcpy(all.data() + (size_t)MT_N * c, w.data(), sizeof(uint32_t) * MT_N);
        }
        if (extra >= 0) {
            if ((uint64_t)extra < at) { /* behind the last regular window: start over */
                w.assign(win, win + MT_N);
                at = 0;
            }
            advance_to((uint64_t)extra);
            memcpy(all.data() + (size_t)MT_N * count, w.data(), sizeof(uint32_t) * MT_N);
        }
        CU(cudaMemcpyAsync(d_ck, all.data(), sizeof(uint32_t) * MT_N * windows, cudaMemcpyHostToDevice, ctx->stream));
        CU(cudaStreamSynchronize(ctx->stream)); /* `all` is pageable and about to go out of scope */
        return PCU_OK;
    }
    std::string jerr;
    if (ctx->jump.run(ctx->stream, ctx->sms, win, first, count, stride, d_ck, &ctx->launches, &jerr, pstep, extra) != 0)
        return fail(ctx, PCU_ERR_CUDA, "jump-ahead failed: %s", jerr.c_str());
    ctx->last_family_ms += ctx->jump.last_build_ms;
    return PCU_OK;
}

void engine_from_window(const uint32_t *win, uint32_t *mt_x, uint32_t *mt_p) {
    /* (window, 624) is an engine state equivalent to libstdc++'s own after the same
     * draws: its next refill computes the same block. */
    memcpy(mt_x, win, sizeof(uint32_t) * MT_N);
    *mt_p = MT_N;
}

int fetch_end_window(pcu_ctx *ctx) {
    CU(cudaMemcpyAsync(ctx->pin_small, ctx->buf[B_ENDWIN].p, sizeof(uint32_t) * MT_N, cudaMemcpyDeviceToHost,
                       ctx->stream));
    ctx->engine_pending = true;
    return PCU_OK;
}

/* Work per model stream (one generator warp).  Every distinct stride needs its own family of
 * jump polynomials (cached, ~80 ms to build), so chunk sizes are quantised to seven significant
 * bits: at most 1/64 of a stream's work is padding (1/16 with five bits cost config A 1.2 %), and the set of
 * strides per depth stays bounded. */
uint64_t pick_chunk(uint64_t units, uint64_t slots, uint64_t min_chunk) {
    uint64_t c = (units + slots - 1) / slots;
    if (c < min_chunk) c = min_chunk;
    if (c < 128) return c;
    int e = 0;
    while ((c >> e) >= 128) ++e; /* c >> e in [64, 128) */
    const uint64_t m = (c + (1ull << e) - 1) >> e;
    return m << e;
}

/* Rollouts per model stream of a uniform-depth batch: a multiple of 128, so that every stream starts on a
 * boundary of the consumers' vector stores (four payoff bytes / one packed byte per lane) and a multi-GPU
 * shard — a whole number of streams — starts on a byte of the packed output. */
uint64_t pick_rollout_chunk(uint64_t n, uint64_t slots, uint64_t min_chunk) {
    const uint64_t c = pick_chunk(n, slots, std::max<uint64_t>(min_chunk, 128));
    return (c + 127) & ~127ull;
}

template <int KIND, typename CodeT>
int launch_stream(pcu_ctx *ctx, const Stream2Params &P, bool tab_smem, size_t smem) {
    auto kern = tab_smem ? k_stream2<KIND, CodeT, true> : k_stream2<KIND, CodeT, false>;
    CU(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const unsigned ctas = (P.streams + GW_PAIRS - 1) / GW_PAIRS;
    CU(cudaEventRecord(ctx->ev0, ctx->stream));
    kern<<<ctas, GW_PAIRS * 64, smem, ctx->stream>>>(P);
    CU(cudaEventRecord(ctx->ev1, ctx->stream));
    ctx->ev_valid = true;
    ctx->launches++;
    CU(cudaGetLastError());
    return PCU_OK;
}

constexpr size_t SMEM_LIMIT = 200 * 1024;

/* Code queue of one stream: a whole number of batches (so a batch never wraps), holding a batch plus the pass being
 * consumed, the pass being generated and the 156 window steps a backward stream starts with. */
uint32_t stream_ring_codes(uint32_t rb, uint32_t depth) {
    const uint64_t b = (uint64_t)rb * depth;
    const uint64_t slots = 2 + (2 * GW_PASS_STEPS + GW_WINDOW_STEPS + b - 1) / b;
    return (uint32_t)std::min<uint64_t>(slots * b, 0xffffffffull);
}

/* Whether the fused kernel's per-stream code queue fits in shared memory; deeper batches take the general path
 * (codes through HBM). */
bool stream_fits(uint32_t depth, uint64_t actions, uint32_t rb = 32) {
    const uint64_t ring = (uint64_t)stream_ring_codes(rb, depth);
    const size_t smem = (size_t)GW_PAIRS * GW_RING * 4 + (size_t)GW_PAIRS * (ring + 16) * (actions <= 256 ? 1 : 2) + 16;
    return smem <= SMEM_LIMIT;
}

/* Model streams (generator warps) resident per SM for a uniform-depth batch of n rollouts on this GPU.  A PAIR of
 * streams costs one jump-ahead checkpoint (~0.4 us of the whole GPU: one runs forwards from it, one backwards) and a
 * second generator warp per scheduler hides the first one's fixed-latency stalls: worth it from ~5e7 rollout-steps
 * per call on (option "dense_streams_from"; tools/sweep_stream.py). */
int streams_per_sm_for(const pcu_ctx *ctx, uint64_t n, uint32_t depth) {
    return ctx->streams_per_sm ? ctx->streams_per_sm : ((double)n * depth >= ctx->dense_streams_from ? 8 : 4);
}

/* Fused path (mode S, uniform depth, shared or absent state stream). */
template <int KIND>
int run_stream(pcu_ctx *ctx, Stream2Params &P, uint32_t rows, pcu_model_rng *rng, const OutDev &od, const pcu_out *out) {
    const uint64_t n = P.n;
    uint64_t first_stream = 0;
    if (ctx->hint.active) { /* one shard of a multi-GPU batch: the parent chose the streams */
        P.chunk = ctx->hint.chunk;
        first_stream = ctx->hint.first_stream;
    } else {
        P.chunk = pick_rollout_chunk(n, (uint64_t)ctx->sms * streams_per_sm_for(ctx, n, P.depth), 128);
    }
    P.streams = (uint32_t)((n + P.chunk - 1) / P.chunk);
    ctx->last_chunk = P.chunk;
    ctx->last_streams = P.streams;
    /* rollouts per consumer batch.  The packed output is one byte per lane: four consecutive rollouts (128).  Byte
     * payoffs stored straight into pinned host memory also want 128 — a warp then writes whole 128-byte lines, i.e.
     * full-size PCIe writes (8 GPUs, r2i: e2e 7.1e10 -> 8.8e10 rollouts/s) — while payoffs that stay in HBM are
     * fastest at 32 (1.181e11 against 1.147e11). */
    P.rb = od.packed ? 128 : (ctx->rb ? ctx->rb : (od.payoff_direct ? 128 : 32));
    if (!stream_fits(P.depth, (uint64_t)rows * P.cols, P.rb)) P.rb = 32;
    P.bcodes = P.rb * P.depth;
    P.ring_codes = stream_ring_codes(P.rb, P.depth);
    if (P.chunk * (uint64_t)P.depth >= (1ull << 32))
        return fail(ctx, PCU_ERR_UNSUPPORTED, "a model stream of %llu x %u steps exceeds 2^32", (unsigned long long)P.chunk, P.depth);
    /* one window per PAIR of streams: even stream 2c runs forwards from it, odd stream 2c - 1 backwards; an odd-indexed
     * last stream runs forwards from a window of its own */
    P.ck_regular = (P.streams + 1) / 2;
    const int64_t extra = (P.streams % 2 == 0) ? (int64_t)(first_stream + P.streams - 1) : -1;
    int rc = make_checkpoints(ctx, rng->mt_x, *rng->mt_p, first_stream, P.ck_regular, 4ull * P.depth * P.chunk, 2, extra);
    if (rc) return rc;
    P.checkpoints = reinterpret_cast<const uint32_t *>(ctx->buf[B_CKPT].p);
    const bool want_end = !ctx->hint.active || ctx->hint.last;
    P.end_window = want_end ? reinterpret_cast<uint32_t *>(ctx->buf[B_ENDWIN].p) : nullptr;
    P.payoff = od.payoff;
    P.payoff_packed2 = od.packed;
    P.traj = od.traj;
    P.traj_hash = od.hash;
    P.sums = od.sums;
    P.vec_ok = (P.chunk % 4 == 0) && ((uintptr_t)od.payoff % 4 == 0) && ((uintptr_t)od.hash % 16 == 0);
    (void)out;
    const bool code8 = (uint64_t)rows * P.cols <= 256;
    const size_t queue_bytes = (size_t)GW_PAIRS * (P.ring_codes + 16) * (code8 ? 1 : 2);
    size_t smem = (size_t)GW_PAIRS * GW_RING * 4 + queue_bytes + 16;
    const size_t tab_bytes = (size_t)P.tab_len * (P.T > 1 ? 2 : 1);
    bool tab_smem = false;
    if (KIND == KIND_TREE && smem + tab_bytes + 16 <= SMEM_LIMIT) {
        tab_smem = true;
        smem += tab_bytes + 16;
    }
    if (smem > SMEM_LIMIT) return fail(ctx, PCU_ERR_UNSUPPORTED, "depth %u needs %zu B of shared memory", P.depth, smem);
    if (code8)
        rc = launch_stream<KIND, uint8_t>(ctx, P, tab_smem, smem);
    else
        rc = launch_stream<KIND, uint16_t>(ctx, P, tab_smem, smem);
    if (rc) return rc;
    return want_end ? fetch_end_window(ctx) : PCU_OK;
}

/* General path helper: joint action codes for every step of a mode-S batch.  `uniform_depth` > 0: every rollout
 * plays that many steps, and the streams are whole numbers of rollouts (so that a multi-GPU shard is a whole
 * number of streams of one global family); 0: ragged depths, streams are cut by steps. */
int gen_codes(pcu_ctx *ctx, uint64_t total_steps, uint32_t rows, uint32_t cols, pcu_model_rng *rng, bool *code16,
              uint32_t uniform_depth = 0, uint64_t n = 0) {
    const bool c16 = (uint64_t)rows * cols > 256;
    *code16 = c16;
    const uint64_t slots = (uint64_t)ctx->sms * (ctx->streams_per_sm ? ctx->streams_per_sm : 4);
    uint64_t chunk, first_stream = 0; /* chunk in steps */
    if (uniform_depth && ctx->hint.active) {
        chunk = ctx->hint.chunk * uniform_depth;
        first_stream = ctx->hint.first_stream;
    } else if (uniform_depth) {
        chunk = pick_rollout_chunk(n, slots, (4096 + uniform_depth - 1) / uniform_depth) * uniform_depth;
    } else {
        chunk = pick_chunk(total_steps ? total_steps : 1, slots, 4096);
    }
    uint64_t streams = (total_steps + chunk - 1) / chunk;
    if (streams == 0) streams = 1;
    ctx->last_chunk = chunk;
    ctx->last_streams = streams;
    int rc = make_checkpoints(ctx, rng->mt_x, *rng->mt_p, first_stream, streams, 4ull * chunk);
    if (rc) return rc;
    CU(ctx->buf[B_CODES].ensure((total_steps + 8) * (c16 ? 2 : 1)));
    const bool want_end = !ctx->hint.active || ctx->hint.last;
    Gen2Params G;
    G.total_steps = total_steps;
    G.chunk_steps = chunk;
    G.streams = (uint32_t)streams;
    G.cols = cols;
    G.mrow = make_fold_const(rows);
    G.mcol = make_fold_const(cols);
    G.checkpoints = reinterpret_cast<const uint32_t *>(ctx->buf[B_CKPT].p);
    G.codes = ctx->buf[B_CODES].p;
    G.end_window = want_end ? reinterpret_cast<uint32_t *>(ctx->buf[B_ENDWIN].p) : nullptr;
    if (total_steps == 0) return PCU_OK; /* engine untouched */
    const unsigned ctas = (unsigned)((streams + 3) / 4);
    if (c16)
        k_gen_codes2<uint16_t><<<ctas, 128, 0, ctx->stream>>>(G);
    else
        k_gen_codes2<uint8_t><<<ctas, 128, 0, ctx->stream>>>(G);
    ctx->launches++;
    CU(cudaGetLastError());
    return want_end ? fetch_end_window(ctx) : PCU_OK;
}

/* Exclusive prefix sum of per-rollout depths -> first step of each rollout. */
__global__ void k_scan_block_sums(const uint16_t *depth, uint64_t n, uint64_t *block_sums) {
    __shared__ unsigned long long sh[32];
    const uint64_t i0 = (uint64_t)blockIdx.x * 4096;
    unsigned long long s = 0;
    for (uint32_t k = threadIdx.x; k < 4096; k += blockDim.x)
        if (i0 + k < n) s += depth[i0 + k];
    s = warp_sum_u64(s);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x < 32) {
        unsigned long long v = threadIdx.x < (blockDim.x >> 5) ? sh[threadIdx.x] : 0;
        v = warp_sum_u64(v);
        if (threadIdx.x == 0) block_sums[blockIdx.x] = v;
    }
}
__global__ void k_scan_finish(const uint16_t *depth, uint64_t n, const uint64_t *block_off, uint64_t *off) {
    /* one warp per 4096-element block: sequential over 128 chunks of 32 */
    const uint64_t i0 = (uint64_t)blockIdx.x * 4096;
    const int lane = threadIdx.x;
    unsigned long long run = block_off[blockIdx.x];
    for (uint32_t k = 0; k < 4096; k += 32) {
        const uint64_t i = i0 + k + lane;
        unsigned long long v = (i < n) ? depth[i] : 0;
        unsigned long long inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            unsigned long long y = __shfl_up_sync(PCU_FULL, inc, o);
            if (lane >= o) inc += y;
        }
        if (i < n) off[i] = run + inc - v;
        run += __shfl_sync(PCU_FULL, inc, 31);
    }
}

int scan_depths(pcu_ctx *ctx, const uint16_t *d_depth, uint64_t n, const uint64_t **d_off, uint64_t *total) {
    const uint64_t blocks = (n + 4095) / 4096;
    CU(ctx->buf[B_STEPOFF].ensure(sizeof(uint64_t) * (n + 1)));
    CU(ctx->buf[B_SCRATCH].ensure(sizeof(uint64_t) * (blocks + 1)));
    uint64_t *off = reinterpret_cast<uint64_t *>(ctx->buf[B_STEPOFF].p);
    uint64_t *bs = reinterpret_cast<uint64_t *>(ctx->buf[B_SCRATCH].p);
    k_scan_block_sums<<<(unsigned)blocks, 256, 0, ctx->stream>>>(d_depth, n, bs);
    ctx->launches++;
    CU(cudaGetLastError());
    std::vector<uint64_t> h(blocks + 1);
    CU(cudaMemcpyAsync(h.data(), bs, sizeof(uint64_t) * blocks, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    uint64_t acc = 0;
    for (uint64_t b = 0; b < blocks; ++b) {
        const uint64_t v = h[b];
        h[b] = acc;
        acc += v;
    }
    *total = acc;
    CU(cudaMemcpyAsync(bs, h.data(), sizeof(uint64_t) * blocks, cudaMemcpyHostToDevice, ctx->stream));
    k_scan_finish<<<(unsigned)blocks, 32, 0, ctx->stream>>>(d_depth, n, bs, off);
    ctx->launches++;
    CU(cudaGetLastError());
    CU(cudaStreamSynchronize(ctx->stream));
    *d_off = off;
    return PCU_OK;
}

/* k_warp over slabs of rollouts: k_seed_pre seeds a slab's private engines lane-per-rollout into an HBM
 * scratch (2496 B per engine), k_warp plays the slab one rollout per warp. */
constexpr uint64_t WARP_SLAB = 1ull << 17;

template <int KIND, int MODE>
int launch_warp(pcu_ctx *ctx, const WarpParams &P0) {
    constexpr int ENGINES = (KIND == KIND_TREE ? 1 : 0) + (MODE == 1 ? 1 : 0);
    const int warps = 8;
    const size_t smem = (size_t)warps * MT_N * 4 * (ENGINES ? ENGINES : 1) + 8 + sizeof(double) * 11 * (P0.T > 1 ? P0.fq_w : 0);
    auto kern = k_warp<KIND, MODE>;
    CU(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const uint64_t slab = ENGINES ? std::min<uint64_t>(P0.n, WARP_SLAB) : P0.n;
    const bool want_pu = KIND == KIND_TREE && P0.T > 1;
    uint32_t *xs = nullptr, *xm = nullptr;
    double *pu = nullptr;
    if (ENGINES) {
        const size_t eng = sizeof(uint32_t) * MT_N * slab;
        CU(ctx->buf[B_SEEDED].ensure(eng * ENGINES + sizeof(double) * slab));
        char *base = reinterpret_cast<char *>(ctx->buf[B_SEEDED].p);
        pu = reinterpret_cast<double *>(base);
        xs = reinterpret_cast<uint32_t *>(base + sizeof(double) * slab);
        xm = (KIND == KIND_TREE) ? xs + MT_N * slab : xs;
    }
    CU(cudaEventRecord(ctx->ev0, ctx->stream));
    for (uint64_t i0 = 0; i0 < P0.n; i0 += slab) {
        WarpParams P = P0;
        P.i0 = i0;
        P.i1 = std::min<uint64_t>(P0.n, i0 + slab);
        const uint64_t cnt = P.i1 - P.i0;
        if (ENGINES) {
            SeedParams S;
            memset(&S, 0, sizeof(S));
            S.i0 = i0;
            S.count = cnt;
            S.state_seed = P0.state_seed;
            S.state_seed_u = P0.state_seed_u;
            S.model_seeds = MODE == 1 ? P0.model_seeds : nullptr;
            S.tseed = P0.tseed;
            S.tseed_u = P0.tseed_u;
            S.want_state = KIND == KIND_TREE;
            S.want_model = MODE == 1;
            S.want_pu = want_pu;
            S.mulc = 1812433253u;
            S.xs = xs;
            S.xm = xm;
            S.pu = pu;
            k_seed_pre<<<(unsigned)((cnt + 127) / 128), 128, 0, ctx->stream>>>(S);
            ctx->launches++;
            P.xs_slab = xs;
            P.xm_slab = xm;
            P.pu_slab = pu;
        }
        uint64_t blocks = std::min<uint64_t>((cnt + warps - 1) / warps, (uint64_t)ctx->sms * 8);
        if (blocks == 0) blocks = 1;
        kern<<<(unsigned)blocks, warps * 32, smem, ctx->stream>>>(P);
        ctx->launches++;
    }
    CU(cudaEventRecord(ctx->ev1, ctx->stream));
    ctx->ev_valid = true;
    CU(cudaGetLastError());
    return PCU_OK;
}

/* k_warp2 (warp2.cuh): one persistent CTA per SM, the private engines seeded and played in shared memory. */
template <int LOCK>
int launch_warp2_lock(pcu_ctx *ctx, const Warp2Params &P, size_t smem, unsigned blocks) {
    auto kern = k_warp2<LOCK>;
    CU(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    CU(cudaEventRecord(ctx->ev0, ctx->stream));
    kern<<<blocks, W2Cfg<LOCK>::THREADS, smem, ctx->stream>>>(P);
    ctx->launches++;
    CU(cudaEventRecord(ctx->ev1, ctx->stream));
    ctx->ev_valid = true;
    CU(cudaGetLastError());
    return PCU_OK;
}

int launch_warp2(pcu_ctx *ctx, const WarpParams &W0) {
    Warp2Params P;
    memset(&P, 0, sizeof(P));
    P.w = W0;
    P.w.i0 = 0;
    P.w.i1 = W0.n;
    P.pu_u = (W0.T > 1 && !W0.tseed) ? host_uniform_from_seed(W0.tseed_u) : 0.0;
    P.mulc = 1812433253u;
    const size_t smem = sizeof(uint32_t) * W2_SLOTS * W2_STRIDE + sizeof(double) * 11 * (W0.T > 1 ? W0.fq_w : 0) + sizeof(W2Ctl);
    const unsigned blocks = (unsigned)std::max<uint64_t>(1, std::min<uint64_t>((W0.n + 31) / 32, (uint64_t)ctx->sms));
    switch (ctx->warp2_lock) {
        case 1: return launch_warp2_lock<1>(ctx, P, smem, blocks);
        case 2: return launch_warp2_lock<2>(ctx, P, smem, blocks);
        default: return launch_warp2_lock<4>(ctx, P, smem, blocks);
    }
}

/* Draw-index bound and largest depth of a batch (k_lane's shared-memory column height).
 * Batch-uniform positions and depths need no kernel. */
int lane_bound(pcu_ctx *ctx, uint64_t n, const uint32_t *d_pos, uint32_t pos_u, const uint16_t *d_depth,
               uint32_t depth_u, uint32_t gen_words, uint32_t per_step, uint64_t *words, uint64_t *dmax,
               uint32_t *pos_min = nullptr, uint32_t *pos_max = nullptr) {
    if (!d_pos && !d_depth) {
        *words = (uint64_t)pos_u - gen_words + (uint64_t)depth_u * per_step;
        *dmax = depth_u;
        if (pos_min) *pos_min = pos_u;
        if (pos_max) *pos_max = pos_u;
        return PCU_OK;
    }
    CU(ctx->buf[B_BOUND].ensure(32));
    unsigned long long *o = reinterpret_cast<unsigned long long *>(ctx->buf[B_BOUND].p);
    CU(cudaMemsetAsync(o, 0, 32, ctx->stream));
    uint64_t blocks = std::min<uint64_t>((n + 255) / 256, (uint64_t)ctx->sms * 8);
    k_lane_bound<<<(unsigned)blocks, 256, 0, ctx->stream>>>(n, d_pos, pos_u, d_depth, depth_u, gen_words, per_step, o);
    ctx->launches++;
    CU(cudaGetLastError());
    unsigned long long h[4];
    CU(cudaMemcpyAsync(h, o, 32, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    *words = h[0];
    *dmax = h[1];
    if (pos_max) *pos_max = (uint32_t)h[2];
    if (pos_min) *pos_min = (uint32_t)(0xffffffffull - h[3]);
    return PCU_OK;
}

/* k_lane: one warp per CTA, 32 rollouts per pass.  Shared memory = (keep + ring) rows x 32 lanes x 4 B:
 * rows [0, keep) are the X[k-227] operands of the rows past 227, the ring holds the window of rows between the
 * lowest draw the next unplayed step can read and the highest one the generation has reached (lane.cuh). */
struct LaneGeom {
    uint32_t words = 0, dmax = 0, pos_min = 0, pos_max = 0;
};

template <int KIND, int MODE>
int launch_lane(pcu_ctx *ctx, const WarpParams &W, const LaneGeom &g) {
    LaneParams LP;
    memset(&LP, 0, sizeof(LP));
    LP.w = W;
    LP.dmax = g.dmax;
    if (KIND == KIND_TREE) {
        const uint32_t rc1 = W.rows * W.cols - 1, gen = W.gen_words;
        const uint32_t base_min = g.pos_min - gen, base_max = g.pos_max - gen;
        LP.words_pad = (g.words + 7) / 8 * 8;
        LP.hstep = gen + rc1 + 2;
        LP.hi0 = base_max + gen + rc1 + 1;
        uint32_t keep = LP.words_pad > MT_LAG ? (LP.words_pad - MT_LAG + 7) / 8 * 8 : 0;
        uint32_t window = 0;
        for (uint32_t s = 0; s < g.dmax; ++s) {
            const uint32_t lo = base_min + (s + 1) * gen + 2 * s, hi = LP.hi0 + s * LP.hstep;
            if (hi >= keep) window = std::max(window, hi - std::max(lo, keep) + 1);
        }
        uint32_t ring = (window + 7 + 7) / 8 * 8;
        if (ctx->lane_full_columns || keep + ring >= LP.words_pad) {
            keep = LP.words_pad;
            ring = 0;
        }
        LP.keep = keep;
        LP.ring = ring;
        LP.ring_m = ring ? (uint32_t)(0x100000000ull / ring) : 0;
    }
    const uint32_t rows = LP.keep + LP.ring;
    LP.mulc = 1812433253u;
    LP.one = 1u;
    CU(ctx->buf[B_L397].ensure(sizeof(uint32_t) * 2 * W.n));
    LP.l397 = reinterpret_cast<uint32_t *>(ctx->buf[B_L397].p);
    LP.lm397 = LP.l397 + W.n;
    const size_t smem = std::max<size_t>((size_t)rows * 32 * 4, 16);
    int per_sm = (int)std::min<size_t>(32, (227 * 1024) / (smem + 1024));
    if (per_sm < 1) per_sm = 1;
    uint64_t blocks = std::min<uint64_t>((W.n + 31) / 32, (uint64_t)ctx->sms * per_sm);
    if (blocks == 0) blocks = 1;
    auto kern = k_lane<KIND, MODE>;
    CU(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    CU(cudaEventRecord(ctx->ev0, ctx->stream));
    k_lane_pre<KIND, MODE><<<(unsigned)((W.n + 255) / 256), 256, 0, ctx->stream>>>(LP);
    ctx->launches++;
    kern<<<(unsigned)blocks, 32, smem, ctx->stream>>>(LP);
    CU(cudaEventRecord(ctx->ev1, ctx->stream));
    ctx->ev_valid = true;
    ctx->launches++;
    CU(cudaGetLastError());
    return PCU_OK;
}

/* <= 32 rollouts in mode S with host buffers: ONE launch whose parameters carry the model engine and the
 * per-rollout fields, outputs written to pinned staging, the host engine advanced on the CPU while the kernel
 * runs.  This is Model::inference(State&&, ModelOutput&) (monte-carlo-model.hh:41-56) as TreeBandit calls it. */
template <int KIND>
int run_small(pcu_ctx *ctx, SmallParams &SP, uint32_t n, pcu_model_rng *rng, const pcu_out *out) {
    if (!ctx->pin_out) CU(cudaMallocHost(reinterpret_cast<void **>(&ctx->pin_out), PIN_OUT_BYTES));
    constexpr size_t ENG = (KIND == KIND_TREE ? 1 : 0) + 1;
    const size_t wstride = MT_N * ENG + 1;
    const bool tab = KIND == KIND_TREE && SP.w.T > 1;
    const size_t smem = ((n * wstride * 4 + 7) & ~(size_t)7) + 32 * sizeof(double) + (tab ? sizeof(double) * 11 * SP.w.fq_w : 0);
    auto kern = k_small<KIND>;
    if (!ctx->small_attr[KIND]) { /* per context: the attribute belongs to this device's copy of the kernel */
        CU(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        ctx->small_attr[KIND] = true;
    }
    SP.x.n = n;
    SP.x.mulc = 1812433253u;
    memcpy(SP.x.win, rng->mt_x, sizeof(uint32_t) * MT_N);
    SP.x.win_p = *rng->mt_p;
    SP.w.n = n;
    SP.w.i0 = 0;
    SP.w.i1 = n;
    SP.w.payoff = ctx->pin_out;
    SP.w.traj_hash = reinterpret_cast<uint32_t *>(ctx->pin_out + PIN_OUT_HASH);
    SP.w.traj = out->traj ? ctx->pin_out + PIN_OUT_TRAJ : nullptr;
    SP.w.traj_stride = out->traj_stride;
    SP.w.sums = nullptr;
    kern<<<1, 32 * n, smem, ctx->stream>>>(SP);
    ctx->launches++;
    ctx->ev_valid = false;
    CU(cudaGetLastError());
    /* the host engine moves on while the kernel runs: 4 draws per step (monte-carlo-model.hh:89-90) */
    uint64_t steps = 0;
    for (uint32_t i = 0; i < n; ++i) steps += SP.x.depth[i];
    /* into a copy: the caller's engine moves only once the kernel is known to have succeeded, as on every other path */
    uint32_t next_x[MT_N], next_p = *rng->mt_p;
    memcpy(next_x, rng->mt_x, sizeof(next_x));
    int rc = pcu_mt_discard(next_x, &next_p, 4 * steps);
    CU(cudaStreamSynchronize(ctx->stream));
    if (rc) return rc;
    memcpy(rng->mt_x, next_x, sizeof(next_x));
    *rng->mt_p = next_p;
    const uint32_t *hash = reinterpret_cast<const uint32_t *>(ctx->pin_out + PIN_OUT_HASH);
    if (out->payoff_x2) memcpy(out->payoff_x2, ctx->pin_out, n);
    if (out->traj_hash) memcpy(out->traj_hash, hash, sizeof(uint32_t) * n);
    if (out->traj) memcpy(out->traj, ctx->pin_out + PIN_OUT_TRAJ, (size_t)n * out->traj_stride * 3);
    if (out->sums) {
        uint64_t sm[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (uint32_t i = 0; i < n; ++i) {
            if (KIND == KIND_TREE)
                sm[2 - ctx->pin_out[i]] += 1; /* payoff_x2 2 = row win, 1 = draw, 0 = row loss */
            else
                sm[1] += 1;
            sm[4] += hash[i];
        }
        sm[3] = steps;
        memcpy(out->sums, sm, sizeof(sm));
    }
    return PCU_OK;
}

int check_common(pcu_ctx *ctx, const void *batch, const pcu_model_rng *rng, const pcu_out *out) {
    if (!ctx) return PCU_ERR_INVALID;
    if (!batch || !rng || !out) return fail(ctx, PCU_ERR_INVALID, "null batch, rng or out");
    if (rng->mode == PCU_MODEL_STREAM && rng->engine != PCU_ENGINE_XOR_SHIFT) {
        if (!rng->mt_x || !rng->mt_p) return fail(ctx, PCU_ERR_INVALID, "mode S needs mt_x and mt_p");
        if (*rng->mt_p > MT_N) return fail(ctx, PCU_ERR_INVALID, "mt_p %u out of range", *rng->mt_p);
    } else if (rng->mode == PCU_MODEL_STREAM) {
        /* xor_shift: checked below */
    } else if (rng->mode == PCU_MODEL_PER_ROLLOUT) {
        if (!rng->seeds) return fail(ctx, PCU_ERR_INVALID, "mode P needs seeds");
    } else if (rng->mode == PCU_MODEL_WORDS) {
        if (!rng->words) return fail(ctx, PCU_ERR_INVALID, "mode W needs words");
        if ((uintptr_t)rng->words % 4) return fail(ctx, PCU_ERR_INVALID, "mode W words must be 4-byte aligned");
    } else {
        return fail(ctx, PCU_ERR_INVALID, "unknown model rng mode %d", rng->mode);
    }
    if (out->mem != PCU_MEM_HOST && out->mem != PCU_MEM_DEVICE) return fail(ctx, PCU_ERR_INVALID, "bad out->mem");
    if (rng->engine == PCU_ENGINE_XOR_SHIFT) {
        if (rng->mode == PCU_MODEL_WORDS) return fail(ctx, PCU_ERR_UNSUPPORTED, "mode W is not offered for xor_shift");
        if (rng->mode == PCU_MODEL_STREAM && !rng->xs_state) return fail(ctx, PCU_ERR_INVALID, "xor_shift mode S needs xs_state");
    } else if (rng->engine != PCU_ENGINE_MT19937) {
        return fail(ctx, PCU_ERR_INVALID, "unknown engine %u", rng->engine);
    }
    return PCU_OK;
}

int upload_fq(pcu_ctx *ctx, uint32_t T, const double **d_fq, uint32_t *fq_w) {
    if (ctx->fq_w != 10 * T + 1) {
        std::vector<double> fq;
        uint32_t w;
        host_build_fq(T, fq, &w);
        CU(ctx->buf[B_FQ].ensure(sizeof(double) * fq.size()));
        CU(cudaMemcpyAsync(ctx->buf[B_FQ].p, fq.data(), sizeof(double) * fq.size(), cudaMemcpyHostToDevice,
                           ctx->stream));
        CU(cudaStreamSynchronize(ctx->stream));
        ctx->fq_w = w;
    }
    *d_fq = reinterpret_cast<const double *>(ctx->buf[B_FQ].p);
    *fq_w = ctx->fq_w;
    return PCU_OK;
}

}  // namespace

/* =============================================================== C ABI ===== */

extern "C" {

int pcu_version(void) { return PCU_VERSION; }

int pcu_create(int device_id, pcu_ctx **out) {
    if (!out) return PCU_ERR_INVALID;
    *out = nullptr;
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || device_id < 0 || device_id >= count) return PCU_ERR_CUDA;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device_id) != cudaSuccess) return PCU_ERR_CUDA;
    if (prop.major != 10) return PCU_ERR_CUDA; /* sm_100a code only */
    if (cudaSetDevice(device_id) != cudaSuccess) return PCU_ERR_CUDA;
    pcu_ctx *ctx = new pcu_ctx();
    ctx->device = device_id;
    ctx->sms = prop.multiProcessorCount;
    if (cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaMallocHost(reinterpret_cast<void **>(&ctx->pin_small), 4096 * sizeof(uint32_t)) != cudaSuccess) {
        delete ctx;
        return PCU_ERR_CUDA;
    }
    ctx->stream = ctx->own_stream;
    if (cudaEventCreate(&ctx->ev0) != cudaSuccess || cudaEventCreate(&ctx->ev1) != cudaSuccess ||
        cudaEventCreateWithFlags(&ctx->ev_in, cudaEventDisableTiming) != cudaSuccess ||
        cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking) != cudaSuccess) {
        delete ctx;
        return PCU_ERR_CUDA;
    }
    *out = ctx;
    return PCU_OK;
}

void pcu_destroy(pcu_ctx *ctx) {
    if (!ctx) return;
    if (ctx->multi) {
        Multi *M = ctx->multi;
        for (pcu_ctx *kid : M->kids) {
            cudaSetDevice(kid->device);
            cudaStreamSynchronize(kid->stream);
        }
        for (ncclComm_t c : M->comms)
            if (c) M->nccl.CommDestroy(c);
        M->workers.clear();
        for (cudaEvent_t e : M->ev_done)
            if (e) cudaEventDestroy(e);
        for (pcu_ctx *kid : M->kids) pcu_destroy(kid);
        if (M->pin_sums) cudaFreeHost(M->pin_sums);
        delete M;
        delete ctx;
        return;
    }
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    for (auto &b : ctx->buf) b.release();
    ctx->jump.release();
    if (ctx->pin_small) cudaFreeHost(ctx->pin_small);
    if (ctx->pin_out) cudaFreeHost(ctx->pin_out);
    if (ctx->ev0) cudaEventDestroy(ctx->ev0);
    if (ctx->ev1) cudaEventDestroy(ctx->ev1);
    if (ctx->ev_in) cudaEventDestroy(ctx->ev_in);
    if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
    if (ctx->own_stream) cudaStreamDestroy(ctx->own_stream);
    delete ctx;
}

const char *pcu_last_error(const pcu_ctx *ctx) { return ctx ? ctx->err.c_str() : "null context"; }
const char *pcu_last_path(const pcu_ctx *ctx) { return ctx ? ctx->path.c_str() : ""; }
uint64_t pcu_kernel_launches(const pcu_ctx *ctx) {
    if (!ctx) return 0;
    uint64_t total = ctx->launches;
    if (ctx->multi)
        for (const pcu_ctx *kid : ctx->multi->kids) total += kid->launches;
    return total;
}

double pcu_last_kernel_ms(pcu_ctx *ctx) {
    if (ctx && ctx->multi) { /* the slowest shard */
        double worst = -1.0;
        for (pcu_ctx *kid : ctx->multi->kids) worst = std::max(worst, pcu_last_kernel_ms(kid));
        return worst;
    }
    if (!ctx || !ctx->ev_valid) return -1.0;
    float ms = 0;
    if (cudaEventSynchronize(ctx->ev1) != cudaSuccess || cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1) != cudaSuccess)
        return -1.0;
    return (double)ms;
}

int pcu_set_stream(pcu_ctx *ctx, void *cuda_stream) {
    if (!ctx) return PCU_ERR_INVALID;
    if (ctx->multi) {
        if (ctx->multi->kids.size() == 1) return pcu_set_stream(ctx->multi->kids[0], cuda_stream);
        return fail(ctx, PCU_ERR_INVALID, "a multi-GPU context runs on its own streams");
    }
    ctx->stream = cuda_stream ? reinterpret_cast<cudaStream_t>(cuda_stream) : ctx->own_stream;
    return PCU_OK;
}

int pcu_set_option(pcu_ctx *ctx, const char *key, int64_t value) {
    if (!ctx || !key) return PCU_ERR_INVALID;
    if (ctx->multi) {
        if (!strcmp(key, "multi_min_rollouts")) {
            if (value < 0) return PCU_ERR_INVALID;
            ctx->multi->min_rollouts = (uint64_t)value;
            return PCU_OK;
        }
        for (pcu_ctx *kid : ctx->multi->kids) {
            const int rc = pcu_set_option(kid, key, value);
            if (rc) {
                ctx->err = kid->err;
                return rc;
            }
        }
        return PCU_OK;
    }
    if (!strcmp(key, "jump_families")) {
        if (value < 1 || value > 1024) return fail(ctx, PCU_ERR_INVALID, "jump_families must be 1..1024");
        ctx->jump.max_families = (size_t)value;
        return PCU_OK;
    }
    if (!strcmp(key, "jump_host_cache_mb")) {
        if (value < 1) return PCU_ERR_INVALID;
        mt_jump_set_host_cache_bytes((size_t)value << 20);
        return PCU_OK;
    }
    if (!strcmp(key, "streams_per_sm") || !strcmp(key, "ctas_per_sm")) { /* second name: r1a spelling */
        if (value < 0 || value > 16) return fail(ctx, PCU_ERR_INVALID, "streams_per_sm must be 0 (auto) or 1..16");
        ctx->streams_per_sm = (int)value;
        return PCU_OK;
    }
    if (!strcmp(key, "force_general")) {
        ctx->force_general = value != 0;
        return PCU_OK;
    }
    if (!strcmp(key, "lane_full_columns")) {
        ctx->lane_full_columns = value != 0;
        return PCU_OK;
    }
    if (!strcmp(key, "rb")) {
        if (value != 0 && value != 32 && value != 64 && value != 128) return PCU_ERR_INVALID;
        ctx->rb = (int)value;
        return PCU_OK;
    }
    if (!strcmp(key, "small_path")) {
        ctx->small_path = value != 0;
        return PCU_OK;
    }
    if (!strcmp(key, "zero_copy")) {
        ctx->zero_copy = value != 0;
        return PCU_OK;
    }
    if (!strcmp(key, "warp2_lock")) {
        if (value != 1 && value != 2 && value != 4) return fail(ctx, PCU_ERR_INVALID, "warp2_lock is 1, 2 or 4");
        ctx->warp2_lock = (int)value;
        return PCU_OK;
    }
    if (!strcmp(key, "warp2")) {
        ctx->warp2 = value != 0;
        return PCU_OK;
    }
    if (!strcmp(key, "force_warp")) {
        ctx->force_warp = value != 0;
        return PCU_OK;
    }
    if (!strcmp(key, "host_checkpoints")) {
        ctx->host_checkpoints = value != 0;
        return PCU_OK;
    }
    return fail(ctx, PCU_ERR_INVALID, "unknown option %s", key);
}

void *pcu_host_alloc(pcu_ctx *ctx, size_t bytes) {
    if (!ctx) return nullptr;
    void *p = nullptr;
    cudaSetDevice(ctx->device);
    /* portable + mapped: every GPU of a multi-GPU context may store its shard's payoffs straight into it */
    if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocPortable | cudaHostAllocMapped) != cudaSuccess) return nullptr;
    return p;
}
void pcu_host_free(pcu_ctx *ctx, void *p) {
    (void)ctx;
    if (p) cudaFreeHost(p);
}

void pcu_mt_seed(uint64_t seed, uint32_t *mt_x, uint32_t *mt_p) { host_mt_seed(seed, mt_x, mt_p); }

uint32_t pcu_mt_next(uint32_t *mt_x, uint32_t *mt_p) {
    if (*mt_p >= MT_N) {
        host_mt_refill(mt_x);
        *mt_p = 0;
    }
    return mt_temper(mt_x[(*mt_p)++]);
}

int pcu_mt_take_raw(uint32_t *mt_x, uint32_t *mt_p, uint64_t n, uint32_t *out) {
    if (!mt_x || !mt_p || *mt_p > MT_N || (n && !out)) return PCU_ERR_INVALID;
    while (n) {
        if (*mt_p >= MT_N) {
            host_mt_refill(mt_x);
            *mt_p = 0;
        }
        uint64_t take = MT_N - *mt_p;
        if (take > n) take = n;
        memcpy(out, mt_x + *mt_p, sizeof(uint32_t) * take);
        out += take;
        *mt_p += (uint32_t)take;
        n -= take;
    }
    return PCU_OK;
}

int pcu_mt_discard(uint32_t *mt_x, uint32_t *mt_p, uint64_t n) {
    if (!mt_x || !mt_p || *mt_p > MT_N) return PCU_ERR_INVALID;
    if (n < (1ull << 22)) {
        while (n) {
            if (*mt_p >= MT_N) {
                host_mt_refill(mt_x);
                *mt_p = 0;
            }
            uint64_t take = MT_N - *mt_p;
            if (take > n) take = n;
            *mt_p += (uint32_t)take;
            n -= take;
        }
        return PCU_OK;
    }
    uint32_t win[MT_N];
    host_window_from_engine(mt_x, *mt_p, win);
    if (!mt_jump_host(win, n)) return PCU_ERR_NOMEM;
    engine_from_window(win, mt_x, mt_p);
    return PCU_OK;
}

int pcu_mt_discard_family(uint32_t *mt_x, uint32_t *mt_p, uint64_t stride, uint64_t c) {
    if (!mt_x || !mt_p || *mt_p > MT_N || c >= 65536) return PCU_ERR_INVALID;
    uint32_t win[MT_N];
    host_window_from_engine(mt_x, *mt_p, win);
    if (!mt_jump_host_family(win, stride, c)) return PCU_ERR_NOMEM;
    engine_from_window(win, mt_x, mt_p);
    return PCU_OK;
}

}  // extern "C"

namespace {

/* Mode W: the batch's raw model words on the device (16-byte aligned for the kernels' uint4 loads) and, for ragged
 * depths, the first step of every rollout.  Host batches: the prefix sum is a host loop over depths the caller holds
 * anyway, uploaded with the words.  Device batches: the scan kernels. */
struct WordsDev {
    const uint32_t *words = nullptr;
    const uint64_t *step_off = nullptr;
    uint64_t total_steps = 0;
    uint32_t dmax = 0;
};

int stage_words(pcu_ctx *ctx, int mem, uint64_t n, const uint16_t *h_depth, const uint16_t *d_depth, uint32_t depth_u,
                const pcu_model_rng *rng, WordsDev &wd) {
    if (mem == PCU_MEM_HOST) {
        wd.dmax = depth_u;
        wd.total_steps = n * (uint64_t)depth_u;
        if (h_depth) {
            CU(ctx->buf[B_STEPOFF].ensure(sizeof(uint64_t) * (n + 1)));
            std::vector<uint64_t> &off = ctx->host_step_off;
            off.resize(n + 1);
            uint64_t acc = 0;
            uint32_t dm = 0;
            for (uint64_t i = 0; i < n; ++i) {
                off[i] = acc;
                acc += h_depth[i];
                dm = std::max<uint32_t>(dm, h_depth[i]);
            }
            off[n] = acc;
            wd.total_steps = acc;
            wd.dmax = dm;
            ctx->pending_in.push_back({ctx->buf[B_STEPOFF].p, off.data(), sizeof(uint64_t) * (n + 1)});
            wd.step_off = reinterpret_cast<const uint64_t *>(ctx->buf[B_STEPOFF].p);
        }
        CU(ctx->buf[B_WORDS].ensure(sizeof(uint32_t) * 4 * wd.total_steps + 16));
        if (wd.total_steps)
            ctx->pending_in.push_back({ctx->buf[B_WORDS].p, rng->words, sizeof(uint32_t) * 4 * wd.total_steps});
        wd.words = reinterpret_cast<const uint32_t *>(ctx->buf[B_WORDS].p);
        return PCU_OK;
    }
    if ((uintptr_t)rng->words % 16) return fail(ctx, PCU_ERR_INVALID, "mode W words in device memory must be 16-byte aligned");
    wd.words = rng->words;
    wd.dmax = depth_u;
    wd.total_steps = n * (uint64_t)depth_u;
    if (d_depth) {
        int rc = scan_depths(ctx, d_depth, n, &wd.step_off, &wd.total_steps);
        if (rc) return rc;
        wd.dmax = 0; /* unknown on the host: the table path is not taken */
    }
    return PCU_OK;
}

int words_to_codes(pcu_ctx *ctx, const WordsDev &wd, uint32_t rows, uint32_t cols, bool *code16) {
    const bool c16 = (uint64_t)rows * cols > 256;
    *code16 = c16;
    CU(ctx->buf[B_CODES].ensure((wd.total_steps + 8) * (c16 ? 2 : 1)));
    if (!wd.total_steps) return PCU_OK;
    const unsigned blocks = (unsigned)((wd.total_steps + 255) / 256);
    const FoldConst mrow = make_fold_const(rows), mcol = make_fold_const(cols);
    if (c16)
        k_words_to_codes<uint16_t><<<blocks, 256, 0, ctx->stream>>>(wd.words, wd.total_steps, cols, mrow, mcol,
                                                                     reinterpret_cast<uint16_t *>(ctx->buf[B_CODES].p));
    else
        k_words_to_codes<uint8_t><<<blocks, 256, 0, ctx->stream>>>(wd.words, wd.total_steps, cols, mrow, mcol,
                                                                    reinterpret_cast<uint8_t *>(ctx->buf[B_CODES].p));
    ctx->launches++;
    CU(cudaGetLastError());
    return PCU_OK;
}

int launch_leaf(pcu_ctx *ctx, const LeafParams &L) {
    CU(cudaEventRecord(ctx->ev0, ctx->stream));
    k_leaf_table<<<(unsigned)((L.n + 127) / 128), 128, 0, ctx->stream>>>(L);
    CU(cudaEventRecord(ctx->ev1, ctx->stream));
    ctx->ev_valid = true;
    ctx->launches++;
    CU(cudaGetLastError());
    return PCU_OK;
}

/* ---- the variant path (thread.cuh): xor_shift type lists and grow-lib growth functions */

/* Columns of A^(2^j), j = 0..63, for the xor_shift step matrix A over GF(2): pw[64 j + b] = A^(2^j) e_b. */
const uint64_t *xs_powers() {
    static uint64_t pw[64 * 64];
    static bool ready = false;
    static std::mutex mu;
    std::lock_guard<std::mutex> lock(mu);
    if (!ready) {
        for (int b = 0; b < 64; ++b) pw[b] = xs_step(1ull << b); /* the step is linear: A e_b */
        for (int j = 1; j < 64; ++j)
            for (int b = 0; b < 64; ++b) { /* A^(2^j) e_b = A^(2^(j-1)) (A^(2^(j-1)) e_b) */
                const uint64_t v = pw[64 * (j - 1) + b];
                uint64_t y = 0;
                for (int c = 0; c < 64; ++c)
                    if ((v >> c) & 1) y ^= pw[64 * (j - 1) + c];
                pw[64 * j + b] = y;
            }
        ready = true;
    }
    return pw;
}

uint64_t host_xs_discard(uint64_t s, uint64_t k) {
    const uint64_t *pw = xs_powers();
    for (int j = 0; k; ++j, k >>= 1) {
        if (!(k & 1)) continue;
        uint64_t y = 0;
        for (int b = 0; b < 64; ++b)
            if ((s >> b) & 1) y ^= pw[64 * j + b];
        s = y;
    }
    return s;
}

/* MT19937 mode S for the variant kernel: the model stream itself, four engine outputs per step, into B_WORDS
 * (the action counts change from step to step, so the generator cannot reduce the draws itself). */
int gen_words(pcu_ctx *ctx, uint64_t total_steps, pcu_model_rng *rng) {
    const uint64_t slots = (uint64_t)ctx->sms * (ctx->streams_per_sm ? ctx->streams_per_sm : 4);
    const uint64_t chunk = pick_chunk(total_steps ? total_steps : 1, slots, 4096);
    uint64_t streams = (total_steps + chunk - 1) / chunk;
    if (streams == 0) streams = 1;
    int rc = make_checkpoints(ctx, rng->mt_x, *rng->mt_p, 0, streams, 4ull * chunk);
    if (rc) return rc;
    CU(ctx->buf[B_WORDS].ensure(16 * (total_steps + 1)));
    Gen2Params G;
    memset(&G, 0, sizeof(G));
    G.total_steps = total_steps;
    G.chunk_steps = chunk;
    G.streams = (uint32_t)streams;
    G.checkpoints = reinterpret_cast<const uint32_t *>(ctx->buf[B_CKPT].p);
    G.codes = ctx->buf[B_WORDS].p;
    G.end_window = reinterpret_cast<uint32_t *>(ctx->buf[B_ENDWIN].p);
    if (total_steps == 0) return PCU_OK;
    k_gen_words<<<(unsigned)((streams + 3) / 4), 128, 0, ctx->stream>>>(G);
    ctx->launches++;
    CU(cudaGetLastError());
    return fetch_end_window(ctx);
}

/* Shared by RandomTree and MoldState batches: everything the fast kernels do not take. */
int run_variant(pcu_ctx *ctx, ThreadParams &P, int mem, const uint16_t *h_depth, const uint16_t *d_depth, pcu_model_rng *rng,
                const pcu_out *out, OutDev &od, const char *path) {
    int rc;
    const uint64_t n = P.n;
    const bool xs = rng->engine == PCU_ENGINE_XOR_SHIFT;
    uint64_t total_steps = n * (uint64_t)P.depth_u;
    if (rng->mode == PCU_MODEL_WORDS) {
        WordsDev wd;
        if ((rc = stage_words(ctx, mem, n, h_depth, d_depth, P.depth_u, rng, wd))) return rc;
        P.model_src = TH_MODEL_WORDS;
        P.words = wd.words;
        P.words_raw = 1;
        P.step_off = wd.step_off;
        if ((rc = inputs_ready(ctx))) return rc;
    } else {
        if ((rc = inputs_ready(ctx))) return rc;
        if (d_depth && rng->mode == PCU_MODEL_STREAM && (rc = scan_depths(ctx, d_depth, n, &P.step_off, &total_steps))) return rc;
        if (rng->mode == PCU_MODEL_PER_ROLLOUT) {
            P.model_src = TH_MODEL_SEEDS;
        } else if (xs) {
            if (!ctx->xs_pow_ready) {
                CU(ctx->buf[B_XSPOW].ensure(sizeof(uint64_t) * 64 * 64));
                CU(cudaMemcpyAsync(ctx->buf[B_XSPOW].p, xs_powers(), sizeof(uint64_t) * 64 * 64, cudaMemcpyHostToDevice, ctx->stream));
                ctx->xs_pow_ready = true;
            }
            P.model_src = TH_MODEL_XS_STREAM;
            P.xs_m0 = *rng->xs_state;
            P.xs_pow = reinterpret_cast<const uint64_t *>(ctx->buf[B_XSPOW].p);
        } else {
            if ((rc = gen_words(ctx, total_steps, rng))) return rc;
            P.model_src = TH_MODEL_WORDS;
            P.words = reinterpret_cast<const uint32_t *>(ctx->buf[B_WORDS].p);
            P.words_raw = 0;
        }
    }
    if ((rc = pack_source(ctx, od, n, &P.payoff))) return rc;
    P.traj = od.traj;
    P.traj_stride = out->traj_stride;
    P.traj_hash = od.hash;
    P.sums = od.sums;
    ctx->path = path;
    CU(cudaEventRecord(ctx->ev0, ctx->stream));
    if (xs)
        k_thread<PCU_ENG_XORSHIFT><<<(unsigned)((n + 127) / 128), 128, 0, ctx->stream>>>(P);
    else
        k_thread<PCU_ENG_MT19937><<<(unsigned)((n + 127) / 128), 128, 0, ctx->stream>>>(P);
    CU(cudaEventRecord(ctx->ev1, ctx->stream));
    ctx->ev_valid = true;
    ctx->launches++;
    CU(cudaGetLastError());
    if ((rc = pack_after(ctx, od, n, P.payoff))) return rc;
    if ((rc = finish_all(ctx, out, n, od, rng))) return rc;
    if (xs && rng->mode == PCU_MODEL_STREAM && !ctx->hint.no_sync) /* two draws per step (monte-carlo-model.hh:89-90) */
        *rng->xs_state = host_xs_discard(*rng->xs_state, 2 * total_steps);
    return PCU_OK;
}

/* One GPU's share of the work: everything below runs on ctx->device. */
int rollout_tree(pcu_ctx *ctx, const pcu_rt_batch *b, pcu_model_rng *rng, pcu_out *out) {
    int rc = check_common(ctx, b, rng, out);
    if (rc) return rc;
    ctx->err.clear();
    ctx->engine_pending = false; /* a call that failed midway must not hand its window to this one */
    ctx->last_chunk = ctx->last_streams = 0;
    ctx->last_family_ms = ctx->last_table_ms = 0;
    if (b->mem != PCU_MEM_HOST && b->mem != PCU_MEM_DEVICE) return fail(ctx, PCU_ERR_INVALID, "bad batch->mem");
    if (b->mem != out->mem) return fail(ctx, PCU_ERR_INVALID, "batch and out must share a memory space");
    if (b->rows < 1 || b->cols < 1 || b->rows > 255 || b->cols > 255)
        return fail(ctx, PCU_ERR_INVALID, "rows/cols must be 1..255");
    if (b->transitions < 1) return fail(ctx, PCU_ERR_INVALID, "transitions must be >= 1");
    if (b->transitions > 8)
        return fail(ctx, PCU_ERR_UNSUPPORTED,
                    "transitions > 8: the reference's Rational<int> sum p/10^T (random-tree.hh:165-184, rational.hh:82-87) "
                    "overflows int once 10^(T-1) * sum(n_k) >= 2^31, so its results are undefined behaviour there");
    if (b->thr_q <= 0 || b->thr_p < 0) return fail(ctx, PCU_ERR_INVALID, "chance_threshold must be p/q with p >= 0, q > 0");
    if (b->growth > PCU_GROWTH_ALTERNATING_CONSTANT) return fail(ctx, PCU_ERR_INVALID, "unknown growth %u", b->growth);
    if (!b->depth && b->depth_u == 0)
        return fail(ctx, PCU_ERR_INVALID, "depth_bound 0 never terminates in the reference (random-tree.hh:95-99)");
    const uint64_t n = b->n;
    const uint32_t T = b->transitions;
    const uint32_t gen_words = 2 * b->rows * b->cols * T;
    const bool is_variant = rng->engine == PCU_ENGINE_XOR_SHIFT || b->growth != PCU_GROWTH_DEFAULT;
    if (rng->engine != PCU_ENGINE_XOR_SHIFT && !b->state_pos && b->state_pos_u < gen_words)
        return fail(ctx, PCU_ERR_INVALID, "state_pos %u < 2*rows*cols*transitions = %u", b->state_pos_u, gen_words);
    if (out->traj && out->traj_stride < (b->depth ? 1u : b->depth_u))
        return fail(ctx, PCU_ERR_INVALID, "traj_stride smaller than depth");
    CU(cudaSetDevice(ctx->device));
    if (ctx->small_path && !ctx->force_general && n >= 1 && n <= 32 && b->mem == PCU_MEM_HOST && !out->payoff_packed2 &&
        !ctx->hint.no_sync && rng->mode == PCU_MODEL_STREAM && !is_variant && (!out->traj || (size_t)n * out->traj_stride * 3 <= PIN_OUT_BYTES - PIN_OUT_TRAJ)) {
        static thread_local SmallParams SP; /* ~3.6 KB: kept off the stack of recursive searches */
        memset(&SP.w, 0, sizeof(SP.w));
        SP.w.rows = b->rows;
        SP.w.cols = b->cols;
        SP.w.T = T;
        SP.w.gen_words = gen_words;
        SP.w.mrow = make_mod_const(b->rows);
        SP.w.mcol = make_mod_const(b->cols);
        SP.w.thr_p = b->thr_p;
        SP.w.thr_q = b->thr_q;
        if (T > 1 && (rc = upload_fq(ctx, T, &SP.w.fq, &SP.w.fq_w))) return rc;
        for (uint32_t i = 0; i < n; ++i) {
            SP.x.state_seed[i] = b->state_seed ? b->state_seed[i] : b->state_seed_u;
            SP.x.tseed[i] = b->transition_seed ? b->transition_seed[i] : b->transition_seed_u;
            SP.x.pos[i] = b->state_pos ? b->state_pos[i] : b->state_pos_u;
            SP.x.bias[i] = b->payoff_bias ? b->payoff_bias[i] : b->payoff_bias_u;
            SP.x.depth[i] = b->depth ? b->depth[i] : b->depth_u;
            if (SP.x.pos[i] < gen_words) return fail(ctx, PCU_ERR_INVALID, "a state_pos is below 2*rows*cols*transitions");
        }
        ctx->path = "small_tree";
        return run_small<KIND_TREE>(ctx, SP, (uint32_t)n, rng, out);
    }
    OutDev od;
    if ((rc = prepare_out(ctx, out, n, od))) return rc;
    if (n == 0) return finish_all(ctx, out, n, od, rng);

    const uint32_t *d_pos;
    const int32_t *d_bias;
    const uint64_t *d_sseed, *d_tseed, *d_mseed = nullptr;
    const uint16_t *d_depth;
    ctx->pending_in.clear();
    if ((rc = stage_in(ctx, b->mem, b->state_pos, n, B_STATE_POS, &d_pos))) return rc;
    if ((rc = stage_in(ctx, b->mem, b->payoff_bias, n, B_BIAS, &d_bias))) return rc;
    if ((rc = stage_in(ctx, b->mem, b->state_seed, n, B_STATE_SEED, &d_sseed))) return rc;
    /* prng{transition_seed} only decides between chance actions (random-tree.hh:118-133): unused when T == 1 */
    if ((rc = stage_in(ctx, b->mem, T > 1 ? b->transition_seed : nullptr, n, B_TSEED, &d_tseed))) return rc;
    if ((rc = stage_in(ctx, b->mem, b->depth, n, B_DEPTH, &d_depth))) return rc;
    if (rng->mode == PCU_MODEL_PER_ROLLOUT)
        if ((rc = stage_in(ctx, b->mem, rng->seeds, n, B_MSEED, &d_mseed))) return rc;
    if (rng->engine == PCU_ENGINE_XOR_SHIFT || b->growth != PCU_GROWTH_DEFAULT) {
        ThreadParams TP;
        memset(&TP, 0, sizeof(TP));
        TP.n = n;
        TP.kind = KIND_TREE;
        TP.rows = b->rows;
        TP.cols = b->cols;
        TP.T = T;
        TP.growth = b->growth;
        TP.thr_p = b->thr_p;
        TP.thr_q = b->thr_q;
        TP.state_seed = d_sseed;
        TP.state_seed_u = b->state_seed_u;
        TP.state_pos = d_pos;
        TP.state_pos_u = b->state_pos_u;
        TP.tseed = d_tseed;
        TP.tseed_u = b->transition_seed_u;
        TP.depth = d_depth;
        TP.depth_u = b->depth_u;
        TP.bias = d_bias;
        TP.bias_u = b->payoff_bias_u;
        TP.model_seeds = d_mseed;
        if (T > 1 && (rc = upload_fq(ctx, T, &TP.fq, &TP.fq_w))) return rc;
        if (b->state_pos && rng->engine != PCU_ENGINE_XOR_SHIFT) { /* positions are range-checked on every path */
            uint64_t lw, ld;
            uint32_t pmin = 0, pmax = 0;
            if ((rc = inputs_ready(ctx))) return rc;
            if ((rc = lane_bound(ctx, n, d_pos, b->state_pos_u, d_depth, b->depth_u, gen_words, 0, &lw, &ld, &pmin, &pmax))) return rc;
            if (pmin < gen_words) return fail(ctx, PCU_ERR_INVALID, "a state_pos is below 2*rows*cols*transitions");
        }
        return run_variant(ctx, TP, b->mem, b->mem == PCU_MEM_HOST ? b->depth : nullptr, d_depth, rng, out, od,
                           rng->engine == PCU_ENGINE_XOR_SHIFT ? "thread_tree_xs" : "thread_tree_growth");
    }
    WordsDev wd;
    if (rng->mode == PCU_MODEL_WORDS && (rc = stage_words(ctx, b->mem, n, b->mem == PCU_MEM_HOST ? b->depth : nullptr,
                                                          d_depth, b->depth_u, rng, wd)))
        return rc;
    /* mode S with uniform depth: the action-code generator needs none of the batch arrays, so the copies above are
     * enqueued behind it and run beside it on the copy stream */
    const bool defer_inputs = rng->mode == PCU_MODEL_STREAM && !b->depth && b->mem == PCU_MEM_HOST;
    if (!defer_inputs && (rc = inputs_ready(ctx))) return rc;

    const double *d_fq = nullptr;
    uint32_t fq_w = 0;
    if (T > 1 && (rc = upload_fq(ctx, T, &d_fq, &fq_w))) return rc;

    /* ---- table paths: one shared state stream, host-visible positions.  Mode S with uniform depth: the fused
     * generator + consumer kernel; mode W (leaf batches of many searches from one root): k_leaf_table */
    const bool leaf = rng->mode == PCU_MODEL_WORDS && (!b->depth || wd.dmax > 0);
    const uint32_t table_depth = leaf ? wd.dmax : b->depth_u;
    bool fused = !ctx->force_general && !b->state_seed && (T == 1 || !b->transition_seed) &&
                 (!b->state_pos || b->mem == PCU_MEM_HOST) &&
                 (leaf || (rng->mode == PCU_MODEL_STREAM && !b->depth && stream_fits(b->depth_u, (uint64_t)b->rows * b->cols)));
    uint32_t pos_min = b->state_pos_u, pos_max = b->state_pos_u;
    if (fused && b->state_pos) {
        pos_min = 0xffffffffu;
        pos_max = 0;
        for (uint64_t i = 0; i < n; ++i) {
            pos_min = std::min(pos_min, b->state_pos[i]);
            pos_max = std::max(pos_max, b->state_pos[i]);
        }
        if (pos_min < gen_words) return fail(ctx, PCU_ERR_INVALID, "a state_pos is below 2*rows*cols*transitions");
    }
    uint64_t span = 0;
    if (fused) {
        const uint64_t per_step = (uint64_t)b->rows * b->cols * T + 2 + gen_words;
        span = (uint64_t)(pos_max - pos_min) + gen_words + per_step * table_depth + 8;
        if (span > (1u << 22)) fused = false;
    }
    if (fused) {
        if (defer_inputs && (rc = inputs_ready(ctx))) return rc;
        uint32_t tab_base = pos_min - gen_words;
        uint32_t tab_len = (uint32_t)span;
        const double pu = (T > 1) ? host_uniform_from_seed(b->transition_seed_u) : 0.0;
        auto &key = ctx->tab_key;
        const bool same_stream = key.valid && key.seed == (uint32_t)b->state_seed_u && key.T == T && key.pu == pu &&
                                 key.thr_p == b->thr_p && key.thr_q == b->thr_q;
        if (same_stream && key.base <= tab_base && (uint64_t)key.base + key.len >= (uint64_t)tab_base + tab_len) {
            /* the cached table COVERS this batch (leaf batches of one search move their position range from call
             * to call): no rebuild, no upload */
            tab_base = key.base;
            tab_len = key.len;
        } else {
            const auto t0 = std::chrono::steady_clock::now();
            if (same_stream) { /* grow to the union, so that a search's table settles after a few batches */
                const uint64_t lo = std::min<uint64_t>(key.base, tab_base);
                const uint64_t hi = std::max<uint64_t>((uint64_t)key.base + key.len, (uint64_t)tab_base + tab_len);
                if (hi - lo <= (1u << 16)) {
                    tab_base = (uint32_t)lo;
                    tab_len = (uint32_t)(hi - lo);
                }
            }
            std::vector<uint8_t> tab;
            host_build_table(b->state_seed_u, tab_base, tab_len, T, pu, b->thr_p, b->thr_q, tab);
            CU(ctx->buf[B_TAB].ensure(tab.size()));
            CU(cudaMemcpyAsync(ctx->buf[B_TAB].p, tab.data(), tab.size(), cudaMemcpyHostToDevice, ctx->stream));
            CU(cudaStreamSynchronize(ctx->stream));
            key.valid = true;
            key.seed = (uint32_t)b->state_seed_u;
            key.base = tab_base;
            key.len = tab_len;
            key.T = T;
            key.pu = pu;
            key.thr_p = b->thr_p;
            key.thr_q = b->thr_q;
            ctx->last_table_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        }
        if (leaf) {
            LeafParams L;
            memset(&L, 0, sizeof(L));
            L.n = n;
            L.kind = KIND_TREE;
            L.cols = b->cols;
            L.T = T;
            L.gen_words = gen_words;
            L.mrow = make_fold_const(b->rows);
            L.mcol = make_fold_const(b->cols);
            L.words = wd.words;
            L.step_off = wd.step_off;
            L.depth = d_depth;
            L.depth_u = b->depth_u;
            L.tab = reinterpret_cast<const uint8_t *>(ctx->buf[B_TAB].p);
            L.tab_base = tab_base;
            L.tab_len = tab_len;
            L.state_pos = d_pos;
            L.state_pos_u = b->state_pos_u;
            L.bias = d_bias;
            L.bias_u = b->payoff_bias_u;
            if ((rc = pack_source(ctx, od, n, &L.payoff))) return rc;
            L.traj = od.traj;
            L.traj_stride = out->traj_stride;
            L.traj_hash = od.hash;
            L.sums = od.sums;
            ctx->path = "leaf_table_W";
            if ((rc = launch_leaf(ctx, L))) return rc;
            if ((rc = pack_after(ctx, od, n, L.payoff))) return rc;
            return finish_all(ctx, out, n, od, rng);
        }
        Stream2Params P;
        memset(&P, 0, sizeof(P));
        P.n = n;
        P.depth = b->depth_u;
        P.cols = b->cols;
        P.T = T;
        P.gen_words = gen_words;
        P.mrow = make_fold_const(b->rows);
        P.mcol = make_fold_const(b->cols);
        P.tab = reinterpret_cast<const uint8_t *>(ctx->buf[B_TAB].p);
        P.tab_base = tab_base;
        P.tab_len = tab_len;
        P.state_pos = d_pos;
        P.state_pos_u = b->state_pos_u;
        P.bias = d_bias;
        P.bias_u = b->payoff_bias_u;
        P.traj_stride = out->traj_stride;
        ctx->path = "stream_table";
        if ((rc = run_stream<KIND_TREE>(ctx, P, b->rows, rng, od, out))) return rc;
        return finish_all(ctx, out, n, od, rng);
    }

    /* ---- general path: one rollout per warp, private state engines */
    WarpParams W;
    memset(&W, 0, sizeof(W));
    W.n = n;
    W.rows = b->rows;
    W.cols = b->cols;
    W.T = T;
    W.gen_words = gen_words;
    W.mrow = make_mod_const(b->rows);
    W.mcol = make_mod_const(b->cols);
    W.state_seed = d_sseed;
    W.state_seed_u = b->state_seed_u;
    W.state_pos = d_pos;
    W.state_pos_u = b->state_pos_u;
    W.tseed = d_tseed;
    W.tseed_u = b->transition_seed_u;
    W.depth = d_depth;
    W.depth_u = b->depth_u;
    W.bias = d_bias;
    W.bias_u = b->payoff_bias_u;
    W.thr_p = b->thr_p;
    W.thr_q = b->thr_q;
    W.fq = d_fq;
    W.fq_w = fq_w;
    if ((rc = pack_source(ctx, od, n, &W.payoff))) return rc;
    W.traj = od.traj;
    W.traj_stride = out->traj_stride;
    W.traj_hash = od.hash;
    W.sums = od.sums;
    /* lane-per-rollout kernel: transitions == 1 (draw positions depend on the action codes only) and every
     * read inside the first refill of the state engine; mode P also needs 4 dmax model words below 227 */
    if (rng->mode == PCU_MODEL_STREAM) {
        uint64_t total_steps = n * (uint64_t)b->depth_u;
        if (d_depth) {
            if ((rc = scan_depths(ctx, d_depth, n, &W.step_off, &total_steps))) return rc;
        }
        bool c16;
        if ((rc = gen_codes(ctx, total_steps, b->rows, b->cols, rng, &c16, d_depth ? 0u : (uint32_t)b->depth_u, n))) return rc;
        W.codes = ctx->buf[B_CODES].p;
        W.code_u16 = c16;
    } else if (rng->mode == PCU_MODEL_WORDS) {
        bool c16;
        if ((rc = words_to_codes(ctx, wd, b->rows, b->cols, &c16))) return rc;
        W.step_off = wd.step_off;
        W.codes = ctx->buf[B_CODES].p;
        W.code_u16 = c16;
    } else {
        W.model_seeds = d_mseed;
    }
    if (defer_inputs && (rc = inputs_ready(ctx))) return rc;
    /* lane-per-rollout kernel: transitions == 1 (draw positions depend on the action codes only) and every
     * read inside the first refill of the state engine; mode P also needs 4 dmax model words below 227 */
    uint64_t lane_words = 0, lane_dmax = 0;
    LaneGeom geom;
    bool lane = !ctx->force_warp && T == 1;
    bool warp2 = ctx->warp2 && rng->mode != PCU_MODEL_PER_ROLLOUT;
    if (lane || d_pos || warp2) { /* per-rollout positions are range-checked on every path (k_warp would index its engine below 0) */
        const uint32_t per_step = gen_words + b->rows * b->cols * T - 1 + 2;
        if ((rc = lane_bound(ctx, n, d_pos, b->state_pos_u, d_depth, b->depth_u, gen_words, per_step, &lane_words,
                             &lane_dmax, &geom.pos_min, &geom.pos_max)))
            return rc;
        if (geom.pos_min < gen_words) return fail(ctx, PCU_ERR_INVALID, "a state_pos is below 2*rows*cols*transitions");
        lane = lane && lane_words <= MT_N - 1 && (rng->mode != PCU_MODEL_PER_ROLLOUT || 4 * lane_dmax <= MT_LAG - 3);
        geom.words = (uint32_t)lane_words;
        geom.dmax = (uint32_t)lane_dmax;
        warp2 = warp2 && lane_words + 16 < (1ull << 31); /* k_warp2 keeps 32-bit draw indices */
    }
    if (rng->mode != PCU_MODEL_PER_ROLLOUT) { /* modes S and W: the kernels take joint action codes */
        const bool w = rng->mode == PCU_MODEL_WORDS;
        ctx->path = lane ? (w ? "lane_private_W" : "lane_private_S") : (w ? "warp_private_W" : "warp_private_S");
        rc = lane ? launch_lane<KIND_TREE, 0>(ctx, W, geom) : warp2 ? launch_warp2(ctx, W) : launch_warp<KIND_TREE, 0>(ctx, W);
    } else {
        ctx->path = lane ? "lane_private_P" : "warp_private_P";
        rc = lane ? launch_lane<KIND_TREE, 1>(ctx, W, geom) : launch_warp<KIND_TREE, 1>(ctx, W);
    }
    if (rc) return rc;
    if ((rc = pack_after(ctx, od, n, W.payoff))) return rc;
    return finish_all(ctx, out, n, od, rng);
}

int rollout_mold(pcu_ctx *ctx, const pcu_mold_batch *b, pcu_model_rng *rng, pcu_out *out) {
    int rc = check_common(ctx, b, rng, out);
    if (rc) return rc;
    ctx->err.clear();
    ctx->engine_pending = false; /* a call that failed midway must not hand its window to this one */
    ctx->last_chunk = ctx->last_streams = 0;
    ctx->last_family_ms = ctx->last_table_ms = 0;
    if (b->mem != PCU_MEM_HOST && b->mem != PCU_MEM_DEVICE) return fail(ctx, PCU_ERR_INVALID, "bad batch->mem");
    if (b->mem != out->mem) return fail(ctx, PCU_ERR_INVALID, "batch and out must share a memory space");
    if (b->n_actions < 1 || b->n_actions > 255) return fail(ctx, PCU_ERR_INVALID, "n_actions must be 1..255");
    if (out->traj && out->traj_stride < (b->depth ? 1u : b->depth_u))
        return fail(ctx, PCU_ERR_INVALID, "traj_stride smaller than depth");
    const uint64_t n = b->n;
    const bool is_variant = rng->engine == PCU_ENGINE_XOR_SHIFT;
    CU(cudaSetDevice(ctx->device));
    if (ctx->small_path && !ctx->force_general && n >= 1 && n <= 32 && b->mem == PCU_MEM_HOST && !out->payoff_packed2 &&
        !ctx->hint.no_sync && rng->mode == PCU_MODEL_STREAM && !is_variant && (!out->traj || (size_t)n * out->traj_stride * 3 <= PIN_OUT_BYTES - PIN_OUT_TRAJ)) {
        static thread_local SmallParams SP;
        memset(&SP.w, 0, sizeof(SP.w));
        SP.w.rows = SP.w.cols = b->n_actions;
        SP.w.T = 1;
        SP.w.mrow = make_mod_const(b->n_actions);
        SP.w.mcol = make_mod_const(b->n_actions);
        for (uint32_t i = 0; i < n; ++i) SP.x.depth[i] = b->depth ? b->depth[i] : b->depth_u;
        ctx->path = "small_mold";
        return run_small<KIND_MOLD>(ctx, SP, (uint32_t)n, rng, out);
    }
    OutDev od;
    if ((rc = prepare_out(ctx, out, n, od))) return rc;
    if (n == 0) return finish_all(ctx, out, n, od, rng);
    const uint16_t *d_depth;
    const uint64_t *d_mseed = nullptr;
    ctx->pending_in.clear();
    if ((rc = stage_in(ctx, b->mem, b->depth, n, B_DEPTH, &d_depth))) return rc;
    if (rng->mode == PCU_MODEL_PER_ROLLOUT)
        if ((rc = stage_in(ctx, b->mem, rng->seeds, n, B_MSEED, &d_mseed))) return rc;
    if (is_variant) {
        ThreadParams TP;
        memset(&TP, 0, sizeof(TP));
        TP.n = n;
        TP.kind = KIND_MOLD;
        TP.rows = TP.cols = b->n_actions;
        TP.T = 1;
        TP.depth = d_depth;
        TP.depth_u = b->depth_u;
        TP.model_seeds = d_mseed;
        return run_variant(ctx, TP, b->mem, b->mem == PCU_MEM_HOST ? b->depth : nullptr, d_depth, rng, out, od, "thread_mold_xs");
    }
    WordsDev wd;
    if (rng->mode == PCU_MODEL_WORDS && (rc = stage_words(ctx, b->mem, n, b->mem == PCU_MEM_HOST ? b->depth : nullptr,
                                                          d_depth, b->depth_u, rng, wd)))
        return rc;
    if ((rc = inputs_ready(ctx))) return rc;
    if (rng->mode == PCU_MODEL_WORDS) { /* no state stream: one thread per rollout folds its words */
        LeafParams L;
        memset(&L, 0, sizeof(L));
        L.n = n;
        L.kind = KIND_MOLD;
        L.cols = b->n_actions;
        L.T = 1;
        L.mrow = make_fold_const(b->n_actions);
        L.mcol = make_fold_const(b->n_actions);
        L.words = wd.words;
        L.step_off = wd.step_off;
        L.depth = d_depth;
        L.depth_u = b->depth_u;
        if ((rc = pack_source(ctx, od, n, &L.payoff))) return rc;
        L.traj = od.traj;
        L.traj_stride = out->traj_stride;
        L.traj_hash = od.hash;
        L.sums = od.sums;
        ctx->path = "leaf_mold_W";
        if ((rc = launch_leaf(ctx, L))) return rc;
        if ((rc = pack_after(ctx, od, n, L.payoff))) return rc;
        return finish_all(ctx, out, n, od, rng);
    }

    if (!ctx->force_general && rng->mode == PCU_MODEL_STREAM && !b->depth && b->depth_u > 0 &&
        stream_fits(b->depth_u, (uint64_t)b->n_actions * b->n_actions)) {
        Stream2Params P;
        memset(&P, 0, sizeof(P));
        P.n = n;
        P.depth = b->depth_u;
        P.cols = b->n_actions;
        P.T = 1;
        P.mrow = make_fold_const(b->n_actions);
        P.mcol = make_fold_const(b->n_actions);
        P.traj_stride = out->traj_stride;
        ctx->path = "stream_mold";
        if ((rc = run_stream<KIND_MOLD>(ctx, P, b->n_actions, rng, od, out))) return rc;
        return finish_all(ctx, out, n, od, rng);
    }
    WarpParams W;
    memset(&W, 0, sizeof(W));
    W.n = n;
    W.rows = b->n_actions;
    W.cols = b->n_actions;
    W.T = 1;
    W.mrow = make_mod_const(b->n_actions);
    W.mcol = make_mod_const(b->n_actions);
    W.depth = d_depth;
    W.depth_u = b->depth_u;
    if ((rc = pack_source(ctx, od, n, &W.payoff))) return rc;
    W.traj = od.traj;
    W.traj_stride = out->traj_stride;
    W.traj_hash = od.hash;
    W.sums = od.sums;
    if (rng->mode == PCU_MODEL_STREAM) {
        uint64_t total_steps = n * (uint64_t)b->depth_u;
        if (d_depth) {
            if ((rc = scan_depths(ctx, d_depth, n, &W.step_off, &total_steps))) return rc;
        }
        bool c16;
        if ((rc = gen_codes(ctx, total_steps, b->n_actions, b->n_actions, rng, &c16, d_depth ? 0u : (uint32_t)b->depth_u, n)))
            return rc;
        W.codes = ctx->buf[B_CODES].p;
        W.code_u16 = c16;
        ctx->path = "warp_mold_S";
        if ((rc = launch_warp<KIND_MOLD, 0>(ctx, W))) return rc;
    } else {
        W.model_seeds = d_mseed;
        uint64_t lw = 0, dmax = 0;
        if ((rc = lane_bound(ctx, n, nullptr, 0, d_depth, b->depth_u, 0, 0, &lw, &dmax))) return rc;
        const bool lane = !ctx->force_warp && 4 * dmax <= MT_LAG - 3;
        ctx->path = lane ? "lane_mold_P" : "warp_mold_P";
        LaneGeom geom;
        geom.dmax = (uint32_t)dmax;
        rc = lane ? launch_lane<KIND_MOLD, 1>(ctx, W, geom) : launch_warp<KIND_MOLD, 1>(ctx, W);
        if (rc) return rc;
    }
    if ((rc = pack_after(ctx, od, n, W.payoff))) return rc;
    return finish_all(ctx, out, n, od, rng);
}

/* ------------------------------------------------------------------ multi-GPU parent (pcu_create_multi) */

}  // namespace

namespace {

struct SumPtrs {
    const unsigned long long *p[16];
    int n;
};
__global__ void k_sum_shards(SumPtrs S, unsigned long long *out) {
    const int t = threadIdx.x;
    if (t < 8) {
        unsigned long long v = 0;
        for (int g = 0; g < S.n; ++g) v += S.p[g][t];
        out[t] = v;
    }
}

struct ShardPlan {
    uint64_t first = 0, count = 0;   /* rollouts [first, first + count) */
    uint64_t first_stream = 0;       /* uniform depth, mode S: global index of the shard's first model stream */
    uint64_t words_before = 0;       /* ragged depth, mode S: model draws consumed by the rollouts before `first` */
    bool last = false;
};

/* Cut a batch into contiguous rollout-index shards.  Uniform depth in mode S: the batch is ONE
 * family of model streams of `chunk` rollouts each (stride 4 * depth * chunk draws) and a shard is a whole number of
 * them, so GPU g applies polynomials x^((g * spg + j) * stride) to the caller's engine window itself — no host-side
 * jump per shard.  Ragged depth: equal rollout counts; the shard's engine is the caller's advanced on the host by
 * 4 * (steps of the rollouts before it).  Mode P: equal rollout counts, nothing else. */
void plan_shards(const Multi &M, uint64_t n, int mode, const uint16_t *depth, uint32_t depth_u, uint64_t *chunk_out,
                 std::vector<ShardPlan> &plan) {
    const size_t G = M.kids.size();
    plan.assign(G, ShardPlan{});
    *chunk_out = 0;
    if (mode == PCU_MODEL_STREAM && !depth) {
        const pcu_ctx *k0 = M.kids[0];
        uint64_t slots = 0;
        for (const pcu_ctx *k : M.kids) slots += (uint64_t)k->sms * streams_per_sm_for(k0, n / G, depth_u);
        const uint64_t min_chunk = std::max<uint64_t>(128, (4096 + depth_u - 1) / std::max<uint32_t>(depth_u, 1));
        const uint64_t chunk = pick_rollout_chunk(n, slots, min_chunk);
        const uint64_t S = (n + chunk - 1) / chunk;
        uint64_t spg = (S + G - 1) / G;
        spg += spg & 1; /* whole forward/backward stream pairs per GPU */
        *chunk_out = chunk;
        for (size_t g = 0; g < G; ++g) {
            const uint64_t s0 = std::min<uint64_t>(S, g * spg), s1 = std::min<uint64_t>(S, s0 + spg);
            plan[g].first_stream = s0;
            plan[g].first = std::min<uint64_t>(n, s0 * chunk);
            plan[g].count = std::min<uint64_t>(n, s1 * chunk) - plan[g].first;
            plan[g].last = s1 == S && s0 < S;
        }
        return;
    }
    const uint64_t per = ((n + G - 1) / G + 3) & ~3ull; /* multiples of four: a shard starts on a byte of the packed output */
    uint64_t words = 0, next = 0;
    for (size_t g = 0; g < G; ++g) {
        plan[g].first = std::min<uint64_t>(n, g * per);
        plan[g].count = std::min<uint64_t>(n, plan[g].first + per) - plan[g].first;
        if (mode != PCU_MODEL_PER_ROLLOUT) { /* model draws before the shard: prefix of the depths (host memory) */
            if (depth)
                for (; next < plan[g].first; ++next) words += 4ull * depth[next];
            else
                words = 4ull * depth_u * plan[g].first;
            plan[g].words_before = words;
        }
    }
    for (size_t g = G; g-- > 0;)
        if (plan[g].count) {
            plan[g].last = true;
            break;
        }
}

template <typename T>
const T *off(const T *p, uint64_t k) {
    return p ? p + k : nullptr;
}
template <typename T>
T *off(T *p, uint64_t k) {
    return p ? p + k : nullptr;
}

/* Run `call(kid, g)` for every non-empty shard on the workers, then reduce the counters over the GPUs (NCCL
 * all-reduce of eight u64 — the path's only collective), synchronise, and hand back the post-batch engine. */
template <typename Call>
int multi_dispatch(pcu_ctx *ctx, const std::vector<ShardPlan> &plan, uint64_t chunk, pcu_model_rng *rng, pcu_out *out,
                   std::vector<std::vector<uint32_t>> &engines, Call call) {
    Multi &M = *ctx->multi;
    const size_t G = M.kids.size();
    std::vector<int> rcs(G, PCU_OK);
    const auto t_begin = std::chrono::steady_clock::now();
    auto ms_since = [](std::chrono::steady_clock::time_point t) {
        return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t).count();
    };
    for (size_t g = 0; g < G; ++g) {
        pcu_ctx *kid = M.kids[g];
        kid->hint = pcu_ctx::ShardHint{};
        if (!plan[g].count) continue;
        kid->hint.active = chunk != 0;
        kid->hint.chunk = chunk;
        kid->hint.first_stream = plan[g].first_stream;
        kid->hint.last = plan[g].last;
        kid->hint.keep_dev_sums = out->sums != nullptr;
        kid->hint.no_sync = true;
        M.workers[g]->start([&, g, kid] { rcs[g] = call(kid, g); });
    }
    for (size_t g = 0; g < G; ++g)
        if (plan[g].count) M.workers[g]->wait();
    int rc = PCU_OK;
    for (size_t g = 0; g < G && !rc; ++g)
        if (rcs[g]) {
            rc = rcs[g];
            ctx->err = "GPU " + std::to_string(M.kids[g]->device) + ": " + M.kids[g]->err;
        }
    M.t_dispatch_ms = ms_since(t_begin);
    const auto t_reduce = std::chrono::steady_clock::now();
    /* ---- the collective: counters of every shard summed on every GPU */
    std::vector<size_t> act;
    for (size_t g = 0; g < G; ++g)
        if (plan[g].count) act.push_back(g);
    if (!rc && out->sums && !act.empty()) {
        /* GPUs without a shard contribute zeros: the communicator spans all of them */
        for (size_t g = 0; g < G && !rc; ++g)
            if (!plan[g].count) {
                pcu_ctx *kid = M.kids[g];
                if (cudaSetDevice(kid->device) != cudaSuccess || kid->buf[B_SUMS].ensure(64) != cudaSuccess ||
                    cudaMemsetAsync(kid->buf[B_SUMS].p, 0, 64, kid->stream) != cudaSuccess)
                    rc = fail(ctx, PCU_ERR_CUDA, "zeroing the counters of an idle GPU failed");
            }
        pcu_ctx *k0 = M.kids[0];
        if (!rc && M.same_device) {
            SumPtrs S;
            S.n = (int)G;
            for (size_t g = 0; g < G; ++g) S.p[g] = reinterpret_cast<const unsigned long long *>(M.kids[g]->buf[B_SUMS].p);
            cudaSetDevice(k0->device);
            for (size_t g = 1; g < G; ++g) {
                cudaEventRecord(M.ev_done[g], M.kids[g]->stream);
                cudaStreamWaitEvent(k0->stream, M.ev_done[g], 0);
            }
            k_sum_shards<<<1, 32, 0, k0->stream>>>(S, reinterpret_cast<unsigned long long *>(k0->buf[B_SUMS].p));
            k0->launches++;
            if (cudaGetLastError() != cudaSuccess) rc = fail(ctx, PCU_ERR_CUDA, "k_sum_shards launch failed");
        } else if (!rc) {
            ncclResult_t nr = M.nccl.GroupStart();
            for (size_t g = 0; g < G && nr == ncclSuccess; ++g) {
                void *buf = M.kids[g]->buf[B_SUMS].p;
                nr = M.nccl.AllReduce(buf, buf, 8, ncclUint64, ncclSum, M.comms[g], M.kids[g]->stream);
            }
            const ncclResult_t ne = M.nccl.GroupEnd();
            if (nr == ncclSuccess) nr = ne;
            if (nr != ncclSuccess) rc = fail(ctx, PCU_ERR_CUDA, "ncclAllReduce failed: %s", M.nccl.GetErrorString(nr));
            M.nccl_calls++;
        }
        if (!rc) {
            cudaSetDevice(k0->device);
            if (cudaMemcpyAsync(M.pin_sums, k0->buf[B_SUMS].p, 64, cudaMemcpyDeviceToHost, k0->stream) != cudaSuccess)
                rc = fail(ctx, PCU_ERR_CUDA, "copying the reduced counters failed");
        }
    }
    M.t_reduce_ms = ms_since(t_reduce);
    const auto t_sync = std::chrono::steady_clock::now();
    /* ---- every shard's outputs are complete once its stream drains */
    for (size_t g = 0; g < G; ++g) {
        pcu_ctx *kid = M.kids[g];
        cudaSetDevice(kid->device);
        const cudaError_t e = cudaStreamSynchronize(kid->stream);
        if (e != cudaSuccess && !rc) rc = fail(ctx, PCU_ERR_CUDA, "GPU %d: %s", kid->device, cudaGetErrorString(e));
    }
    M.t_sync_ms = ms_since(t_sync);
    if (!rc) {
        if (out->sums) {
            if (act.empty())
                memset(out->sums, 0, 64);
            else
                memcpy(out->sums, M.pin_sums, 64);
        }
        if (rng->mode == PCU_MODEL_STREAM)
            for (size_t g = G; g-- > 0;)
                if (plan[g].count) { /* the shard that owns the end of the batch holds the post-batch window */
                    if (M.kids[g]->engine_pending) engine_from_window(M.kids[g]->pin_small, rng->mt_x, rng->mt_p);
                    break;
                }
    }
    for (pcu_ctx *kid : M.kids) {
        kid->engine_pending = false;
        kid->hint = pcu_ctx::ShardHint{};
    }
    (void)engines;
    ctx->path = M.kids[act.empty() ? 0 : act[0]]->path + " x" + std::to_string(act.size()) + "gpu";
    return rc;
}

/* Engine of a ragged-depth shard: the caller's, advanced by the draws of the rollouts before it. */
int shard_engine(const pcu_model_rng *rng, uint64_t words_before, std::vector<uint32_t> &e) {
    e.resize(MT_N + 1);
    memcpy(e.data(), rng->mt_x, sizeof(uint32_t) * MT_N);
    e[MT_N] = *rng->mt_p;
    return pcu_mt_discard(e.data(), e.data() + MT_N, words_before);
}

int multi_rollout_tree(pcu_ctx *ctx, const pcu_rt_batch *b, pcu_model_rng *rng, pcu_out *out) {
    Multi &M = *ctx->multi;
    int rc = check_common(ctx, b, rng, out);
    if (rc) return rc;
    ctx->err.clear();
    if (M.kids.size() == 1 || b->mem != PCU_MEM_HOST || out->mem != PCU_MEM_HOST || b->n < M.min_rollouts ||
        rng->engine != PCU_ENGINE_MT19937 || b->growth != PCU_GROWTH_DEFAULT) {
        /* small batches, batches that already live on one GPU and the variant type lists stay on the first one */
        rc = rollout_tree(M.kids[0], b, rng, out);
        ctx->err = M.kids[0]->err;
        ctx->path = M.kids[0]->path;
        return rc;
    }
    if (!b->depth && b->depth_u == 0)
        return fail(ctx, PCU_ERR_INVALID, "depth_bound 0 never terminates in the reference (random-tree.hh:95-99)");
    std::vector<ShardPlan> plan;
    uint64_t chunk = 0;
    plan_shards(M, b->n, rng->mode, b->depth, b->depth_u, &chunk, plan);
    std::vector<std::vector<uint32_t>> engines(M.kids.size());
    return multi_dispatch(ctx, plan, chunk, rng, out, engines, [&](pcu_ctx *kid, size_t g) -> int {
        const ShardPlan &sp = plan[g];
        pcu_rt_batch sb = *b;
        sb.n = sp.count;
        sb.state_seed = off(b->state_seed, sp.first);
        sb.state_pos = off(b->state_pos, sp.first);
        sb.transition_seed = off(b->transition_seed, sp.first);
        sb.depth = off(b->depth, sp.first);
        sb.payoff_bias = off(b->payoff_bias, sp.first);
        pcu_model_rng sr = *rng;
        sr.seeds = off(rng->seeds, sp.first);
        if (rng->mode == PCU_MODEL_WORDS) sr.words = rng->words + sp.words_before;
        if (rng->mode == PCU_MODEL_STREAM && !chunk) {
            const int e = shard_engine(rng, sp.words_before, engines[g]);
            if (e) return e;
            sr.mt_x = engines[g].data();
            sr.mt_p = engines[g].data() + MT_N;
        }
        pcu_out so = *out;
        so.payoff_x2 = off(out->payoff_x2, sp.first);
        so.payoff_packed2 = off(out->payoff_packed2, sp.first / 4);
        so.traj = off(out->traj, sp.first * (uint64_t)out->traj_stride * 3);
        so.traj_hash = off(out->traj_hash, sp.first);
        so.sums = nullptr;
        return rollout_tree(kid, &sb, &sr, &so);
    });
}

int multi_rollout_mold(pcu_ctx *ctx, const pcu_mold_batch *b, pcu_model_rng *rng, pcu_out *out) {
    Multi &M = *ctx->multi;
    int rc = check_common(ctx, b, rng, out);
    if (rc) return rc;
    ctx->err.clear();
    if (M.kids.size() == 1 || b->mem != PCU_MEM_HOST || out->mem != PCU_MEM_HOST || b->n < M.min_rollouts ||
        (!b->depth && b->depth_u == 0) || rng->engine != PCU_ENGINE_MT19937) {
        rc = rollout_mold(M.kids[0], b, rng, out);
        ctx->err = M.kids[0]->err;
        ctx->path = M.kids[0]->path;
        return rc;
    }
    std::vector<ShardPlan> plan;
    uint64_t chunk = 0;
    plan_shards(M, b->n, rng->mode, b->depth, b->depth_u, &chunk, plan);
    std::vector<std::vector<uint32_t>> engines(M.kids.size());
    return multi_dispatch(ctx, plan, chunk, rng, out, engines, [&](pcu_ctx *kid, size_t g) -> int {
        const ShardPlan &sp = plan[g];
        pcu_mold_batch sb = *b;
        sb.n = sp.count;
        sb.depth = off(b->depth, sp.first);
        pcu_model_rng sr = *rng;
        sr.seeds = off(rng->seeds, sp.first);
        if (rng->mode == PCU_MODEL_WORDS) sr.words = rng->words + sp.words_before;
        if (rng->mode == PCU_MODEL_STREAM && !chunk) {
            const int e = shard_engine(rng, sp.words_before, engines[g]);
            if (e) return e;
            sr.mt_x = engines[g].data();
            sr.mt_p = engines[g].data() + MT_N;
        }
        pcu_out so = *out;
        so.payoff_x2 = off(out->payoff_x2, sp.first);
        so.payoff_packed2 = off(out->payoff_packed2, sp.first / 4);
        so.traj = off(out->traj, sp.first * (uint64_t)out->traj_stride * 3);
        so.traj_hash = off(out->traj_hash, sp.first);
        so.sums = nullptr;
        return rollout_mold(kid, &sb, &sr, &so);
    });
}

}  // namespace

extern "C" {

int pcu_rollout_random_tree(pcu_ctx *ctx, const pcu_rt_batch *b, pcu_model_rng *rng, pcu_out *out) {
    if (!ctx) return PCU_ERR_INVALID;
    return ctx->multi ? multi_rollout_tree(ctx, b, rng, out) : rollout_tree(ctx, b, rng, out);
}

int pcu_rollout_mold(pcu_ctx *ctx, const pcu_mold_batch *b, pcu_model_rng *rng, pcu_out *out) {
    if (!ctx) return PCU_ERR_INVALID;
    return ctx->multi ? multi_rollout_mold(ctx, b, rng, out) : rollout_mold(ctx, b, rng, out);
}

int pcu_create_multi(const int *device_ids, int n, pcu_ctx **out) {
    if (!out) return PCU_ERR_INVALID;
    *out = nullptr;
    if (!device_ids || n < 1 || n > 16) return PCU_ERR_INVALID;
    bool all_same = true, all_distinct = true;
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < i; ++j) {
            if (device_ids[i] == device_ids[j]) all_distinct = false;
            else all_same = false;
        }
    if (n > 1 && !all_same && !all_distinct) return PCU_ERR_INVALID; /* a communicator cannot hold a GPU twice */
    pcu_ctx *ctx = new pcu_ctx();
    Multi *M = new Multi();
    ctx->multi = M;
    ctx->device = device_ids[0];
    M->same_device = n > 1 && all_same;
    auto bail = [&](int code) {
        pcu_destroy(ctx);
        return code;
    };
    for (int i = 0; i < n; ++i) {
        pcu_ctx *kid = nullptr;
        const int rc = pcu_create(device_ids[i], &kid);
        if (rc) return bail(rc);
        M->kids.push_back(kid);
        M->workers.emplace_back(new Worker());
    }
    ctx->sms = M->kids[0]->sms;
    M->min_rollouts = (uint64_t)n << 16;
    cudaSetDevice(device_ids[0]);
    if (cudaHostAlloc(reinterpret_cast<void **>(&M->pin_sums), 64, cudaHostAllocPortable) != cudaSuccess) return bail(PCU_ERR_CUDA);
    if (M->same_device) {
        M->ev_done.resize(n);
        for (int i = 0; i < n; ++i)
            if (cudaEventCreateWithFlags(&M->ev_done[i], cudaEventDisableTiming) != cudaSuccess) return bail(PCU_ERR_CUDA);
    } else if (n > 1) {
        /* one process, one communicator per GPU; there is no non-NCCL path between distinct GPUs */
        if (!M->nccl.load()) return bail(PCU_ERR_UNSUPPORTED);
        M->comms.assign(n, nullptr);
        if (M->nccl.CommInitAll(M->comms.data(), n, device_ids) != ncclSuccess) {
            M->comms.clear();
            return bail(PCU_ERR_CUDA);
        }
    }
    *out = ctx;
    return PCU_OK;
}

uint64_t pcu_xs_discard(uint64_t state, uint64_t n) { return host_xs_discard(state, n); }

int pcu_visible_devices(void) {
    int count = 0, ok = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    for (int i = 0; i < count; ++i) {
        int major = 0;
        if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, i) == cudaSuccess && major == 10) ++ok;
        else break; /* device ids must stay contiguous from 0 */
    }
    return ok;
}

int pcu_device_count(const pcu_ctx *ctx) { return !ctx ? 0 : ctx->multi ? (int)ctx->multi->kids.size() : 1; }

int pcu_get_stat(pcu_ctx *ctx, const char *key, double *value) {
    if (!ctx || !key || !value) return PCU_ERR_INVALID;
    const pcu_ctx *c = ctx->multi ? ctx->multi->kids[0] : ctx;
    if (!strcmp(key, "last_chunk")) *value = (double)c->last_chunk;
    else if (!strcmp(key, "last_streams")) *value = (double)c->last_streams;
    else if (!strcmp(key, "last_family_build_ms")) {
        *value = 0;
        if (ctx->multi) { for (const pcu_ctx *k : ctx->multi->kids) *value = std::max(*value, k->last_family_ms); }
        else *value = c->last_family_ms;
    } else if (!strcmp(key, "last_table_build_ms")) *value = c->last_table_ms;
    else if (!strcmp(key, "nccl_calls")) *value = ctx->multi ? (double)ctx->multi->nccl_calls : 0.0;
    else if (!strcmp(key, "multi_dispatch_ms")) *value = ctx->multi ? ctx->multi->t_dispatch_ms : 0.0;
    else if (!strcmp(key, "multi_reduce_ms")) *value = ctx->multi ? ctx->multi->t_reduce_ms : 0.0;
    else if (!strcmp(key, "multi_sync_ms")) *value = ctx->multi ? ctx->multi->t_sync_ms : 0.0;
    else return fail(ctx, PCU_ERR_INVALID, "unknown statistic %s", key);
    return PCU_OK;
}

/* Build (and upload) now what the first call of a new batch shape would otherwise build inside the call: the family of
 * jump polynomials for `n` uniform-depth rollouts in mode S (~80 ms of host GF(2) arithmetic per new stride). */
int pcu_warmup(pcu_ctx *ctx, uint64_t n, uint32_t depth) {
    if (!ctx || depth == 0) return PCU_ERR_INVALID;
    std::string jerr;
    if (ctx->multi) {
        Multi &M = *ctx->multi;
        if (M.kids.size() == 1 || n < M.min_rollouts) return pcu_warmup(M.kids[0], n, depth);
        std::vector<ShardPlan> plan;
        uint64_t chunk = 0;
        plan_shards(M, n, PCU_MODEL_STREAM, nullptr, depth, &chunk, plan);
        for (size_t g = 0; g < M.kids.size(); ++g) {
            if (!plan[g].count) continue;
            pcu_ctx *kid = M.kids[g];
            cudaSetDevice(kid->device);
            const uint64_t streams = (plan[g].count + chunk - 1) / chunk;
            const int64_t extra = streams % 2 == 0 ? (int64_t)(plan[g].first_stream + streams - 1) : -1;
            if (kid->jump.prepare(kid->stream, plan[g].first_stream, (streams + 1) / 2, 4ull * depth * chunk, &jerr, 2, extra))
                return fail(ctx, PCU_ERR_CUDA, "warm-up failed: %s", jerr.c_str());
        }
        return PCU_OK;
    }
    CU(cudaSetDevice(ctx->device));
    const uint64_t chunk = pick_rollout_chunk(n, (uint64_t)ctx->sms * streams_per_sm_for(ctx, n, depth), 128);
    const uint64_t streams = (n + chunk - 1) / chunk;
    if (ctx->jump.prepare(ctx->stream, 0, (streams + 1) / 2, 4ull * depth * chunk, &jerr, 2,
                          streams % 2 == 0 ? (int64_t)(streams - 1) : -1))
        return fail(ctx, PCU_ERR_CUDA, "warm-up failed: %s", jerr.c_str());
    return PCU_OK;
}

int pcu_measure_int_peak(pcu_ctx *ctx, int variant, int reps, double *ops_per_second) {
    if (!ctx || !ops_per_second) return PCU_ERR_INVALID;
    if (ctx->multi) return pcu_measure_int_peak(ctx->multi->kids[0], variant, reps, ops_per_second);
    CU(cudaSetDevice(ctx->device));
    const int iters = 4096;
    const int blocks = ctx->sms * 8, threads = 256;
    CU(ctx->buf[B_SCRATCH].ensure(sizeof(uint32_t) * blocks * threads));
    uint32_t *d = reinterpret_cast<uint32_t *>(ctx->buf[B_SCRATCH].p);
    cudaEvent_t e0, e1;
    CU(cudaEventCreate(&e0));
    CU(cudaEventCreate(&e1));
    double best = 0;
    for (int r = 0; r < reps + 1; ++r) {
        CU(cudaEventRecord(e0, ctx->stream));
        switch (variant) {
            case 0: k_int_peak<0><<<blocks, threads, 0, ctx->stream>>>(d, 12345u + r, iters); break;
            case 1: k_int_peak<1><<<blocks, threads, 0, ctx->stream>>>(d, 12345u + r, iters); break;
            case 2: k_int_peak<2><<<blocks, threads, 0, ctx->stream>>>(d, 12345u + r, iters); break;
            case 3: k_int_peak<3><<<blocks, threads, 0, ctx->stream>>>(d, 12345u + r, iters); break;
            case 4: k_int_peak<4><<<blocks, threads, 0, ctx->stream>>>(d, 12345u + r, iters); break;
            case 5: k_int_peak<5><<<blocks, threads, 0, ctx->stream>>>(d, 12345u + r, iters); break;
            case 6: k_int_peak<6><<<blocks, threads, 0, ctx->stream>>>(d, 12345u + r, iters); break;
            case 7: k_int_peak<7><<<blocks, threads, 0, ctx->stream>>>(d, 12345u + r, iters); break;
            case 8: k_int_peak<8><<<blocks, threads, 0, ctx->stream>>>(d, 12345u + r, iters); break;
            default: k_int_peak<9><<<blocks, threads, 0, ctx->stream>>>(d, 12345u + r, iters); break;
        }
        ctx->launches++;
        CU(cudaEventRecord(e1, ctx->stream));
        CU(cudaEventSynchronize(e1));
        float ms = 0;
        CU(cudaEventElapsedTime(&ms, e0, e1));
        const double ops = (double)blocks * threads * iters * INT_PEAK_OPS_PER_ITER;
        if (r > 0 && ms > 0) best = std::max(best, ops / (ms * 1e-3));
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    *ops_per_second = best;
    return PCU_OK;
}

}  // extern "C"

/* ---- pcu_leaf_batch_noalloc: a leaf batch with the state-stream table in STATIC device memory ---------------------
 * No counterpart in the reference; added for callers that cannot call cudaMalloc
 * (CUDA graph capture, tight serving loops, harnesses that forbid it).  It is the "leaf_table_W" path (k_leaf_table,
 * leaf.cuh) with its one piece of device scratch, the state-stream table, in the file-scope array below instead of a
 * grown-on-demand context buffer.  Every GPU that loads the library holds the array until the process exits.  All calls
 * on a device share it and it keeps the last call's table: each call rewrites the whole range it reads (nothing has to
 * start at zero; the caller's sums are cleared here), and calls that overlap on one GPU race. */
static __device__ __align__(256) uint8_t g_noalloc_scratch[PCU_NOALLOC_SCRATCH_BYTES];

namespace {
constexpr int NOALLOC_MAX_DEVICES = 64;
std::mutex g_noalloc_mu;
uint8_t *g_noalloc_dev_ptr[NOALLOC_MAX_DEVICES]; /* DEVICE address of g_noalloc_scratch per device (the name itself is
                                                    a host-side stand-in in host code, not a device address) */

/* once per device, with that device current */
int noalloc_scratch(pcu_ctx *ctx, uint8_t **out) {
    if (ctx->device < 0 || ctx->device >= NOALLOC_MAX_DEVICES)
        return fail(ctx, PCU_ERR_INVALID, "pcu_leaf_batch_noalloc: device index %d out of range", ctx->device);
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    if (!g_noalloc_dev_ptr[ctx->device]) {
        CU(cudaSetDevice(ctx->device));
        uint8_t *p = nullptr;
        CU(cudaGetSymbolAddress((void **)&p, g_noalloc_scratch));
        if (!p) return fail(ctx, PCU_ERR_CUDA, "cudaGetSymbolAddress returned a null address");
        g_noalloc_dev_ptr[ctx->device] = p;
    }
    *out = g_noalloc_dev_ptr[ctx->device];
    return PCU_OK;
}
}  // namespace

extern "C" int pcu_leaf_batch_noalloc(pcu_ctx *ctx, const pcu_rt_batch *b, const uint32_t *d_words, uint8_t *d_payoff_x2,
                                      uint64_t *d_sums) {
    if (!ctx) return PCU_ERR_INVALID;
    if (ctx->multi) return fail(ctx, PCU_ERR_INVALID, "pcu_leaf_batch_noalloc needs a single-GPU context (pcu_create)");
    if (!b || !d_words || (!d_payoff_x2 && !d_sums))
        return fail(ctx, PCU_ERR_INVALID, "pcu_leaf_batch_noalloc: null batch, words or outputs");
    if (b->state_seed || b->state_pos || b->transition_seed || b->depth || b->payoff_bias)
        return fail(ctx, PCU_ERR_INVALID, "pcu_leaf_batch_noalloc takes batch-uniform fields only (leaves of ONE state)");
    if (b->growth != PCU_GROWTH_DEFAULT) return fail(ctx, PCU_ERR_UNSUPPORTED, "pcu_leaf_batch_noalloc: default growth only");
    const uint32_t T = b->transitions;
    if (b->rows < 1 || b->rows > 255 || b->cols < 1 || b->cols > 255 || T < 1 || b->depth_u < 1 || b->thr_q < 1)
        return fail(ctx, PCU_ERR_INVALID, "pcu_leaf_batch_noalloc: rows, cols in 1..255, transitions, depth, thr_q >= 1");
    if (T > 8) return fail(ctx, PCU_ERR_UNSUPPORTED, "transitions > 8");
    if (b->n == 0) return PCU_OK;
    if (b->n > PCU_NOALLOC_MAX_LEAVES)
        return fail(ctx, PCU_ERR_INVALID, "pcu_leaf_batch_noalloc: %llu leaves exceed the maximum of %llu",
                    (unsigned long long)b->n, (unsigned long long)PCU_NOALLOC_MAX_LEAVES);
    const uint32_t gen_words = 2u * b->rows * b->cols * T;
    if (b->state_pos_u < gen_words)
        return fail(ctx, PCU_ERR_INVALID, "state_pos is below 2*rows*cols*transitions");
    const uint64_t per_step = (uint64_t)b->rows * b->cols * T + 2 + gen_words;
    const uint64_t span = (uint64_t)gen_words + per_step * b->depth_u + 8;
    const uint64_t bytes = span * (T > 1 ? 2 : 1);
    if (bytes > PCU_NOALLOC_SCRATCH_BYTES)
        return fail(ctx, PCU_ERR_INVALID,
                    "pcu_leaf_batch_noalloc: the state-stream table of this shape and depth needs %llu bytes, the static "
                    "scratch holds %llu (use pcu_rollout_random_tree)",
                    (unsigned long long)bytes, (unsigned long long)PCU_NOALLOC_SCRATCH_BYTES);
    uint8_t *scratch = nullptr;
    int rc;
    if ((rc = noalloc_scratch(ctx, &scratch))) return rc;
    CU(cudaSetDevice(ctx->device));

    const uint32_t tab_base = b->state_pos_u - gen_words, tab_len = (uint32_t)span;
    const double pu = (T > 1) ? host_uniform_from_seed(b->transition_seed_u) : 0.0;
    std::vector<uint8_t> tab; /* host heap only */
    host_build_table(b->state_seed_u, tab_base, tab_len, T, pu, b->thr_p, b->thr_q, tab);
    if (tab.size() != bytes) return fail(ctx, PCU_ERR_INVALID, "internal: table size mismatch");
    CU(cudaMemcpyAsync(scratch, tab.data(), tab.size(), cudaMemcpyHostToDevice, ctx->stream));
    if (d_sums) CU(cudaMemsetAsync(d_sums, 0, sizeof(uint64_t) * 8, ctx->stream));

    LeafParams L;
    memset(&L, 0, sizeof(L));
    L.n = b->n;
    L.kind = KIND_TREE;
    L.cols = b->cols;
    L.T = T;
    L.gen_words = gen_words;
    L.mrow = make_fold_const(b->rows);
    L.mcol = make_fold_const(b->cols);
    L.words = d_words;
    L.depth_u = b->depth_u;
    L.tab = scratch;
    L.tab_base = tab_base;
    L.tab_len = tab_len;
    L.state_pos_u = b->state_pos_u;
    L.bias_u = b->payoff_bias_u;
    L.payoff = d_payoff_x2;
    L.sums = reinterpret_cast<unsigned long long *>(d_sums);
    k_leaf_table<<<(unsigned)((L.n + 127) / 128), 128, 0, ctx->stream>>>(L);
    ctx->launches++;
    CU(cudaGetLastError());
    CU(cudaStreamSynchronize(ctx->stream)); /* the host table must outlive the copy; outputs are complete on return */
    ctx->path = "leaf_table_noalloc";
    return PCU_OK;
}
