/* libpinyon_cuda — C ABI of the B200-native Monte-Carlo leaf evaluator for Pinyon.
 *
 * Pinyon (baskuit/pinyon) has no FFI of its own: its model boundary is a C++ type
 * list (src/model/model.hh:8-79).  This header is the boundary a maintainer binds
 * that type list to; include/pinyon_cuda/monte_carlo_model.hh is that binding
 * (header-only C++23, no CUDA types).  Every entry point names the reference
 * interface it replaces, file:line relative to the reference tree.
 *
 * Conventions
 *   - extern "C", plain pointers and sizes; no CUDA or torch types.
 *   - every function returning int returns PCU_OK (0) or a negative PCU_ERR_*;
 *     nothing throws; pcu_last_error() gives a human-readable reason.
 *   - a context is bound to one GPU and may be used by one thread at a time.
 *   - there is NO CPU fallback: without a usable sm_100 device pcu_create fails.
 *
 * Semantics reproduced bit-for-bit:
 *   prng            = libstdc++ std::mt19937 + its uniform_int<uint64_t> / generate_canonical<double,53>
 *                     arithmetic                       src/types/random.hh:8-64
 *   RandomTree      = default growth functions only    src/state/random-tree.hh:12-200
 *   MoldState                                          src/state/test-states.hh:13-67
 *   MonteCarloModel rollout / batched inference        src/model/monte-carlo-model.hh:41-96
 */
#ifndef PINYON_CUDA_H
#define PINYON_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PCU_VERSION 200

#define PCU_OK 0
#define PCU_ERR_INVALID (-1)     /* bad argument (null pointer, depth 0, rows or cols > 255, ...) */
#define PCU_ERR_CUDA (-2)        /* a CUDA runtime call failed */
#define PCU_ERR_UNSUPPORTED (-3) /* legal in Pinyon but outside what the device path reproduces (transitions > 8: the
                                  * reference's Rational<int> arithmetic overflows int there, random-tree.hh:165-184) */
#define PCU_ERR_NOMEM (-4)

/* Where the array pointers of a batch / output live. */
#define PCU_MEM_HOST 0   /* host memory; the call copies in and out (pinned memory is faster) */
#define PCU_MEM_DEVICE 1 /* device memory of the context's GPU; nothing is copied */

typedef struct pcu_ctx pcu_ctx;

/* ---- context ------------------------------------------------------------------ */

/* Create a context on CUDA device `device_id`.  Fails with PCU_ERR_CUDA if there is
 * no such device or it is not compute capability 10.x. */
int pcu_create(int device_id, pcu_ctx **out);
/* Create ONE context over `n` GPUs of this box (north_star (d), SURVEY.md 8(e): one process, ncclCommInitAll).
 * A rollout call on it is cut into contiguous rollout-index shards, one per GPU, each played by that GPU's own
 * streams from its slice of the ONE model stream (the shard's jump-ahead polynomials are applied on the GPU to
 * the caller's engine: no host-side jump), the outputs land in the caller's arrays at the shard's offset, and the
 * eight u64 counters of pcu_out.sums are summed over the GPUs with one ncclAllReduce(ncclUint64, ncclSum) — the
 * path's only collective.  Results, including the post-batch engine, are byte-identical to the same call on one GPU.
 * Batches must live in HOST memory (pinned, from pcu_host_alloc, for zero-copy stores from every GPU); batches
 * below 65 536 rollouts per GPU (option "multi_min_rollouts") and device-memory batches run on device_ids[0] alone.
 * NCCL is loaded at run time (libnccl.so.2); PCU_ERR_UNSUPPORTED if it is absent and n > 1.  Listing one GPU n
 * times is allowed (a test configuration: the shards then share the GPU and the counters are summed by a kernel). */
int pcu_create_multi(const int *device_ids, int n, pcu_ctx **out);
/* Number of CUDA devices of compute capability 10.x visible to this process (0 without a usable GPU). */
int pcu_visible_devices(void);
/* GPUs behind a context: 1 for pcu_create. */
int pcu_device_count(const pcu_ctx *ctx);
void pcu_destroy(pcu_ctx *ctx);
const char *pcu_last_error(const pcu_ctx *ctx);
int pcu_version(void);
/* Run on an existing cudaStream_t (passed as void*) instead of the context's own
 * stream; NULL restores the private stream.  Calls stay synchronous for the caller. */
int pcu_set_stream(pcu_ctx *ctx, void *cuda_stream);
/* Number of kernels launched by this context since creation (bench.py gpu_launches). */
uint64_t pcu_kernel_launches(const pcu_ctx *ctx);
/* Device time in ms of the rollout kernel (k_stream2 / k_warp) of the last rollout call,
 * from CUDA events recorded on the launching stream around that launch; -1 if none. */
double pcu_last_kernel_ms(pcu_ctx *ctx);
/* Statistics of the last rollout call, for tests and benchmarks: "last_chunk" (rollouts, or steps for ragged batches,
 * per model stream), "last_streams", "last_family_build_ms" and "last_table_build_ms" (host time the call spent
 * building a jump-polynomial family / the state-stream table: 0 when cached), "nccl_calls". */
int pcu_get_stat(pcu_ctx *ctx, const char *key, double *value);
/* Pay the first-call costs of a batch shape now: builds and uploads the jump-polynomial family that `n`
 * uniform-depth mode-S rollouts need (~80 ms of host GF(2) arithmetic per new stride; cached afterwards: the
 * host keeps families up to 256 MiB, option "jump_host_cache_mb", each context `jump_families` = 8 on its GPU). */
int pcu_warmup(pcu_ctx *ctx, uint64_t n, uint32_t depth);
/* Tuning knobs, mainly for benchmarks: key "streams_per_sm" (1..16; default 0 = 4, or 8 for calls of >= 2.5e8 rollout-steps) sets how many
 * model-stream generator warps the fused kernels keep resident per SM (each needs one
 * jump-ahead checkpoint per call).  Returns PCU_ERR_INVALID for unknown keys. */
int pcu_set_option(pcu_ctx *ctx, const char *key, int64_t value);
/* Pinned host memory for batch inputs/outputs (optional; any host pointer works). */
void *pcu_host_alloc(pcu_ctx *ctx, size_t bytes);
void pcu_host_free(pcu_ctx *ctx, void *p);

/* ---- model PRNG -----------------------------------------------------------------
 * Replaces MonteCarloModel::Model::device, a Types::PRNG (monte-carlo-model.hh:37),
 * i.e. prng = seed + std::mt19937 (random.hh:8-13).
 *
 * PCU_MODEL_STREAM ("mode S") is the reference's batched inference verbatim
 * (monte-carlo-model.hh:58-67): ONE engine consumed by the rollouts in index order,
 * rollout i starting where rollout i-1 stopped.  The engine is passed in libstdc++'s
 * own representation, 624 state words + the position 0..624 of the next word, and is
 * advanced in place to its post-batch state, so the host model stays in lock-step
 * with a CPU model that ran the same batch.
 *
 * PCU_MODEL_PER_ROLLOUT ("mode P"): rollout i runs on a fresh prng{seeds[i]}; what the
 * reference computes when called as Model{prng{seeds[i]}}.inference(state_i).
 *
 * PCU_MODEL_WORDS ("mode W"): every rollout brings the engine words its own model would draw: words[] holds, rollout
 * after rollout, the next 4 * depth_i RAW (untempered) std::mt19937 state words of that rollout's engine (two
 * random_int per step, two words each, monte-carlo-model.hh:89-90).  This is how single-leaf inferences of MANY
 * models (one per search tree, each with its own sequential stream; tree-bandit.hh:103,111) are evaluated as one
 * batch: the caller takes the words with pcu_mt_take_raw, which advances its engine exactly as the rollout would;
 * tempering, the uniform_int arithmetic (random.hh:26,28) and the rollout run on the GPU. */
#define PCU_MODEL_STREAM 0
#define PCU_MODEL_PER_ROLLOUT 1
#define PCU_MODEL_WORDS 2

/* Which PRNG type the type list uses (Types::PRNG, src/types/types.hh:27-76) — for the model AND the states of the
 * batch.  PCU_ENGINE_XOR_SHIFT is Pinyon's xor_shift (random.hh:66-121; it needs a random_seed() member added to
 * satisfy IsPRNG): 64 bits of state, one step per draw.  With it, mode S takes the model engine's
 * state in *xs_state (updated in place), mode P takes states in seeds[], mode W is not offered, and a RandomTree
 * batch carries in state_seed the state engine's CURRENT 64-bit state (state_pos is ignored). */
#define PCU_ENGINE_MT19937 0
#define PCU_ENGINE_XOR_SHIFT 1

typedef struct {
    int32_t mode;
    uint32_t *mt_x;        /* mode S: [624] engine words, HOST memory, updated in place */
    uint32_t *mt_p;        /* mode S: next-word position 0..624, HOST memory, updated in place */
    const uint64_t *seeds; /* mode P: [n] seeds, in the memory space of the batch */
    const uint32_t *words; /* mode W: [4 * sum of depths] raw engine words, in the memory space of the batch */
    uint32_t engine;       /* PCU_ENGINE_* */
    uint64_t *xs_state;    /* xor_shift, mode S: the model engine's state, HOST memory, updated in place */
} pcu_model_rng;

/* Host-side engine helpers (no GPU needed).  pcu_mt_seed == std::mt19937::seed as
 * prng(seed) calls it (random.hh:16); pcu_mt_discard == prng::discard (random.hh:63),
 * using polynomial jump-ahead for large n. */
void pcu_mt_seed(uint64_t seed, uint32_t *mt_x, uint32_t *mt_p);
int pcu_mt_discard(uint32_t *mt_x, uint32_t *mt_p, uint64_t n);
uint32_t pcu_mt_next(uint32_t *mt_x, uint32_t *mt_p);
/* The next n RAW (untempered) state words of the engine, which is advanced by n draws: the words a mode-W rollout of
 * n / 4 steps consumes.  pcu_mt_next(...) == temper(raw word). */
int pcu_mt_take_raw(uint32_t *mt_x, uint32_t *mt_p, uint64_t n, uint32_t *out);
/* xor_shift::discard (random.hh:107-111) in O(64^2 log n): the state after n steps. */
uint64_t pcu_xs_discard(uint64_t state, uint64_t n);
/* Diagnostic: the engine advanced by c * stride draws through the cached polynomial family
 * x^(c * stride) mod phi that the device jump-ahead uses for its checkpoints (c < 65536). */
int pcu_mt_discard_family(uint32_t *mt_x, uint32_t *mt_p, uint64_t stride, uint64_t c);

/* ---- RandomTree batch -----------------------------------------------------------
 * Structure-of-arrays image of n RandomTree::State objects (random-tree.hh:14-27).
 * A per-rollout array pointer may be NULL, in which case the matching *_u scalar
 * holds for every rollout (a batch of copies of one root needs no arrays at all).
 *
 *   state_seed       device.get_seed()            (random.hh:18) — the engine uses seed mod 2^32
 *   state_pos        draws the state's device has made since seeding, INCLUDING the
 *                    2*rows*cols*transitions of its latest get_chance_strategies()
 *                    (ctor :31-41 or apply_actions :105-107)
 *   transition_seed  (:15,59-61)
 *   depth            depth_bound (:16), >= 1.  A zero ENTRY of a per-rollout depth array plays no step and
 *                    returns the payoff of its payoff_bias (the drop-in model never sends such states: terminal
 *                    states are answered on the host); a batch-uniform depth of 0 is refused, because the
 *                    reference never terminates from a depth_bound == 0 root (:95-99)
 *   payoff_bias      (:20)
 *   rows, cols, transitions, chance_threshold = thr_p/thr_q  (:17-21) batch-uniform
 *   growth           which growth functions the states carry (:25-27): PCU_GROWTH_DEFAULT (:141-146), or
 *                    PCU_GROWTH_ALTERNATING_CONSTANT = GrowLib::alternating_constant_actions (libpinyon/grow-lib.hh:16-24)
 *                    as actions_func with the default depth and bias functions — rows and cols then change from step to
 *                    step.  (grow-lib's other function calls a prng member that does not exist and cannot be instantiated.)
 * Other growth functions are host function pointers and are refused by the drop-in model. */
#define PCU_GROWTH_DEFAULT 0
#define PCU_GROWTH_ALTERNATING_CONSTANT 1

typedef struct {
    uint64_t n;
    int32_t mem;
    const uint64_t *state_seed;
    uint64_t state_seed_u;
    const uint32_t *state_pos;
    uint32_t state_pos_u;
    const uint64_t *transition_seed;
    uint64_t transition_seed_u;
    const uint16_t *depth;
    uint16_t depth_u;
    const int32_t *payoff_bias;
    int32_t payoff_bias_u;
    uint32_t rows, cols, transitions;
    int32_t thr_p, thr_q;
    uint32_t growth;
} pcu_rt_batch;

/* MoldState batch (test-states.hh:13-24): n_actions uniform, max_depth per rollout. */
typedef struct {
    uint64_t n;
    int32_t mem;
    const uint16_t *depth;
    uint16_t depth_u;
    uint32_t n_actions;
} pcu_mold_batch;

/* ---- outputs ---------------------------------------------------------------------
 * Replaces ModelOutput::value / ModelBatchOutput (monte-carlo-model.hh:6-32).
 *   payoff_x2[i]  2 * row payoff: 0, 1 (= 1/2) or 2 (RandomTree :99-103); MoldState's
 *                 payoff is never written by the reference, so it is 0 = (0,0).
 *   payoff_packed2  optional; the same values at 2 bits per rollout, rollout i in bits 2 (i % 4) .. of byte
 *                 i / 4, (n + 3) / 4 bytes: a quarter of the bytes that cross PCIe per batch
 *   traj          optional; 3 bytes per step (row action, col action, chance action),
 *                 traj_stride steps reserved per rollout, zero filled past the end
 *   traj_hash[i]  optional; FNV-1a style fold of the trajectory:
 *                 h = 2166136261; per step h = (h ^ ((row*cols+col) | chance << 16)) * 16777619
 *   sums[8]       optional; [0] row wins [1] draws (MoldState: all) [2] row losses [3] steps
 *                 [4] sum of traj_hash mod 2^64, [5..7] zero.  All exact integers, so batch
 *                 means are bit-exact: mean row value = (2*sums[0] + sums[1]) / (2.0 * n).
 * All pointers are in out->mem space and may be NULL. */
typedef struct {
    int32_t mem;
    uint8_t *payoff_x2;
    uint8_t *traj;
    uint32_t traj_stride;
    uint32_t *traj_hash;
    uint64_t *sums;
    uint8_t *payoff_packed2;
} pcu_out;

/* ---- the hot path ------------------------------------------------------------------
 * Batched inference: replaces Model::inference(ModelBatchInput&, ModelBatchOutput&)
 * (monte-carlo-model.hh:58-67) and, with n == 1, Model::inference(State&&, ModelOutput&)
 * (:41-56) for RandomTree / MoldState states.  Blocks until outputs are complete. */
int pcu_rollout_random_tree(pcu_ctx *ctx, const pcu_rt_batch *batch, pcu_model_rng *rng, pcu_out *out);
int pcu_rollout_mold(pcu_ctx *ctx, const pcu_mold_batch *batch, pcu_model_rng *rng, pcu_out *out);

/* NO COUNTERPART IN THE REFERENCE: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * A leaf batch of ONE RandomTree state in model-rng mode W (what pcu_rollout_random_tree does with rng->mode ==
 * PCU_MODEL_WORDS, batch->mem == PCU_MEM_DEVICE and batch-uniform fields: path "leaf_table_W") that allocates no device
 * memory: the state-stream table, that path's only scratch, lives in a static device array of
 * PCU_NOALLOC_SCRATCH_BYTES that every GPU which loads the library holds until the process exits.
 *   batch        every per-rollout pointer NULL (the *_u scalars describe the one state), growth default, mem ignored
 *   d_words      DEVICE u32[n * 4 * depth_u]: each leaf's raw engine words (pcu_mt_take_raw), leaf-major
 *   d_payoff_x2  DEVICE u8[n], may be NULL;  d_sums  DEVICE u64[8] as pcu_out.sums, may be NULL (cleared by the call)
 * Limits, rejected with PCU_ERR_INVALID and a message in pcu_last_error: n <= PCU_NOALLOC_MAX_LEAVES, and the table
 * ((2*R*C*T + (3*R*C*T + 2) * depth + 8) bytes, twice that when T > 1) must fit the scratch; single-GPU contexts only.
 * Calls no cudaMalloc / cudaFree; runs on the context's stream and returns when the outputs are complete.  All calls
 * on a device share the array, which keeps the last call's table: calls that overlap on one GPU race. */
#define PCU_NOALLOC_SCRATCH_BYTES (64u * 1024u)
#define PCU_NOALLOC_MAX_LEAVES (1ull << 24)
int pcu_leaf_batch_noalloc(pcu_ctx *ctx, const pcu_rt_batch *batch, const uint32_t *d_words, uint8_t *d_payoff_x2,
                           uint64_t *d_sums);

/* Name of the kernel path the last rollout call took ("stream_table", "stream_mold",
 * "warp_private", ...); for tests and benchmarks. */
const char *pcu_last_path(const pcu_ctx *ctx);

/* Integer-issue peak microbenchmark (eight cross-coupled chains per thread, best of
 * `reps`, CUDA events), the roofline denominator of SURVEY.md §8(d).  variant 0: LOP3
 * only; 1: LOP3 + SHF + IADD3 (ALU pipe); 2: LOP3 + SHF + IMAD (ALU + FMA pipes); 3..6 are
 * single-pipe and mixed probes (IMAD, IMAD.HI, LOP3:IMAD 1:1, LOP3:IMAD:IMAD.HI).
 * Writes int32 thread-operations per second. */
int pcu_measure_int_peak(pcu_ctx *ctx, int variant, int reps, double *ops_per_second);

#ifdef __cplusplus
}
#endif
#endif /* PINYON_CUDA_H */
