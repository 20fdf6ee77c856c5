/* Profiling shim — NOT part of the product, the tests or the bench, built by nothing in this repository, and NOT a
 * CPU fallback: it computes no rollout at all (every payoff is the constant 1).  It exists so that the HOST-side cost of
 * the batching adapter (include/pinyon_cuda/batcher.hh) can be profiled on a machine without a GPU:
 *   gcc -O2 -shared -fPIC -I include -o /tmp/libfake.so tools/fake_gpu.c
 *   LD_PRELOAD=/tmp/libfake.so tests/cpp/_bin/bench_search gpu_tb_batched 0 0 512 150 0 1 512
 * (profiles/r3b_host_profile_batched_treebandit.txt). */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include "pinyon_cuda.h"
struct pcu_ctx { int dummy; };
int pcu_create(int device_id, pcu_ctx **out) { (void)device_id; *out = (pcu_ctx *)calloc(1, sizeof(struct pcu_ctx)); return 0; }
int pcu_create_multi(const int *ids, int n, pcu_ctx **out) { (void)ids; (void)n; return pcu_create(0, out); }
void pcu_destroy(pcu_ctx *c) { free(c); }
const char *pcu_last_error(const pcu_ctx *c) { (void)c; return ""; }
int pcu_visible_devices(void) { return 1; }
int pcu_rollout_random_tree(pcu_ctx *c, const pcu_rt_batch *b, pcu_model_rng *r, pcu_out *o) {
    (void)c; (void)r;
    if (o->payoff_x2) memset(o->payoff_x2, 1, b->n);
    if (o->sums) { memset(o->sums, 0, 64); o->sums[1] = b->n; }
    return 0;
}
int pcu_rollout_mold(pcu_ctx *c, const pcu_mold_batch *b, pcu_model_rng *r, pcu_out *o) {
    (void)c; (void)r;
    if (o->payoff_x2) memset(o->payoff_x2, 0, b->n);
    return 0;
}
