/* Profiling shim (NOT part of the product or the tests): a fake libpinyon_cuda front for the host-side batching
 * adapter, so that its overhead can be measured on a machine without a GPU.  LD_PRELOAD it in front of the real library. */
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
