// TEST INFRASTRUCTURE ONLY — stand-in for the C interface of baskuit/lrslib @ 2414808d (extern/lrslib/include/lib.h),
// the one native dependency of Pinyon, which is absent from the reference checkout (.SUBMODULES.json:8-15).
// It declares exactly what src/libpinyon/lrslib.hh:23-26,60-63,97-100,149-154 calls, over the mpz_t of the parse-only
// gmpxx.h stand-in (oracle/shim/gmpxx.h).  The definitions are in tests/cpp/lrs_shim/lrs_shim.cc: a Nash equilibrium
// of a small bimatrix game by support enumeration — NOT lrslib's vertex enumeration; the same solver serves the run
// over the reference model and the run over the GPU model of a parity test, so their results are comparable
// ("results at that boundary are unpinned").
#pragma once
#include <gmpxx.h>

mpz_t *alloc(size_t n);
void dealloc(mpz_t *data, size_t n);
/* payoff_data[2 i], payoff_data[2 i + 1]: row / column payoff of entry i (row-major), integers in [0, den].
 * row_solution_data[0] = common denominator, [1 .. rows] = strategy numerators, [rows + 1] = numerator of the
 * COLUMN player's equilibrium payoff / den; col_solution_data likewise with the ROW player's payoff. */
void solve_gmp_float(int rows, int cols, const long *payoff_data, long den, mpz_t *row_solution_data,
                     mpz_t *col_solution_data);
void solve_gmp_pointer(int rows, int cols, const mpq_t **row_payoff_data, const mpq_t **col_payoff_data,
                       mpz_t *row_solution_data, mpz_t *col_solution_data);
void solve_gmp_pointer_constant_sum(int rows, int cols, const mpq_t **row_payoff_data, mpz_t *row_solution_data,
                                    mpz_t *col_solution_data, int payoff_sum_num, int payoff_sum_den);
