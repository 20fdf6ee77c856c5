// TEST INFRASTRUCTURE ONLY — see extern/lrslib/include/lib.h next to this file.
//
// One Nash equilibrium of an m x n bimatrix game (A, B) by support enumeration (Porter, Nudelman & Shoham 2004):
// for supports I, J of equal size k solve  sum_{j in J} y_j A[i][j] = u (i in I), sum y = 1  and the mirror system
// for x, and accept when x, y >= 0 and no action outside the supports does better.  Payoff ties are everywhere in
// RandomTree games (payoffs are 0, 1/2, 1), so the search runs on a copy perturbed by a fixed tiny pattern, which
// makes the game nondegenerate and the result deterministic; the equilibrium is then re-solved on the original
// payoffs for the supports found.  Strategies are returned as numerators over 2^40.
#include <cmath>
#include <cstddef>
#include <vector>

#include "extern/lrslib/include/lib.h"

namespace {

typedef long double R;

bool solve_linear(std::vector<std::vector<R>> &a, std::vector<R> &x) { /* Gaussian elimination, partial pivoting */
    const size_t n = a.size();
    for (size_t c = 0; c < n; ++c) {
        size_t piv = c;
        for (size_t r = c + 1; r < n; ++r)
            if (std::fabs(a[r][c]) > std::fabs(a[piv][c])) piv = r;
        if (std::fabs(a[piv][c]) < 1e-13L) return false;
        std::swap(a[piv], a[c]);
        for (size_t r = 0; r < n; ++r) {
            if (r == c) continue;
            const R f = a[r][c] / a[c][c];
            if (f == 0) continue;
            for (size_t k = c; k <= n; ++k) a[r][k] -= f * a[c][k];
        }
    }
    x.resize(n);
    for (size_t r = 0; r < n; ++r) x[r] = a[r][n] / a[r][r];
    return true;
}

/* Mixed strategy of the OTHER player over `mine` that makes this player indifferent between the actions in `rows`:
 * M[r][c] = payoff of this player's action r against the other's action c. */
bool indifference(const std::vector<std::vector<R>> &M, const std::vector<int> &rows, const std::vector<int> &cols,
                  std::vector<R> &strategy, R &value) {
    const size_t k = rows.size();
    std::vector<std::vector<R>> a(k + 1, std::vector<R>(k + 2, 0));
    for (size_t i = 0; i < k; ++i) {
        for (size_t j = 0; j < k; ++j) a[i][j] = M[rows[i]][cols[j]];
        a[i][k] = -1; /* - u */
    }
    for (size_t j = 0; j < k; ++j) a[k][j] = 1;
    a[k][k + 1] = 1;
    std::vector<R> sol;
    if (!solve_linear(a, sol)) return false;
    strategy.assign(sol.begin(), sol.begin() + k);
    value = sol[k];
    return true;
}

struct Equilibrium {
    std::vector<R> x, y;
    R row_payoff = 0, col_payoff = 0;
    bool ok = false;
};

bool next_subset(std::vector<int> &s, int n) {
    const int k = (int)s.size();
    for (int i = k - 1; i >= 0; --i)
        if (s[i] < n - k + i) {
            ++s[i];
            for (int j = i + 1; j < k; ++j) s[j] = s[j - 1] + 1;
            return true;
        }
    return false;
}

Equilibrium enumerate(const std::vector<std::vector<R>> &A, const std::vector<std::vector<R>> &Bt, int m, int n, R tol,
                      std::vector<int> *I_out, std::vector<int> *J_out) {
    /* A[i][j]: row player's payoff; Bt[j][i]: column player's payoff */
    Equilibrium e;
    for (int k = 1; k <= std::min(m, n); ++k) {
        std::vector<int> I(k);
        for (int i = 0; i < k; ++i) I[i] = i;
        do {
            std::vector<int> J(k);
            for (int j = 0; j < k; ++j) J[j] = j;
            do {
                std::vector<R> y, x;
                R u, v;
                if (!indifference(A, I, J, y, u) || !indifference(Bt, J, I, x, v)) continue;
                bool good = true;
                for (R p : y) good = good && p >= -tol;
                for (R p : x) good = good && p >= -tol;
                if (!good) continue;
                for (int i = 0; i < m && good; ++i) { /* no row outside I does better against y */
                    R w = 0;
                    for (int j = 0; j < k; ++j) w += A[i][J[j]] * y[j];
                    good = w <= u + tol;
                }
                for (int j = 0; j < n && good; ++j) {
                    R w = 0;
                    for (int i = 0; i < k; ++i) w += Bt[j][I[i]] * x[i];
                    good = w <= v + tol;
                }
                if (!good) continue;
                e.x.assign(m, 0);
                e.y.assign(n, 0);
                for (int i = 0; i < k; ++i) e.x[I[i]] = x[i] < 0 ? 0 : x[i];
                for (int j = 0; j < k; ++j) e.y[J[j]] = y[j] < 0 ? 0 : y[j];
                e.row_payoff = u;
                e.col_payoff = v;
                e.ok = true;
                if (I_out) *I_out = I;
                if (J_out) *J_out = J;
                return e;
            } while (next_subset(J, n));
        } while (next_subset(I, m));
    }
    return e;
}

Equilibrium nash(const std::vector<std::vector<R>> &A, const std::vector<std::vector<R>> &B, int m, int n) {
    std::vector<std::vector<R>> Ap(m, std::vector<R>(n)), Btp(n, std::vector<R>(m)), Bt(n, std::vector<R>(m));
    R scale = 1;
    for (int i = 0; i < m; ++i)
        for (int j = 0; j < n; ++j) scale = std::max(scale, std::max(std::fabs(A[i][j]), std::fabs(B[i][j])));
    for (int i = 0; i < m; ++i)
        for (int j = 0; j < n; ++j) {
            const R eps = scale * 1e-7L;
            Ap[i][j] = A[i][j] + eps * (R)((i * 37 + j * 101 + 13) % 211) / 211;
            Btp[j][i] = B[i][j] + eps * (R)((i * 59 + j * 43 + 7) % 223) / 223;
            Bt[j][i] = B[i][j];
        }
    std::vector<int> I, J;
    Equilibrium e = enumerate(Ap, Btp, m, n, scale * 1e-12L, &I, &J);
    if (!e.ok) { /* cannot happen for a nondegenerate game; keep the caller alive with the uniform profile */
        e.x.assign(m, (R)1 / m);
        e.y.assign(n, (R)1 / n);
        e.ok = true;
    }
    /* payoffs of the profile on the ORIGINAL matrices */
    e.row_payoff = e.col_payoff = 0;
    for (int i = 0; i < m; ++i)
        for (int j = 0; j < n; ++j) {
            e.row_payoff += e.x[i] * e.y[j] * A[i][j];
            e.col_payoff += e.x[i] * e.y[j] * B[i][j];
        }
    return e;
}

const long DEN = 1L << 40;

void write_solution(const Equilibrium &e, int m, int n, R payoff_scale, mpz_t *row_sol, mpz_t *col_sol) {
    row_sol[0][0].v = DEN;
    col_sol[0][0].v = DEN;
    for (int i = 0; i < m; ++i) row_sol[i + 1][0].v = (long)std::llround((double)(e.x[i] * DEN));
    for (int j = 0; j < n; ++j) col_sol[j + 1][0].v = (long)std::llround((double)(e.y[j] * DEN));
    /* lrslib.hh reads the ROW payoff from the column solution and vice versa (:171-172) */
    col_sol[n + 1][0].v = (long)std::llround((double)(e.row_payoff / payoff_scale * DEN));
    row_sol[m + 1][0].v = (long)std::llround((double)(e.col_payoff / payoff_scale * DEN));
}

}  // namespace

mpz_t *alloc(size_t n) { return new mpz_t[n](); }
void dealloc(mpz_t *data, size_t) { delete[] data; }

void solve_gmp_float(int rows, int cols, const long *payoff_data, long den, mpz_t *row_sol, mpz_t *col_sol) {
    std::vector<std::vector<R>> A(rows, std::vector<R>(cols)), B(rows, std::vector<R>(cols));
    for (int i = 0; i < rows; ++i)
        for (int j = 0; j < cols; ++j) {
            A[i][j] = (R)payoff_data[2 * (i * cols + j)];
            B[i][j] = (R)payoff_data[2 * (i * cols + j) + 1];
        }
    write_solution(nash(A, B, rows, cols), rows, cols, (R)den, row_sol, col_sol);
}

void solve_gmp_pointer(int rows, int cols, const mpq_t **rp, const mpq_t **cp, mpz_t *row_sol, mpz_t *col_sol) {
    std::vector<std::vector<R>> A(rows, std::vector<R>(cols)), B(rows, std::vector<R>(cols));
    for (int i = 0; i < rows; ++i)
        for (int j = 0; j < cols; ++j) {
            const int e = i * cols + j;
            A[i][j] = (R)(*rp[e])[0].num / (R)(*rp[e])[0].den;
            B[i][j] = (R)(*cp[e])[0].num / (R)(*cp[e])[0].den;
        }
    write_solution(nash(A, B, rows, cols), rows, cols, 1, row_sol, col_sol);
}

void solve_gmp_pointer_constant_sum(int rows, int cols, const mpq_t **rp, mpz_t *row_sol, mpz_t *col_sol, int sum_num,
                                    int sum_den) {
    std::vector<std::vector<R>> A(rows, std::vector<R>(cols)), B(rows, std::vector<R>(cols));
    for (int i = 0; i < rows; ++i)
        for (int j = 0; j < cols; ++j) {
            const int e = i * cols + j;
            A[i][j] = (R)(*rp[e])[0].num / (R)(*rp[e])[0].den;
            B[i][j] = (R)sum_num / (R)sum_den - A[i][j];
        }
    write_solution(nash(A, B, rows, cols), rows, cols, 1, row_sol, col_sol);
}
