// TEST INFRASTRUCTURE ONLY — parse-only stand-in for GMP's C++ header.
//
// The reference (baskuit/pinyon) includes <gmpxx.h> from types/random.hh:3,
// types/rational.hh:3 and types/value.hh:5, but the GMP development headers
// are absent from this image.  The type lists on the Monte-Carlo hot path
// (SimpleTypes, RandomTreeFloatTypes; types/types.hh:229-240) are double-typed
// and never execute an mpq operation, so a stub that merely lets the headers
// parse is enough to compile the reference's own MonteCarloModel unmodified.
// Every arithmetic entry point below works on a long double quotient, which is
// NOT GMP semantics; nothing on the measured path calls it.
#pragma once

#include <algorithm>
#include <cfloat>
#include <cstring>
#include <iostream>
#include <limits>
#include <stdexcept>
#include <string>
#include <utility>

struct pinyon_shim_mpz_struct { long v; };
struct pinyon_shim_mpq_struct { long num; long den; };
typedef pinyon_shim_mpz_struct mpz_t[1];
typedef pinyon_shim_mpq_struct mpq_t[1];
typedef pinyon_shim_mpz_struct *mpz_ptr;
typedef pinyon_shim_mpq_struct *mpq_ptr;

class mpz_class {
    mpz_t z_;

   public:
    mpz_class() { z_[0].v = 0; }
    mpz_class(long v) { z_[0].v = v; }
    mpz_class(int v) { z_[0].v = v; }
    mpz_class(unsigned long v) { z_[0].v = static_cast<long>(v); }
    mpz_class(double v) { z_[0].v = static_cast<long>(v); }
    mpz_class(const pinyon_shim_mpz_struct *z) { z_[0].v = z->v; } /* from an mpz_t, as libpinyon/lrslib.hh:157 does */
    long get_si() const { return z_[0].v; }
    double get_d() const { return static_cast<double>(z_[0].v); }
    std::string get_str() const { return std::to_string(z_[0].v); }
    mpz_ptr get_mpz_t() { return z_; }
    const pinyon_shim_mpz_struct *get_mpz_t() const { return z_; }
};

class mpq_class {
    mpq_t q_;

    static long gcd_(long a, long b) {
        a = a < 0 ? -a : a;
        b = b < 0 ? -b : b;
        while (b) {
            long c = a % b;
            a = b;
            b = c;
        }
        return a ? a : 1;
    }

   public:
    mpq_class() { q_[0] = {0, 1}; }
    mpq_class(int v) { q_[0] = {v, 1}; }
    mpq_class(long v) { q_[0] = {v, 1}; }
    mpq_class(unsigned long v) { q_[0] = {static_cast<long>(v), 1}; }
    mpq_class(double v) {
        long den = 1L << 30;
        q_[0] = {static_cast<long>(v * static_cast<double>(den)), den};
        canonicalize();
    }
    mpq_class(float v) : mpq_class(static_cast<double>(v)) {}
    mpq_class(long a, long b) { q_[0] = {a, b}; }
    mpq_class(int a, int b) { q_[0] = {a, b}; }
    mpq_class(const mpz_class &a, const mpz_class &b) { q_[0] = {a.get_si(), b.get_si()}; }
    mpq_class(const mpz_class &a) { q_[0] = {a.get_si(), 1}; }

    void canonicalize() {
        long g = gcd_(q_[0].num, q_[0].den);
        q_[0].num /= g;
        q_[0].den /= g;
        if (q_[0].den < 0) {
            q_[0].num = -q_[0].num;
            q_[0].den = -q_[0].den;
        }
    }
    double get_d() const { return static_cast<double>(q_[0].num) / static_cast<double>(q_[0].den); }
    std::string get_str() const { return std::to_string(q_[0].num) + "/" + std::to_string(q_[0].den); }
    mpq_ptr get_mpq_t() { return q_; }
    const pinyon_shim_mpq_struct *get_mpq_t() const { return q_; }
    mpz_class get_num() const { return mpz_class{q_[0].num}; }
    mpz_class get_den() const { return mpz_class{q_[0].den}; }

    explicit operator double() const { return get_d(); }
    explicit operator float() const { return static_cast<float>(get_d()); }

    friend mpq_class operator+(const mpq_class &a, const mpq_class &b) {
        mpq_class r{a.q_[0].num * b.q_[0].den + b.q_[0].num * a.q_[0].den, a.q_[0].den * b.q_[0].den};
        r.canonicalize();
        return r;
    }
    friend mpq_class operator-(const mpq_class &a, const mpq_class &b) {
        mpq_class r{a.q_[0].num * b.q_[0].den - b.q_[0].num * a.q_[0].den, a.q_[0].den * b.q_[0].den};
        r.canonicalize();
        return r;
    }
    friend mpq_class operator*(const mpq_class &a, const mpq_class &b) {
        mpq_class r{a.q_[0].num * b.q_[0].num, a.q_[0].den * b.q_[0].den};
        r.canonicalize();
        return r;
    }
    friend mpq_class operator/(const mpq_class &a, const mpq_class &b) {
        mpq_class r{a.q_[0].num * b.q_[0].den, a.q_[0].den * b.q_[0].num};
        r.canonicalize();
        return r;
    }
    mpq_class operator-() const { return mpq_class{-q_[0].num, q_[0].den}; }
    mpq_class &operator+=(const mpq_class &b) { return *this = *this + b; }
    mpq_class &operator-=(const mpq_class &b) { return *this = *this - b; }
    mpq_class &operator*=(const mpq_class &b) { return *this = *this * b; }
    mpq_class &operator/=(const mpq_class &b) { return *this = *this / b; }

    friend bool operator==(const mpq_class &a, const mpq_class &b) {
        return a.q_[0].num * b.q_[0].den == b.q_[0].num * a.q_[0].den;
    }
    friend bool operator!=(const mpq_class &a, const mpq_class &b) { return !(a == b); }
    friend bool operator<(const mpq_class &a, const mpq_class &b) {
        return a.q_[0].num * b.q_[0].den < b.q_[0].num * a.q_[0].den;
    }
    friend bool operator>(const mpq_class &a, const mpq_class &b) { return b < a; }
    friend bool operator<=(const mpq_class &a, const mpq_class &b) { return !(b < a); }
    friend bool operator>=(const mpq_class &a, const mpq_class &b) { return !(a < b); }
    friend std::ostream &operator<<(std::ostream &os, const mpq_class &q) { return os << q.get_str(); }
};
