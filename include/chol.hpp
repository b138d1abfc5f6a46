// chol.hpp -- drop-in host API of the B200-native sparse Cholesky (C++23).
//
// Same public names, types and error behaviour as the reference's include/chol.hpp
// (evanwporter/SparseCholesky), so callers such as the reference's tests/test_chol.cpp and the
// README snippet compile unchanged:
//
//     auto A = triplet_to_csc_matrix(ti, tj, tx, n);     // reference chol.hpp:309
//     auto S = schol(A);                                   // reference chol.hpp:874
//     auto L = chol(A, S).value();                         // reference README.md:26-27
//     auto L2 = chol(A);                                   // reference chol.hpp:750
//
// Nothing numeric happens in this header: it forwards to the C-ABI of libspchol.so
// (include/spchol.h), whose numeric phase runs hand-written sm_100a kernels.  There is no CPU
// fallback; without a GPU chol() returns std::unexpected with the library's error text.
// This file is written from scratch against the C-ABI; it contains no reference code.
#pragma once

#include <algorithm>
#include <cassert>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <expected>
#include <memory>
#include <random>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "spchol.h"

// The reference's header declares LAPACK's dpotrf_ (include/chol.hpp:20-22) and its tests call it through that
// declaration (tests/test_chol.cpp:73,113) as the dense cross-check; callers link their own LAPACK as before.
// Nothing in this header or in libspchol.so calls it.
extern "C" {
void dpotrf_(const char* uplo, const int* n, double* a, const int* lda, int* info);
}

enum class sym { none, upper, lower };          // reference chol.hpp:26

using elimination_tree = std::vector<int>;      // reference chol.hpp:36

namespace internal {
// reference chol.hpp:39-96
class csc_storage {
protected:
    std::size_t m_ = 0, n_ = 0, nnz_ = 0;
    std::vector<int> p_, i_;
    csc_storage() = default;
    csc_storage(std::size_t m, std::size_t n, std::size_t nnz) : m_(m), n_(n), nnz_(nnz), p_(n + 1, 0), i_(nnz) {}

public:
    std::size_t rows() const { return m_; }
    std::size_t cols() const { return n_; }
    std::size_t size() const { return n_; }
    std::size_t capacity() const { return nnz_; }
    std::vector<int>& p() { return p_; }
    std::vector<int>& i() { return i_; }
    const std::vector<int>& p() const { return p_; }
    const std::vector<int>& i() const { return i_; }
    // index of entry (i, j) in the column-compressed arrays, -1 if absent (binary search in column j)
    int find_index(std::size_t i, std::size_t j) const {
        const auto b = i_.begin() + p_[j], e = i_.begin() + p_[j + 1];
        const auto it = std::lower_bound(b, e, static_cast<int>(i));
        return (it != e && *it == static_cast<int>(i)) ? static_cast<int>(it - i_.begin()) : -1;
    }
};
struct symbolic_deleter { void operator()(spchol_symbolic* s) const { spchol_symbolic_destroy(s); } };
struct plan_deleter { void operator()(spchol_plan* p) const { spchol_plan_destroy(p); } };
}  // namespace internal

// reference chol.hpp:99-132 -- the pattern of L plus the elimination tree.  New: opaque handles so
// chol(A, S) reuses the symbolic analysis and the device-resident plan across calls.
struct SChol : public internal::csc_storage {
    SChol(std::size_t n, std::size_t nnz) : internal::csc_storage(n, n, nnz) { assert(n > 0 && nnz > 0); }
    SChol() = default;

    std::vector<int> parent;

    bool operator[](int i, int j) const {
        if (i < j) std::swap(i, j);
        return find_index(i, j) != -1;
    }
    void set_capacity(std::size_t nnz) { nnz_ = nnz; i_.resize(nnz_); }
    std::size_t size() const { return p_.size() - 1; }
    std::size_t capacity() const { return i_.size(); }

    std::shared_ptr<spchol_symbolic> symbolic;             // host analysis (supernodes, row lists)
    mutable std::shared_ptr<spchol_plan> plan;             // device plan, created on first chol(A, S)
    mutable int plan_device = -1;
};

// reference chol.hpp:134-299
template <typename T, sym S = sym::upper>
class csc_matrix : public internal::csc_storage {
    std::vector<T> x_;

    T* find_entry(std::size_t i, std::size_t j) {
        if constexpr (S == sym::upper) { if (j < i) std::swap(i, j); }
        else if constexpr (S == sym::lower) { if (i < j) std::swap(i, j); }
        const int idx = find_index(i, j);
        return idx < 0 ? nullptr : &x_[idx];
    }

public:
    csc_matrix(std::size_t m, std::size_t n, std::size_t nnz) : internal::csc_storage(m, n, nnz), x_(nnz, T{}) {
        assert(m > 0 && n > 0 && nnz > 0);
        if constexpr (S != sym::none) assert(m == n);
    }
    csc_matrix(const SChol& s) : internal::csc_storage(s.size(), s.size(), s.capacity()), x_(s.capacity(), T{}) {
        p_ = s.p(); i_ = s.i();
    }
    csc_matrix(std::size_t n, std::size_t nnz) requires(S != sym::none)
        : internal::csc_storage(n, n, nnz), x_(nnz, T{}) { assert(n > 0 && nnz > 0); }
    // reference chol.hpp:190-217: the pattern of `schol`, holding A's value wherever A stores that very (row, column)
    // slot and 0 in the fill-in slots.  As in the reference the look-up is the raw find_index(row, col) -- no swap by
    // the symmetry tag -- so from an upper-stored A only the diagonal carries over.
    template <sym S_A>
    csc_matrix(const csc_matrix<T, S_A>& A, const SChol& schol)
        : internal::csc_storage(schol.size(), schol.size(), schol.capacity()), x_(schol.capacity(), T(0)) {
        p_ = schol.p(); i_ = schol.i();
        for (int j = 0; j < static_cast<int>(A.cols()); ++j)
            for (int q = p_[j]; q < p_[j + 1]; ++q) {
                const int at = A.find_index(static_cast<std::size_t>(i_[q]), static_cast<std::size_t>(j));
                x_[q] = at != -1 ? A.x()[at] : T(0);
            }
    }

    T& operator[](std::size_t i, std::size_t j) {
        assert(i < m_ && j < n_);
        T* e = find_entry(i, j);
        if (!e) throw std::out_of_range("Element not present in CSC structure");
        return *e;
    }
    const T operator[](std::size_t i, std::size_t j) const {
        assert(i < m_ && j < n_);
        const T* e = const_cast<csc_matrix*>(this)->find_entry(i, j);
        return e ? *e : T(0);
    }
    std::vector<T>& x() { return x_; }
    const std::vector<T>& x() const { return x_; }

    auto transpose() const {
        const int m = static_cast<int>(m_), n = static_cast<int>(n_);
        const int nz = p_[n];
        std::vector<int> tp(m + 1, 0), ti(nz);
        std::vector<T> tx(nz);
        for (int k = 0; k < nz; ++k) ++tp[i_[k] + 1];
        for (int r = 0; r < m; ++r) tp[r + 1] += tp[r];
        std::vector<int> head(tp.begin(), tp.end() - 1);
        for (int j = 0; j < n; ++j)
            for (int k = p_[j]; k < p_[j + 1]; ++k) { const int d = head[i_[k]]++; ti[d] = j; tx[d] = x_[k]; }
        constexpr sym ST = S == sym::upper ? sym::lower : S == sym::lower ? sym::upper : sym::none;
        csc_matrix<T, ST> At(n, m, std::max(nz, 1));
        At.p() = std::move(tp); At.i() = std::move(ti); At.x() = std::move(tx);
        return At;
    }
};

namespace internal {
inline void throw_last(const char* what) { throw std::runtime_error(std::string(what) + ": " + spchol_last_error()); }
}

// reference chol.hpp:309-369: COO -> upper CSC, duplicates summed, explicit zeros kept.
template <typename T>
csc_matrix<T, sym::upper> triplet_to_csc_matrix(const std::vector<int>& ti, const std::vector<int>& tj,
                                                const std::vector<T>& tx, int n) {
    static_assert(std::is_same_v<T, double>, "the B200 path is fp64 (the only instantiation the reference uses)");
    assert(ti.size() == tj.size() && tj.size() == tx.size());
    const std::int64_t nz = static_cast<std::int64_t>(ti.size());
    std::vector<int> Ap(n + 1), Ai(std::max<std::int64_t>(nz, 1));
    std::vector<double> Ax(std::max<std::int64_t>(nz, 1));
    std::int64_t m = 0;
    if (spchol_triplet_to_csc(nz, ti.data(), tj.data(), tx.data(), n, Ap.data(), Ai.data(), Ax.data(), &m) != SPCHOL_OK)
        internal::throw_last("triplet_to_csc_matrix");
    Ai.resize(m); Ax.resize(m);
    csc_matrix<T, sym::upper> A(n, std::max<std::int64_t>(m, 1));
    A.p() = std::move(Ap); A.i() = std::move(Ai); A.x() = std::move(Ax);
    return A;
}

// reference chol.hpp:416-435
template <typename T>
csc_matrix<T, sym::upper> build_csc_matrix_from_pattern(const std::vector<std::vector<int>>& pattern) {
    std::vector<int> ti, tj; std::vector<T> tx;
    for (int i = 0; i < static_cast<int>(pattern.size()); ++i)
        for (int c : pattern[i]) { ti.push_back(std::min(i, c)); tj.push_back(std::max(i, c)); tx.push_back(T(1)); }
    return triplet_to_csc_matrix(ti, tj, tx, static_cast<int>(pattern.size()));
}

// reference chol.hpp:378 / :466 / :567 / :725 / :737 / :741 / :954 / :1221
template <typename T>
elimination_tree etree(const csc_matrix<T, sym::upper>& A) {
    elimination_tree parent(A.size());
    spchol_etree(static_cast<int>(A.size()), A.p().data(), A.i().data(), parent.data());
    return parent;
}
inline std::vector<int> post_order(const elimination_tree& parent) {
    std::vector<int> post(parent.size());
    spchol_post_order(static_cast<int>(parent.size()), parent.data(), post.data());
    return post;
}
template <typename T>
std::vector<int> col_count(const csc_matrix<T>& A, const std::vector<int>& parent, const std::vector<int>& post) {
    std::vector<int> cc(A.size());
    spchol_col_count(static_cast<int>(A.size()), A.p().data(), A.i().data(), parent.data(), post.data(), cc.data());
    return cc;
}
template <typename T>
std::size_t ereach(const csc_matrix<T>& A, std::size_t k, const std::vector<int>& parent, std::vector<int>& s,
                   std::vector<int>& w, std::vector<T>& x, std::size_t top) {
    return spchol_ereach(static_cast<int>(A.size()), A.p().data(), A.i().data(), A.x().data(), static_cast<int>(k),
                         parent.data(), s.data(), w.data(), x.data(), static_cast<int>(top));
}
template <typename T>
std::size_t ereach(const csc_matrix<T>& A, std::size_t k, const std::vector<int>& parent, std::vector<int>& s,
                   std::vector<int>& w, std::size_t top) {
    return spchol_ereach(static_cast<int>(A.size()), A.p().data(), A.i().data(), nullptr, static_cast<int>(k),
                         parent.data(), s.data(), w.data(), nullptr, static_cast<int>(top));
}
inline std::vector<std::vector<int>> compute_levels(const std::vector<int>& parent) {
    const int n = static_cast<int>(parent.size());
    std::vector<int> lp(n + 1), ln(n);
    const int nl = spchol_compute_levels(n, parent.data(), lp.data(), ln.data());
    std::vector<std::vector<int>> out(nl);
    for (int l = 0; l < nl; ++l) out[l].assign(ln.begin() + lp[l], ln.begin() + lp[l + 1]);
    return out;
}
inline std::vector<int> compute_supernodes(const SChol& S, std::vector<std::size_t>& supernodes) {
    const int n = static_cast<int>(S.parent.size());
    std::vector<std::int64_t> Lp(S.p().begin(), S.p().end());
    std::vector<int> sn_id(n), st(n + 1);
    const int ns = spchol_compute_supernodes(n, S.parent.data(), Lp.data(), sn_id.data(), st.data());
    supernodes.assign(st.begin(), st.begin() + ns + 1);
    return sn_id;
}
inline std::vector<int> atree(const SChol& S, const std::vector<int>& sn_id, const std::vector<std::size_t>& supernodes) {
    const int ns = static_cast<int>(supernodes.size()) - 1;
    std::vector<int> st(supernodes.begin(), supernodes.end()), sp(ns);
    spchol_atree(static_cast<int>(S.parent.size()), S.parent.data(), sn_id.data(), ns, st.data(), sp.data());
    return sp;
}

// reference chol.hpp:874-946: symbolic factorization (pattern of L + etree).
template <typename T>
SChol schol(const csc_matrix<T, sym::upper>& A) {
    spchol_symbolic* h = nullptr;
    if (spchol_symbolic_create(static_cast<int>(A.size()), A.p().data(), A.i().data(), &h) != SPCHOL_OK)
        internal::throw_last("schol");
    std::shared_ptr<spchol_symbolic> sp(h, internal::symbolic_deleter{});
    const int n = static_cast<int>(A.size());
    SChol S(n, static_cast<std::size_t>(spchol_symbolic_nnzL(h)));
    S.symbolic = sp;
    S.parent.resize(n);
    spchol_symbolic_parent(h, S.parent.data());
    if (spchol_symbolic_Lp(h, S.p().data()) != SPCHOL_OK || spchol_symbolic_Li(h, S.i().data()) != SPCHOL_OK)
        internal::throw_last("schol");      // nnz(L) > INT_MAX cannot be carried by the reference's 32-bit CSC
    return S;
}

// README.md:26-27 (documented) -- numeric factorization reusing S.
template <typename T>
std::expected<csc_matrix<T, sym::none>, std::string> chol(const csc_matrix<T, sym::upper>& A, const SChol& S, int device = 0) {
    static_assert(std::is_same_v<T, double>, "the B200 path is fp64");
    if (!S.symbolic) return std::unexpected(std::string("chol: S was not produced by schol()"));
    if (A.size() != S.size() || static_cast<std::int64_t>(A.p()[A.size()]) != spchol_symbolic_nnzA(S.symbolic.get()))
        return std::unexpected(std::string("chol: A does not have the pattern S was built from"));
    if (!S.plan || S.plan_device != device) {
        spchol_plan* p = nullptr;
        if (spchol_plan_create(S.symbolic.get(), device, &p) != SPCHOL_OK) return std::unexpected(std::string(spchol_last_error()));
        S.plan.reset(p, internal::plan_deleter{});
        S.plan_device = device;
    }
    csc_matrix<T, sym::none> L(S);
    int info = 0;
    if (spchol_factor(S.plan.get(), A.x().data(), L.x().data(), &info) != SPCHOL_OK)
        return std::unexpected(std::string(spchol_last_error()));
    if (info > 0) return std::unexpected(std::string("A is not positive definite."));      // reference chol.hpp:850
    return L;
}

// reference chol.hpp:750 (actual signature): symbolic + numeric.
template <typename T>
std::expected<csc_matrix<T, sym::none>, std::string> chol(const csc_matrix<T, sym::upper>& A) {
    return chol(A, schol(A));
}

// reference chol.hpp:1407-1446: the supernodal entry point.  The reference's own body is unfinished (it returns
// diag(sqrt(a_ii)) scattered into wrong slots); here it is the same supernodal GPU factorization
// chol() runs -- the thing that function was meant to be -- so tests/test_chol.cpp:99-136 passes.
template <typename T>
std::expected<csc_matrix<T, sym::none>, std::string> chol_sn(csc_matrix<T, sym::upper>& A) {
    return chol(A);
}

// reference chol.hpp:628-678: pattern of L from one ereach per column.  As in the reference the column's own marker is
// not preset, so the walk runs on past j and the column also lists the etree ancestors of j (an upper bound of the
// true pattern: 1172 vs 877 entries on bcsstk01); kept for callers that use it, schol() is the exact one.
template <typename T>
SChol schol_etree(const csc_matrix<T, sym::upper>& A) {
    const int n = static_cast<int>(A.size());
    SChol S;
    S.parent = etree(A);
    auto& Sp = S.p();
    Sp.assign(n + 1, 0);
    std::vector<int> s(n), w(n, -1), rows;
    std::vector<T> x(n, T(0));
    for (int j = 0; j < n; ++j) {
        const std::size_t top = ereach(A, static_cast<std::size_t>(j), S.parent, s, w, x, static_cast<std::size_t>(n));
        const std::size_t start = rows.size();
        rows.push_back(j);
        for (std::size_t t = top; t < static_cast<std::size_t>(n); ++t) if (s[t] > j) rows.push_back(s[t]);
        std::sort(rows.begin() + static_cast<std::ptrdiff_t>(start), rows.end());
        Sp[j + 1] = static_cast<int>(rows.size());
    }
    S.i() = std::move(rows);
    S.set_capacity(S.i().size());
    return S;
}

namespace internal {
// pcg32 (PCG XSH-RR 64/32, default stream) -- the generator the reference takes from its pcg-cpp submodule
// (pcg_random.hpp, `pcg32 gen(seed)`): state = (seed + inc) * mult + inc, output from the state before the step.
class pcg32_engine {
    std::uint64_t state_;
    static constexpr std::uint64_t mult = 6364136223846793005ULL, inc = 1442695040888963407ULL;
public:
    using result_type = std::uint32_t;
    explicit pcg32_engine(std::uint64_t seed) : state_((seed + inc) * mult + inc) {}
    static constexpr result_type min() { return 0u; }
    static constexpr result_type max() { return 0xffffffffu; }
    result_type operator()() {
        const std::uint64_t old = state_;
        state_ = old * mult + inc;
        const std::uint32_t xs = static_cast<std::uint32_t>(((old >> 18u) ^ old) >> 27u), rot = static_cast<std::uint32_t>(old >> 59u);
        return (xs >> rot) | (xs << ((32u - rot) & 31u));
    }
};
inline int iabs(double v) { const int t = static_cast<int>(v); return t < 0 ? -t : t; }      // int abs(int) on a double argument
}  // namespace internal

// reference chol.hpp:964-996: random symmetric sparse matrix (upper stored), seed 21, diagonal kept and boosted by n.
// The reference writes an unqualified abs(v) here and in the overload below; in its translation units that resolves to the
// C library's int abs(int), i.e. the value is truncated towards zero first.  Reproduced on purpose (internal::iabs): the
// matrices come out bit for bit as the reference's (tests/golden/boundary.npz).
template <typename T>
csc_matrix<T, sym::upper> random_sparse(int n, double density = 0.25, bool positive_definite = true) {
    internal::pcg32_engine gen(21);
    std::uniform_real_distribution<double> dist(-n, n);
    std::vector<int> ti, tj;
    std::vector<T> tx;
    for (int i = 0; i < n; ++i)
        for (int j = i; j < n; ++j) {
            const double v = dist(gen);
            if (internal::iabs(v) < density * n || i == j) {
                T val = static_cast<T>(dist(gen));
                if (i == j && positive_definite) val += static_cast<T>(n);
                ti.push_back(i); tj.push_back(j); tx.push_back(val);
            }
        }
    return triplet_to_csc_matrix(ti, tj, tx, n);
}

// reference chol.hpp:1004-1034: random SPD values on the pattern of S (returned lower, callers transpose), seed 42.
template <typename T>
csc_matrix<T, sym::lower> random_sparse(const SChol& S, bool positive_definite = true) {
    const int n = static_cast<int>(S.size());
    csc_matrix<T, sym::lower> A(S);
    internal::pcg32_engine gen(42);
    std::uniform_real_distribution<double> dist(-1.0, 1.0);
    for (int j = 0; j < n; ++j)
        for (int q = S.p()[j]; q < S.p()[j + 1]; ++q)
            A.x()[q] = S.i()[q] == j ? static_cast<T>(1.5 * (std::fabs(dist(gen)) + (positive_definite ? n : 0)))
                                     : static_cast<T>(internal::iabs(static_cast<T>(dist(gen)) * 10));
    return A;
}

// ---- beyond the reference: factors with more than INT_MAX nonzeros --------------------------------
// The reference's containers are 32-bit (std::vector<int> p_, i_; chol.hpp:52,55; the nnz accumulators at :765, :890
// overflow at the 128^3 Laplacian, nnz(L) = 2.83e9).  SChol64 is the symbolic handle without the 32-bit pattern
// arrays, csc_matrix64 the same CSC layout with 64-bit column pointers, chol64 the numeric factorization into it;
// chol_resident keeps the factor on the device (for chol_solve / chol_residual) when even that is too much to move.
struct SChol64 {
    std::size_t n = 0;
    std::int64_t nnz = 0;
    std::vector<int> parent;
    std::shared_ptr<spchol_symbolic> symbolic;
    mutable std::shared_ptr<spchol_plan> plan;
    mutable int plan_device = -1;
    std::size_t size() const { return n; }
    std::int64_t capacity() const { return nnz; }
};
template <typename T>
class csc_matrix64 {
    std::size_t n_ = 0;
    std::vector<std::int64_t> p_;
    std::vector<int> i_;
    std::vector<T> x_;
public:
    csc_matrix64() = default;
    csc_matrix64(std::size_t n, std::int64_t nnz) : n_(n), p_(n + 1, 0), i_(static_cast<std::size_t>(nnz)), x_(static_cast<std::size_t>(nnz)) {}
    std::size_t rows() const { return n_; }
    std::size_t cols() const { return n_; }
    std::size_t size() const { return n_; }
    std::int64_t capacity() const { return static_cast<std::int64_t>(i_.size()); }
    std::vector<std::int64_t>& p() { return p_; }
    std::vector<int>& i() { return i_; }
    std::vector<T>& x() { return x_; }
    const std::vector<std::int64_t>& p() const { return p_; }
    const std::vector<int>& i() const { return i_; }
    const std::vector<T>& x() const { return x_; }
    std::int64_t find_index(std::size_t i, std::size_t j) const {
        const auto b = i_.begin() + p_[j], e = i_.begin() + p_[j + 1];
        const auto it = std::lower_bound(b + 1 <= e ? b + 1 : e, e, static_cast<int>(i));      // diagonal first, then ascending
        if (b != e && *b == static_cast<int>(i)) return p_[j];
        return (it != e && *it == static_cast<int>(i)) ? static_cast<std::int64_t>(it - i_.begin()) : -1;
    }
    const T operator[](std::size_t i, std::size_t j) const { const auto q = find_index(i, j); return q < 0 ? T(0) : x_[static_cast<std::size_t>(q)]; }
};
template <typename T>
SChol64 schol64(const csc_matrix<T, sym::upper>& A) {
    spchol_symbolic* h = nullptr;
    if (spchol_symbolic_create(static_cast<int>(A.size()), A.p().data(), A.i().data(), &h) != SPCHOL_OK) internal::throw_last("schol64");
    SChol64 S;
    S.symbolic.reset(h, internal::symbolic_deleter{});
    S.n = A.size(); S.nnz = spchol_symbolic_nnzL(h);
    S.parent.resize(S.n);
    spchol_symbolic_parent(h, S.parent.data());
    return S;
}
namespace internal {
template <typename T, typename SC>
std::expected<bool, std::string> ensure_plan(const csc_matrix<T, sym::upper>& A, const SC& S, int device) {
    if (!S.symbolic) return std::unexpected(std::string("chol: S was not produced by schol()"));
    if (A.size() != S.size() || static_cast<std::int64_t>(A.p()[A.size()]) != spchol_symbolic_nnzA(S.symbolic.get()))
        return std::unexpected(std::string("chol: A does not have the pattern S was built from"));
    if (!S.plan || S.plan_device != device) {
        spchol_plan* p = nullptr;
        if (spchol_plan_create(S.symbolic.get(), device, &p) != SPCHOL_OK) return std::unexpected(std::string(spchol_last_error()));
        S.plan.reset(p, plan_deleter{});
        S.plan_device = device;
    }
    return true;
}
}  // namespace internal
// numeric factorization into the 64-bit CSC: the supernodal panels come back through spchol_factor_lbuf (read back
// while the upper levels still compute when the buffer is pinned) and are gathered on the host
template <typename T>
std::expected<csc_matrix64<T>, std::string> chol64(const csc_matrix<T, sym::upper>& A, const SChol64& S, int device = 0) {
    static_assert(std::is_same_v<T, double>, "the B200 path is fp64");
    if (auto ok = internal::ensure_plan(A, S, device); !ok) return std::unexpected(ok.error());
    std::int64_t entries = 0;
    spchol_symbolic_panel_layout(S.symbolic.get(), nullptr, nullptr, nullptr, nullptr, &entries);
    void* host = nullptr;
    if (spchol_host_alloc(entries * 8, &host) != SPCHOL_OK) return std::unexpected(std::string(spchol_last_error()));
    std::unique_ptr<void, void (*)(void*)> guard(host, [](void* q) { spchol_host_free(q); });
    int info = 0;
    if (spchol_factor_lbuf(S.plan.get(), A.x().data(), static_cast<double*>(host), &info) != SPCHOL_OK)
        return std::unexpected(std::string(spchol_last_error()));
    if (info > 0) return std::unexpected(std::string("A is not positive definite."));      // reference chol.hpp:850
    csc_matrix64<T> L(S.n, S.nnz);
    if (spchol_symbolic_gather_csc64(S.symbolic.get(), static_cast<const double*>(host), L.p().data(), L.i().data(), L.x().data()) != SPCHOL_OK)
        return std::unexpected(std::string(spchol_last_error()));
    return L;
}
// the handle-returning form: the factor stays on the device in S's plan; chol_solve / chol_residual consume it
template <typename T, typename SC>
std::expected<bool, std::string> chol_resident(const csc_matrix<T, sym::upper>& A, const SC& S, int device = 0) {
    static_assert(std::is_same_v<T, double>, "the B200 path is fp64");
    if (auto ok = internal::ensure_plan(A, S, device); !ok) return std::unexpected(ok.error());
    int info = 0;
    if (spchol_factor_resident(S.plan.get(), A.x().data(), &info) != SPCHOL_OK) return std::unexpected(std::string(spchol_last_error()));
    if (info > 0) return std::unexpected(std::string("A is not positive definite."));
    return true;
}

// ---- beyond the reference: consumers of the device-resident factor of the last chol(A, S) ----
// x = A^-1 b: forward and backward substitution on the supernodal panels on the GPU.  b: n x nrhs, column-major.
template <typename T, typename SC = SChol>
std::expected<std::vector<T>, std::string> chol_solve(const SC& S, const std::vector<T>& b, int nrhs = 1) {
    static_assert(std::is_same_v<T, double>, "the B200 path is fp64");
    if (!S.plan) return std::unexpected(std::string("chol_solve: call chol(A, S) first"));
    if (b.size() != S.size() * static_cast<std::size_t>(nrhs)) return std::unexpected(std::string("chol_solve: b must hold n x nrhs values"));
    std::vector<T> x(b.size());
    if (spchol_solve(S.plan.get(), b.data(), x.data(), nrhs) != SPCHOL_OK) return std::unexpected(std::string(spchol_last_error()));
    return x;
}
// ||A - L L^T||_F / ||A||_F of that factor, estimated on the GPU (what the reference's tests do densely, tests/test_chol.cpp:87)
template <typename T, typename SC = SChol>
std::expected<double, std::string> chol_residual(const csc_matrix<T, sym::upper>& A, const SC& S) {
    static_assert(std::is_same_v<T, double>, "the B200 path is fp64");
    if (!S.plan) return std::unexpected(std::string("chol_residual: call chol(A, S) first"));
    double r = 0;
    if (spchol_residual(S.plan.get(), A.x().data(), &r) != SPCHOL_OK) return std::unexpected(std::string(spchol_last_error()));
    return r;
}

// ---- beyond the reference: fill-reducing ordering (the reference has none) ----
// perm[new] = old: approximate minimum degree on the pattern of A, composed with an etree postorder
template <typename T>
std::vector<int> order_amd(const csc_matrix<T, sym::upper>& A) {
    std::vector<int> perm(A.size());
    if (spchol_order_amd(static_cast<int>(A.size()), A.p().data(), A.i().data(), perm.data()) != SPCHOL_OK) internal::throw_last("order_amd");
    return perm;
}
// triu(P A P^T), perm[new] = old
template <typename T>
csc_matrix<T, sym::upper> permute(const csc_matrix<T, sym::upper>& A, const std::vector<int>& perm) {
    static_assert(std::is_same_v<T, double>, "the B200 path is fp64");
    const int n = static_cast<int>(A.size());
    if (perm.size() != A.size()) throw std::invalid_argument("permute: perm must hold n entries");
    csc_matrix<T, sym::upper> B(A.size(), A.capacity());
    if (spchol_permute_upper(n, A.p().data(), A.i().data(), A.x().data(), perm.data(), B.p().data(), B.i().data(), B.x().data(), nullptr) != SPCHOL_OK)
        internal::throw_last("permute");
    return B;
}

// reference chol.hpp:1449-1479
template <typename T, sym S>
std::vector<T> csc_to_dense(const csc_matrix<T, S>& A) {
    const int m = static_cast<int>(A.rows()), n = static_cast<int>(A.cols());
    std::vector<T> dense(static_cast<std::size_t>(m) * n, T(0));
    for (int j = 0; j < n; ++j)
        for (int k = A.p()[j]; k < A.p()[j + 1]; ++k) {
            const int i = A.i()[k];
            dense[i + static_cast<std::size_t>(j) * m] = A.x()[k];
            if constexpr (S != sym::none) { if (i != j) dense[j + static_cast<std::size_t>(i) * m] = A.x()[k]; }
        }
    return dense;
}
