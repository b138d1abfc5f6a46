// oracle/ref_shim.cpp -- TEST INFRASTRUCTURE ONLY.
// C-ABI wrapper around the UNMODIFIED reference (include/chol.hpp, src/chol.cpp,
// include/mtx_reader.hpp) so tests and the bench's reference arm can call the reference's own
// functions through ctypes.  Built by oracle/Makefile into oracle/_ref/libref.so from the
// sources where they lie under /root/reference (nothing of the reference is copied into the
// repository; oracle/_ref/ is git-ignored).  The only accommodation is the one SURVEY.md §8(c)
// documents: empty Eigen/cblas shim headers and eliding the never-called factorize_sn_eigen.
#include <chrono>
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

#include "chol.hpp"
#include "mtx_reader.hpp"

namespace {
csc_matrix<double, sym::upper> make_A(int n, const int* Ap, const int* Ai, const double* Ax) {
    const int nz = Ap[n];
    csc_matrix<double, sym::upper> A(n, nz > 0 ? nz : 1);
    A.p().assign(Ap, Ap + n + 1);
    A.i().assign(Ai, Ai + nz);
    if (Ax) A.x().assign(Ax, Ax + nz); else A.x().assign(nz, 1.0);
    return A;
}
SChol make_S(int n, const int* parent, const int* Lp, const int* Li) {
    SChol S(n, Lp[n]);
    S.p().assign(Lp, Lp + n + 1);
    if (Li) S.i().assign(Li, Li + Lp[n]);
    S.parent.assign(parent, parent + n);
    return S;
}
thread_local std::string g_err;
}  // namespace

extern "C" {

const char* ref_last_error() { return g_err.c_str(); }

// triplet_to_csc_matrix (include/chol.hpp:309). Returns nnz; outputs sized by caller (>= nz).
int ref_triplet_to_csc(int nz, const int* ti, const int* tj, const double* tx, int n,
                       int* Ap, int* Ai, double* Ax) {
    std::vector<int> vi(ti, ti + nz), vj(tj, tj + nz);
    std::vector<double> vx(tx, tx + nz);
    auto A = triplet_to_csc_matrix<double>(vi, vj, vx, n);
    std::memcpy(Ap, A.p().data(), sizeof(int) * (n + 1));
    const int m = A.p()[n];
    std::memcpy(Ai, A.i().data(), sizeof(int) * m);
    std::memcpy(Ax, A.x().data(), sizeof(double) * m);
    return m;
}

// load_matrix_market_to_csc (include/mtx_reader.hpp:17). Two-call protocol: Ap==NULL -> sizes.
int ref_load_mtx(const char* path, int* n_out, int* nz_out, int* Ap, int* Ai, double* Ax) {
    try {
        auto A = load_matrix_market_to_csc<double>(path);
        const int n = static_cast<int>(A.size());
        *n_out = n; *nz_out = A.p()[n];
        if (Ap) {
            std::memcpy(Ap, A.p().data(), sizeof(int) * (n + 1));
            std::memcpy(Ai, A.i().data(), sizeof(int) * A.p()[n]);
            std::memcpy(Ax, A.x().data(), sizeof(double) * A.p()[n]);
        }
        return 0;
    } catch (const std::exception& e) { g_err = e.what(); return -1; }
}

void ref_etree(int n, const int* Ap, const int* Ai, int* parent) {
    auto p = etree(make_A(n, Ap, Ai, nullptr));
    std::memcpy(parent, p.data(), sizeof(int) * n);
}
void ref_post_order(int n, const int* parent, int* post) {
    auto p = post_order(std::vector<int>(parent, parent + n));
    std::memcpy(post, p.data(), sizeof(int) * n);
}
void ref_col_count(int n, const int* Ap, const int* Ai, const int* parent, const int* post, int* cc) {
    auto c = col_count(make_A(n, Ap, Ai, nullptr), std::vector<int>(parent, parent + n),
                       std::vector<int>(post, post + n));
    std::memcpy(cc, c.data(), sizeof(int) * n);
}
// ereach (include/chol.hpp:725): s, w, x are in/out length-n arrays exactly as in the reference.
int ref_ereach(int n, const int* Ap, const int* Ai, const double* Ax, int k, const int* parent,
               int* s, int* w, double* x, int top) {
    auto A = make_A(n, Ap, Ai, Ax);
    std::vector<int> vs(s, s + n), vw(w, w + n), vp(parent, parent + n);
    std::size_t t;
    if (x) {
        std::vector<double> vx(x, x + n);
        t = ereach(A, static_cast<std::size_t>(k), vp, vs, vw, vx, static_cast<std::size_t>(top));
        std::memcpy(x, vx.data(), sizeof(double) * n);
    } else {
        t = ereach(A, static_cast<std::size_t>(k), vp, vs, vw, static_cast<std::size_t>(top));
    }
    std::memcpy(s, vs.data(), sizeof(int) * n);
    std::memcpy(w, vw.data(), sizeof(int) * n);
    return static_cast<int>(t);
}
// schol (include/chol.hpp:874). Li==NULL: only parent/Lp. seconds_out (optional) = wall time.
int ref_schol(int n, const int* Ap, const int* Ai, int* parent, int* Lp, int* Li, double* seconds_out) {
    auto A = make_A(n, Ap, Ai, nullptr);
    auto t0 = std::chrono::steady_clock::now();
    SChol S = schol(A);
    auto t1 = std::chrono::steady_clock::now();
    if (seconds_out) *seconds_out = std::chrono::duration<double>(t1 - t0).count();
    std::memcpy(parent, S.parent.data(), sizeof(int) * n);
    std::memcpy(Lp, S.p().data(), sizeof(int) * (n + 1));
    if (Li) std::memcpy(Li, S.i().data(), sizeof(int) * S.p()[n]);
    return S.p()[n];
}
// chol (include/chol.hpp:750). Returns 0 ok, 1 = unexpected(...) (message via ref_last_error).
// Lp/Li/Lx sized from ref_schol. seconds_out = wall time of chol() only.
int ref_chol(int n, const int* Ap, const int* Ai, const double* Ax, int* Lp, int* Li, double* Lx,
             double* seconds_out) {
    auto A = make_A(n, Ap, Ai, Ax);
    auto t0 = std::chrono::steady_clock::now();
    auto L = chol(A);
    auto t1 = std::chrono::steady_clock::now();
    if (seconds_out) *seconds_out = std::chrono::duration<double>(t1 - t0).count();
    if (!L.has_value()) { g_err = L.error(); return 1; }
    const int nz = L->p()[n];
    if (Lp) std::memcpy(Lp, L->p().data(), sizeof(int) * (n + 1));
    if (Li) std::memcpy(Li, L->i().data(), sizeof(int) * nz);
    if (Lx) std::memcpy(Lx, L->x().data(), sizeof(double) * nz);
    return 0;
}
// compute_levels (src/chol.cpp:7). level_ptr needs n+1 slots. Returns nlev.
int ref_compute_levels(int n, const int* parent, int* level_ptr, int* level_nodes) {
    auto lv = compute_levels(std::vector<int>(parent, parent + n));
    int k = 0; level_ptr[0] = 0;
    for (std::size_t l = 0; l < lv.size(); ++l) {
        for (int v : lv[l]) level_nodes[k++] = v;
        level_ptr[l + 1] = k;
    }
    return static_cast<int>(lv.size());
}
// compute_supernodes (src/chol.cpp:42). sn_start needs n+1 slots. Returns ns.
int ref_compute_supernodes(int n, const int* parent, const int* Lp, int* sn_id, int* sn_start) {
    SChol S = make_S(n, parent, Lp, nullptr);
    std::vector<std::size_t> sn;
    auto id = compute_supernodes(S, sn);
    std::memcpy(sn_id, id.data(), sizeof(int) * n);
    for (std::size_t k = 0; k < sn.size(); ++k) sn_start[k] = static_cast<int>(sn[k]);
    return static_cast<int>(sn.size()) - 1;
}
// atree (src/chol.cpp:102).
void ref_atree(int n, const int* parent, const int* Lp, const int* Li, const int* sn_id, int ns,
               const int* sn_start, int* super_parent) {
    SChol S = make_S(n, parent, Lp, Li);
    std::vector<std::size_t> sn(sn_start, sn_start + ns + 1);
    auto sp = atree(S, std::vector<int>(sn_id, sn_id + n), sn);
    std::memcpy(super_parent, sp.data(), sizeof(int) * ns);
}
// random_sparse<T>(n, density, positive_definite) (include/chol.hpp:964).  Two-call protocol: Ap == NULL -> returns nnz only.
int ref_random_sparse(int n, double density, int pd, int* Ap, int* Ai, double* Ax) {
    auto A = random_sparse<double>(n, density, pd != 0);
    const int nz = A.p()[n];
    if (Ap) { std::memcpy(Ap, A.p().data(), sizeof(int) * (n + 1)); std::memcpy(Ai, A.i().data(), sizeof(int) * nz); std::memcpy(Ax, A.x().data(), sizeof(double) * nz); }
    return nz;
}
// random_sparse<T>(S, positive_definite) (include/chol.hpp:1004): values on the pattern of S (Lp[n] doubles)
void ref_random_sparse_S(int n, const int* parent, const int* Lp, const int* Li, int pd, double* x) {
    SChol S = make_S(n, parent, Lp, Li);
    auto B = random_sparse<double>(S, pd != 0);
    std::memcpy(x, B.x().data(), sizeof(double) * Lp[n]);
}
// csc_matrix(A, S) constructor (include/chol.hpp:190): values on the pattern of S
void ref_csc_from_A_S(int n, const int* Ap, const int* Ai, const double* Ax, const int* parent, const int* Lp, const int* Li, double* x) {
    auto A = make_A(n, Ap, Ai, Ax);
    SChol S = make_S(n, parent, Lp, Li);
    csc_matrix<double, sym::upper> W(A, S);
    std::memcpy(x, W.x().data(), sizeof(double) * Lp[n]);
}
// schol_etree (include/chol.hpp:628).  Two-call protocol: Li == NULL -> Lp only; returns nnz.
int ref_schol_etree(int n, const int* Ap, const int* Ai, int* Lp, int* Li) {
    auto A = make_A(n, Ap, Ai, nullptr);
    SChol S = schol_etree(A);
    std::memcpy(Lp, S.p().data(), sizeof(int) * (n + 1));
    if (Li) std::memcpy(Li, S.i().data(), sizeof(int) * S.p()[n]);
    return S.p()[n];
}
// csc_to_dense for sym::none (include/chol.hpp:1449).
void ref_csc_to_dense_none(int n, const int* Lp, const int* Li, const double* Lx, double* dense) {
    csc_matrix<double, sym::none> L(n, n, Lp[n]);
    L.p().assign(Lp, Lp + n + 1); L.i().assign(Li, Li + Lp[n]); L.x().assign(Lx, Lx + Lp[n]);
    auto d = csc_to_dense(L);
    std::memcpy(dense, d.data(), sizeof(double) * n * n);
}
}  // extern "C"
