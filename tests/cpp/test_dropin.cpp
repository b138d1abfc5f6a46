// The reference's tests/test_chol.cpp restated as plain asserts (GoogleTest is not available
// offline) against the drop-in header.  Usage: test_dropin <data dir> [gpu]
// Without "gpu" only the host-side cases run (the CPU container); with it chol() runs on the GPU.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "chol.hpp"
#include "mtx_reader.hpp"

#define CHECK(c) do { if (!(c)) { std::printf("FAILED %s:%d: %s\n", __FILE__, __LINE__, #c); std::exit(1); } } while (0)

static void dense_cholesky(std::vector<double>& a, int n) {      // stand-in for the test's dpotrf_ call
    for (int k = 0; k < n; ++k) {
        a[k + k * n] = std::sqrt(a[k + k * n]);
        for (int i = k + 1; i < n; ++i) a[i + k * n] /= a[k + k * n];
        for (int j = k + 1; j < n; ++j) for (int i = j; i < n; ++i) a[i + j * n] -= a[i + k * n] * a[j + k * n];
    }
}

extern "C" void dpotrf_(const char*, const int* n, double* a, const int* lda, int* info) {   // the header only declares LAPACK's
    std::vector<double> t(a, a + static_cast<std::size_t>(*lda) * *n);
    dense_cholesky(t, *n);
    std::copy(t.begin(), t.end(), a);
    *info = 0;
}

template <class V>
static void dump(std::FILE* f, const char* name, const V& v) {
    std::fprintf(f, "%s %zu", name, v.size());
    for (const auto& e : v) std::fprintf(f, " %.17g", static_cast<double>(e));
    std::fprintf(f, "\n");
}

int main(int argc, char** argv) {
    const std::string data = argc > 1 ? argv[1] : "tests/data";
    const bool gpu = argc > 2 && !std::strcmp(argv[2], "gpu");
    if (argc > 3 && !std::strcmp(argv[2], "dump")) {
        // the remaining public helpers of the reference header, written out for a bit-for-bit comparison with what the
        // reference itself produced (tests/golden/boundary.npz, generated from oracle/_ref/libref.so)
        std::FILE* f = std::fopen(argv[3], "w");
        CHECK(f != nullptr);
        for (auto [n, dens] : {std::pair<int, double>{10, 0.25}, {37, 0.1}}) {
            auto R = random_sparse<double>(n, dens);                       // reference chol.hpp:964
            const std::string tag = "rs" + std::to_string(n);
            dump(f, (tag + "_p").c_str(), R.p()); dump(f, (tag + "_i").c_str(), R.i()); dump(f, (tag + "_x").c_str(), R.x());
        }
        auto A = load_matrix_market_to_csc<double>(data + "/bcsstk01.mtx");
        auto S = schol(A);
        auto B = random_sparse<double>(S);                                 // :1004 (lower, on the pattern of S)
        dump(f, "rsS_x", B.x());
        auto Bt = B.transpose();                                           // usage of reference main.cpp:323
        CHECK(Bt.p()[Bt.size()] == B.p()[B.size()]);
        csc_matrix<double, sym::upper> Wk(A, S);                           // :190 (only the diagonal carries over from an upper A)
        dump(f, "ctor_x", Wk.x());
        auto E = schol_etree(A);                                           // :628
        dump(f, "se_p", E.p()); dump(f, "se_i", E.i());
        std::fclose(f);
        std::printf("test_dropin dump OK\n");
        return 0;
    }
    const std::vector<std::vector<int>> pattern = {{0}, {1}, {0, 2}, {3}, {0, 2, 4}, {0, 1, 3, 5}, {0, 2, 5, 6}};
    {   // CholeskyTest.EliminationTree (reference tests/test_chol.cpp:6-25)
        auto A = build_csc_matrix_from_pattern<double>(pattern);
        auto parent = etree(A);
        CHECK(parent.size() == A.size());
        CHECK((parent == std::vector<int>{2, 5, 4, 5, 5, 6, -1}));
    }
    {   // CholeskyTest.ColumnReach (:27-57)
        auto A = build_csc_matrix_from_pattern<double>(pattern);
        const auto n = A.size();
        std::vector<int> w(n, -1), s(n);
        std::vector<double> x(n);
        auto parent = etree(A);
        auto t = ereach(A, 5, parent, s, w, x, n); (void)t;
        CHECK((s == std::vector<int>{3, 1, 0, 2, 4, 5, 6}));
        t = ereach(A, 5, parent, s, w, n);
        CHECK((s == std::vector<int>{3, 1, 0, 2, 4, 5, 6}));
    }
    {   // symbolic known answers: bcsstk01 through the reader
        auto A = load_matrix_market_to_csc<double>(data + "/bcsstk01.mtx");
        CHECK(A.size() == 48 && A.p()[48] == 224);
        auto S = schol(A);
        CHECK(S.capacity() == 877 && S.size() == 48);
        std::vector<std::size_t> sn;
        auto id = compute_supernodes(S, sn);
        CHECK(sn.size() == 16 && sn.back() == 48);
        auto at = atree(S, id, sn);
        CHECK(compute_levels(at).size() == 13 && compute_levels(S.parent).size() == 46);
        // ordering (beyond the reference): AMD cuts nnz(L) 877 -> about 493 and permute keeps the matrix symmetric-equal
        auto perm = order_amd(A);
        auto B = permute(A, perm);
        CHECK(B.p()[48] == 224 && schol(B).capacity() <= 520);
        const auto& Ac = A; const auto& Bc = B;          // const access: absent entries read as 0
        for (int a = 0; a < 48; ++a) for (int b = a; b < 48; ++b) CHECK((Bc[a, b] == Ac[perm[a], perm[b]]));
        bool threw = false;
        try { load_matrix_market_to_csc<double>("/nonexistent.mtx"); } catch (const std::runtime_error&) { threw = true; }
        CHECK(threw);
    }
    if (gpu) {
        {   // CholeskyTest.SimplicialCholesky (:59-97)
            std::vector<double> expected = {4.0, 1.0, 1.0, 1.0, 3.0, 0.0, 1.0, 0.0, 2.0};
            dense_cholesky(expected, 3);
            std::vector<int> ti = {0, 0, 0, 1, 1, 2}, tj = {0, 1, 2, 1, 2, 2};
            std::vector<double> tx = {4.0, 1.0, 1.0, 3.0, 0.0, 2.0};
            auto A = triplet_to_csc_matrix(ti, tj, tx, 3);
            auto L = chol(A);
            CHECK(L.has_value());
            auto result = csc_to_dense(*L);
            for (int j = 0; j < 3; ++j) for (int i = j; i < 3; ++i) CHECK(std::fabs(result[i + j * 3] - expected[i + j * 3]) < 1e-9);
        }
        {   // README usage: S reused across numeric calls; known first values of L
            auto A = load_matrix_market_to_csc<double>(data + "/bcsstk01.mtx");
            auto S = schol(A);
            auto L = chol(A, S).value();
            CHECK(L.p() == S.p() && L.i() == S.i());
            CHECK(std::fabs(L.x()[0] - 1682.9344962059574) < 1e-9 && std::fabs(L.x()[3] - (-1.9806673051415467)) < 1e-9);
            auto L2 = chol(A, S).value();
            CHECK(L2.x() == L.x());
            // consumers of the resident factor: residual estimate and solve (A x = A e, e = ones)
            auto res = chol_residual(A, S);
            CHECK(res.has_value() && *res <= 1e-12);
            const int n = static_cast<int>(A.size());
            std::vector<double> b(n, 0.0);
            for (int j = 0; j < n; ++j)
                for (int k = A.p()[j]; k < A.p()[j + 1]; ++k) { const int i = A.i()[k]; b[i] += A.x()[k]; if (i != j) b[j] += A.x()[k]; }
            auto x = chol_solve(S, b);
            CHECK(x.has_value());
            for (int i = 0; i < n; ++i) CHECK(std::fabs((*x)[i] - 1.0) < 1e-8);
        }
        {   // CholeskyTest.SupernodalCholesky (:99-136): chol_sn is the same supernodal GPU factorization
            std::vector<int> ti = {0, 0, 0, 1, 1, 2}, tj = {0, 1, 2, 1, 2, 2};
            std::vector<double> tx = {4.0, 1.0, 1.0, 3.0, 0.0, 2.0}, expected = {4.0, 1.0, 1.0, 1.0, 3.0, 0.0, 1.0, 0.0, 2.0};
            dense_cholesky(expected, 3);
            auto A = triplet_to_csc_matrix(ti, tj, tx, 3);
            auto L = chol_sn(A);
            CHECK(L.has_value());
            auto result = csc_to_dense(*L);
            for (int j = 0; j < 3; ++j) for (int i = j; i < 3; ++i) CHECK(std::fabs(result[i + j * 3] - expected[i + j * 3]) < 1e-9);
        }
        {   // 64-bit container: same factor as the 32-bit path, entry for entry; resident factor + solve
            auto A = load_matrix_market_to_csc<double>(data + "/1138_bus.mtx");
            auto S = schol(A);
            auto L = chol(A, S).value();
            auto S64 = schol64(A);
            CHECK(S64.nnz == static_cast<std::int64_t>(S.capacity()) && S64.parent == S.parent);
            auto L64 = chol64(A, S64);
            CHECK(L64.has_value());
            CHECK(L64->p().size() == L.p().size() && L64->i() == L.i());
            for (std::size_t j = 0; j < L.p().size(); ++j) CHECK(L64->p()[j] == L.p()[j]);
            CHECK(L64->x() == L.x());
            CHECK(((*L64)[5, 5] == L.x()[L.p()[5]]));
            auto ok = chol_resident(A, S64);
            CHECK(ok.has_value());
            auto res = chol_residual(A, S64);
            CHECK(res.has_value() && *res <= 1e-12);
            // a matrix with another pattern is refused instead of being read out of bounds
            auto A2 = load_matrix_market_to_csc<double>(data + "/bcsstk01.mtx");
            CHECK(!chol(A2, S).has_value());
        }
        {   // error path (reference chol.hpp:850)
            auto A = triplet_to_csc_matrix<double>({0, 0, 1}, {0, 1, 1}, {1.0, 2.0, 1.0}, 2);
            auto L = chol(A);
            CHECK(!L.has_value() && L.error() == "A is not positive definite.");
        }
        for (const char* name : {"/bcsstk01.mtx", "/1138_bus.mtx"}) {   // the reader with parse and COO -> CSC on the GPU: same matrix, bit for bit
            auto A = load_matrix_market_to_csc<double>(data + name);
            auto B = load_matrix_market_to_csc_gpu<double>(data + name);
            CHECK(B.size() == A.size() && B.p() == A.p() && B.i() == A.i() && B.x() == A.x());
        }
        {
            bool threw = false;
            try { load_matrix_market_to_csc_gpu<double>("/nonexistent.mtx"); } catch (const std::runtime_error&) { threw = true; }
            CHECK(threw);
        }
    } else {
        auto A = triplet_to_csc_matrix<double>({0, 0, 1}, {0, 1, 1}, {4.0, 1.0, 3.0}, 2);
        auto L = chol(A);
        if (spchol_device_count() == 0) CHECK(!L.has_value() && L.error().find("no CUDA device") != std::string::npos);
        auto S0 = schol(A);
        CHECK(!chol_solve(S0, std::vector<double>{1.0, 2.0}).has_value());       // no factor yet: refused, not guessed
        CHECK(!chol_residual(A, S0).has_value());
    }
    std::printf("test_dropin OK (%s)\n", gpu ? "gpu" : "host only");
    return 0;
}
