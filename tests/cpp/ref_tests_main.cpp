// TEST INFRASTRUCTURE ONLY.  main() for the reference's own tests/test_chol.cpp compiled verbatim against the drop-in
// header (see gtest_shim/gtest/gtest.h), plus the dense stand-in for LAPACK's dpotrf_ that those tests call as their
// cross-check (reference tests/test_chol.cpp:73,113; the reference links lapack-reference through vcpkg).
#include <cmath>
#include <cstdio>
#include <cstring>

#include "gtest/gtest.h"

extern "C" void dpotrf_(const char* uplo, const int* n, double* a, const int* lda, int* info) {
    const int N = *n, ld = *lda;
    *info = 0;
    if (*uplo != 'L' && *uplo != 'l') { *info = -1; return; }
    for (int k = 0; k < N; ++k) {
        const double d = a[k + k * ld];
        if (!(d > 0.0)) { *info = k + 1; return; }
        const double p = std::sqrt(d);
        a[k + k * ld] = p;
        for (int i = k + 1; i < N; ++i) a[i + k * ld] /= p;
        for (int j = k + 1; j < N; ++j) for (int i = j; i < N; ++i) a[i + j * ld] -= a[i + k * ld] * a[j + k * ld];
    }
}

int main(int argc, char** argv) {
    // "host" : only the cases that need no GPU (the CPU container); default: all four
    const bool host_only = argc > 1 && !std::strcmp(argv[1], "host");
    int ran = 0;
    for (const auto& c : gtest_shim::registry()) {
        const bool needs_gpu = std::strstr(c.name, "Cholesky") != nullptr;       // SimplicialCholesky, SupernodalCholesky call chol()
        if (host_only && needs_gpu) continue;
        std::printf("[ RUN ] %s.%s\n", c.suite, c.name);
        const int before = gtest_shim::failures();
        c.fn();
        std::printf("[ %s ] %s.%s\n", gtest_shim::failures() == before ? " OK " : "FAIL", c.suite, c.name);
        ++ran;
    }
    std::printf("ref_tests: %d ran, %d failed checks\n", ran, gtest_shim::failures());
    return gtest_shim::failures() ? 1 : 0;
}
