// TEST INFRASTRUCTURE ONLY.  A minimal stand-in for <gtest/gtest.h> (GoogleTest is not available offline) with just the
// macros the reference's tests/test_chol.cpp uses, so that file compiles VERBATIM against the drop-in include/chol.hpp:
// TEST, ASSERT_EQ / EXPECT_EQ, ASSERT_TRUE, EXPECT_NEAR (with a streamed message).  Failures are counted and printed;
// ASSERT_* additionally leaves the test body.  tests/cpp/ref_tests_main.cpp holds main() and the dense dpotrf_ stand-in.
#pragma once
#include <cmath>
#include <cstdio>
#include <functional>
#include <sstream>
#include <string>
#include <vector>

namespace gtest_shim {
struct Case { const char* suite; const char* name; std::function<void()> fn; };
inline std::vector<Case>& registry() { static std::vector<Case> r; return r; }
inline int& failures() { static int f = 0; return f; }
struct Registrar { Registrar(const char* s, const char* n, std::function<void()> f) { registry().push_back({s, n, std::move(f)}); } };
// collects the streamed message of a failed check and reports in its destructor
struct Report {
    bool failed; std::ostringstream msg; const char* file; int line; const char* what;
    Report(bool f, const char* fi, int li, const char* w) : failed(f), file(fi), line(li), what(w) {}
    template <class T> Report& operator<<(const T& v) { if (failed) msg << v; return *this; }
    ~Report() { if (failed) { ++failures(); std::printf("  FAILED %s:%d: %s %s\n", file, line, what, msg.str().c_str()); } }
};
struct Voidify { void operator&(const Report&) const {} };
}  // namespace gtest_shim

#define TEST(suite, name)                                                                          \
    static void suite##_##name##_body();                                                           \
    static ::gtest_shim::Registrar suite##_##name##_reg(#suite, #name, suite##_##name##_body);     \
    static void suite##_##name##_body()

#define GTEST_SHIM_CHECK_(cond, what) ::gtest_shim::Report(!(cond), __FILE__, __LINE__, what)
#define EXPECT_TRUE(c) GTEST_SHIM_CHECK_((c), "EXPECT_TRUE(" #c ")")
#define EXPECT_EQ(a, b) GTEST_SHIM_CHECK_((a) == (b), "EXPECT_EQ(" #a ", " #b ")")
#define EXPECT_NEAR(a, b, tol) GTEST_SHIM_CHECK_(std::fabs((a) - (b)) <= (tol), "EXPECT_NEAR(" #a ", " #b ", " #tol ")")
// ASSERT_*: report and leave the test body (the streamed message is dropped -- the reference's asserts stream nothing)
#define ASSERT_TRUE(c) if (!(c)) { ::gtest_shim::Report(true, __FILE__, __LINE__, "ASSERT_TRUE(" #c ")"); return; } else (void)0
#define ASSERT_EQ(a, b) if (!((a) == (b))) { ::gtest_shim::Report(true, __FILE__, __LINE__, "ASSERT_EQ(" #a ", " #b ")"); return; } else (void)0
