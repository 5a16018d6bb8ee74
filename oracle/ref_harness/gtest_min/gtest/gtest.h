// TEST INFRASTRUCTURE.  googletest is an unpinned submodule of the reference that is absent from this image; this is the
// small subset of its interface the reference's own unit tests (tests/test_reads.cpp, tests/test_spacer_ordering.cpp) use,
// so that those files compile UNMODIFIED (oracle/Makefile: _ref/ref_unit_tests) and can be run against the stand-ins for
// kseq++ / cft under mcaat_b200/host/absent_deps.  A failed expectation is printed and makes main() return 1.
#pragma once
#include <cstdio>
#include <functional>
#include <iostream>
#include <string>
#include <vector>
namespace testing_min {
struct Registry {
  static std::vector<std::pair<std::string, std::function<void()>>> &tests() {
    static std::vector<std::pair<std::string, std::function<void()>>> t;
    return t;
  }
  static int &failures() {
    static int f = 0;
    return f;
  }
};
struct Adder {
  Adder(const char *n, std::function<void()> f) { Registry::tests().push_back({n, std::move(f)}); }
};
struct Sink {  // EXPECT_x(...) << "message"
  template <class T> Sink &operator<<(const T &) { return *this; }
};
inline Sink check(bool ok, const char *expr, const char *file, int line) {
  if (!ok) {
    ++Registry::failures();
    std::printf("  FAILED %s:%d: %s\n", file, line, expr);
  }
  return Sink();
}
}  // namespace testing_min
#define TEST(suite, name)                                                                       \
  static void suite##_##name##_body();                                                         \
  static testing_min::Adder suite##_##name##_adder(#suite "." #name, suite##_##name##_body);   \
  static void suite##_##name##_body()
#define EXPECT_TRUE(c) testing_min::check(static_cast<bool>(c), #c, __FILE__, __LINE__)
#define EXPECT_FALSE(c) testing_min::check(!static_cast<bool>(c), "!(" #c ")", __FILE__, __LINE__)
#define EXPECT_EQ(a, b) testing_min::check((a) == (b), #a " == " #b, __FILE__, __LINE__)
#define EXPECT_NE(a, b) testing_min::check(!((a) == (b)), #a " != " #b, __FILE__, __LINE__)
#define EXPECT_LE(a, b) testing_min::check((a) <= (b), #a " <= " #b, __FILE__, __LINE__)
#define EXPECT_LT(a, b) testing_min::check((a) < (b), #a " < " #b, __FILE__, __LINE__)
#define EXPECT_GE(a, b) testing_min::check((a) >= (b), #a " >= " #b, __FILE__, __LINE__)
#define EXPECT_GT(a, b) testing_min::check((a) > (b), #a " > " #b, __FILE__, __LINE__)
#define ASSERT_TRUE EXPECT_TRUE
#define ASSERT_FALSE EXPECT_FALSE
#define ASSERT_EQ EXPECT_EQ
#define ASSERT_NE EXPECT_NE
#define ASSERT_LE EXPECT_LE
#define ASSERT_LT EXPECT_LT
#ifdef GTEST_MIN_MAIN
int main() {
  for (auto &t : testing_min::Registry::tests()) {
    int before = testing_min::Registry::failures();
    t.second();
    std::printf("[%s] %s\n", testing_min::Registry::failures() == before ? "  OK  " : " FAIL ", t.first.c_str());
  }
  std::printf("UNIT_TESTS %zu run, %d failed expectations\n", testing_min::Registry::tests().size(), testing_min::Registry::failures());
  return testing_min::Registry::failures() ? 1 : 0;
}
#endif
