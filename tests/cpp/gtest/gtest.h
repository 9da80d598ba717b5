// gtest/gtest.h -- a minimal stand-in for the parts of GoogleTest the reference's tests/tests.cpp uses
// (::testing::Test fixtures, TEST_F, EXPECT_TRUE/EXPECT_NEAR/... with a streamed message, RUN_ALL_TESTS).
// Test infrastructure only: GoogleTest itself is fetched over the network by the reference's build
// (tests/CMakeLists.txt:10-18), which is not possible here.  Same macro names and semantics for what is
// covered, nothing else.  A test program built with it prints one "[ OK ]" / "[ FAILED ]" line per test
// and returns the number of failed tests.
#ifndef RTP_TESTS_GTEST_SHIM_H
#define RTP_TESTS_GTEST_SHIM_H

#include <cmath>
#include <functional>
#include <iostream>
#include <sstream>
#include <string>
#include <vector>

namespace testing {

class Test {
  public:
    virtual ~Test() {}
    virtual void SetUp() {}
    virtual void TearDown() {}
    virtual void TestBody() = 0;
};

namespace internal {

struct Registry {
    struct Entry { std::string name; std::function<Test*()> make; };
    std::vector<Entry> entries;
    int failures_in_current = 0;
    static Registry& get() { static Registry r; return r; }
};

struct Registrar {
    Registrar(const char* suite, const char* name, std::function<Test*()> make) {
        Registry::get().entries.push_back({std::string(suite) + "." + name, std::move(make)});
    }
};

// collects the `<< message` that may follow an expectation and reports on destruction
class Failure {
  public:
    Failure(bool failed, const char* file, int line, const std::string& what) : failed_(failed) {
        if (failed_) { ++Registry::get().failures_in_current; os_ << file << ":" << line << ": Failure\n" << what << "\n"; }
    }
    ~Failure() { if (failed_) std::cerr << os_.str() << std::endl; }
    template <typename T> Failure& operator<<(const T& v) { if (failed_) os_ << v; return *this; }
  private:
    bool failed_;
    std::ostringstream os_;
};

template <typename A, typename B> std::string pair_text(const char* ea, const char* eb, const A& a, const B& b, const char* op) {
    std::ostringstream os;
    os.precision(17);
    os << "Expected: (" << ea << ") " << op << " (" << eb << "), actual: " << a << " vs " << b;
    return os.str();
}

}  // namespace internal

inline void InitGoogleTest(int*, char**) {}
inline void InitGoogleTest() {}

}  // namespace testing

#define RTP_GTEST_CLASS_(suite, name) suite##_##name##_Test

#define RTP_GTEST_DEFINE_(suite, name, parent)                                                               \
    class RTP_GTEST_CLASS_(suite, name) : public parent {                                                    \
      public:                                                                                                \
        void TestBody() override;                                                                            \
    };                                                                                                       \
    static ::testing::internal::Registrar rtp_gtest_reg_##suite##_##name(                                    \
        #suite, #name, []() -> ::testing::Test* { return new RTP_GTEST_CLASS_(suite, name); });              \
    void RTP_GTEST_CLASS_(suite, name)::TestBody()

#define TEST_F(fixture, name) RTP_GTEST_DEFINE_(fixture, name, fixture)
#define TEST(suite, name) RTP_GTEST_DEFINE_(suite, name, ::testing::Test)

#define EXPECT_TRUE(c) ::testing::internal::Failure(!(c), __FILE__, __LINE__, "Value of: " #c "\n  Actual: false\nExpected: true")
#define EXPECT_FALSE(c) ::testing::internal::Failure(!!(c), __FILE__, __LINE__, "Value of: " #c "\n  Actual: true\nExpected: false")
#define RTP_GTEST_CMP_(a, b, op) ::testing::internal::Failure(!((a) op (b)), __FILE__, __LINE__, ::testing::internal::pair_text(#a, #b, (a), (b), #op))
#define EXPECT_EQ(a, b) RTP_GTEST_CMP_(a, b, ==)
#define EXPECT_NE(a, b) RTP_GTEST_CMP_(a, b, !=)
#define EXPECT_LT(a, b) RTP_GTEST_CMP_(a, b, <)
#define EXPECT_LE(a, b) RTP_GTEST_CMP_(a, b, <=)
#define EXPECT_GT(a, b) RTP_GTEST_CMP_(a, b, >)
#define EXPECT_GE(a, b) RTP_GTEST_CMP_(a, b, >=)
#define EXPECT_NEAR(a, b, tol) ::testing::internal::Failure(!(std::fabs(double(a) - double(b)) <= double(tol)), __FILE__, __LINE__, ::testing::internal::pair_text(#a, #b, (a), (b), "~="))
#define ASSERT_TRUE(c) EXPECT_TRUE(c)
#define ASSERT_EQ(a, b) EXPECT_EQ(a, b)

inline int RUN_ALL_TESTS() {
    auto& reg = ::testing::internal::Registry::get();
    int failed = 0;
    for (auto& e : reg.entries) {
        std::cerr << "[ RUN      ] " << e.name << std::endl;
        reg.failures_in_current = 0;
        try {
            ::testing::Test* t = e.make();
            t->SetUp();
            t->TestBody();
            t->TearDown();
            delete t;
        } catch (const std::exception& ex) {
            ++reg.failures_in_current;
            std::cerr << "unexpected exception: " << ex.what() << std::endl;
        }
        if (reg.failures_in_current) { ++failed; std::cerr << "[  FAILED  ] " << e.name << std::endl; }
        else std::cerr << "[       OK ] " << e.name << std::endl;
    }
    std::cerr << "[==========] " << reg.entries.size() << " tests, " << failed << " failed" << std::endl;
    return failed;
}

#endif  // RTP_TESTS_GTEST_SHIM_H
