// A few lines of test runner with googletest's spelling (TEST_F, ASSERT_*), so that the reference's own test files can
// be compiled and run against this library where googletest is not installed.  Test infrastructure only.
#pragma once
#include <functional>
#include <iostream>
#include <string>
#include <vector>

namespace testing
{
    struct Test
    {
        virtual ~Test() = default;
        virtual void SetUp() {}
        virtual void TearDown() {}
        virtual void TestBody() {}
    };
    struct Registry
    {
        std::vector<std::pair<std::string, std::function<Test*()>>> tests;
        int failures = 0;
        static Registry& get() { static Registry r; return r; }
    };
    struct Registrar { Registrar(const char* name, std::function<Test*()> make) { Registry::get().tests.emplace_back(name, std::move(make)); } };
    struct Message
    {
        bool failed;
        Message(bool f, const char* file, int line, const char* what) : failed(f) { if (f) { ++Registry::get().failures; std::cerr << file << ":" << line << ": Failure: " << what << " "; } }
        ~Message() { if (failed) std::cerr << std::endl; }
        template <class T> Message& operator<<(const T& v) { if (failed) std::cerr << v; return *this; }
    };
    inline void InitGoogleTest(int*, char**) {}
    inline int RunAll(const std::string& filter)
    {
        int ran = 0;
        for (auto& t : Registry::get().tests)
        {
            if (!filter.empty() && t.first.find(filter) == std::string::npos) continue;
            const int before = Registry::get().failures;
            std::cout << "[ RUN      ] " << t.first << std::endl;
            Test* test = t.second();
            test->SetUp();
            test->TestBody();
            test->TearDown();
            delete test;
            std::cout << (Registry::get().failures == before ? "[       OK ] " : "[  FAILED  ] ") << t.first << std::endl;
            ++ran;
        }
        std::cout << "[  TOTAL   ] " << ran << " tests, " << Registry::get().failures << " failures" << std::endl;
        return Registry::get().failures ? 1 : 0;
    }
}

#define TEST_F(fixture, name)                                                                                         \
    struct fixture##_##name##_Test : fixture { void TestBody() override; };                                           \
    static ::testing::Registrar fixture##_##name##_registrar(#fixture "." #name, [] { return static_cast<::testing::Test*>(new fixture##_##name##_Test); }); \
    void fixture##_##name##_Test::TestBody()
#define MINI_CHECK(cond, text) ::testing::Message(!(cond), __FILE__, __LINE__, text)
#define ASSERT_EQ(a, b) MINI_CHECK((a) == (b), #a " == " #b)
#define ASSERT_NE(a, b) MINI_CHECK((a) != (b), #a " != " #b)
#define ASSERT_LE(a, b) MINI_CHECK((a) <= (b), #a " <= " #b)
#define ASSERT_GE(a, b) MINI_CHECK((a) >= (b), #a " >= " #b)
#define ASSERT_LT(a, b) MINI_CHECK((a) < (b), #a " < " #b)
#define ASSERT_GT(a, b) MINI_CHECK((a) > (b), #a " > " #b)
#define ASSERT_TRUE(a) MINI_CHECK((a), #a)
#define ASSERT_FALSE(a) MINI_CHECK(!(a), "!" #a)
#define EXPECT_EQ ASSERT_EQ
#define EXPECT_TRUE ASSERT_TRUE
#define RUN_ALL_TESTS() ::testing::RunAll(std::getenv("MINI_GTEST_FILTER") ? std::getenv("MINI_GTEST_FILTER") : "")
