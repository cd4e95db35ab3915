// gtest_main for tests/cpp/mini_gtest/gtest/gtest.h
#include <cstdlib>
#include <gtest/gtest.h>

int main(int argc, char** argv)
{
    ::testing::InitGoogleTest(&argc, argv);
    return RUN_ALL_TESTS();
}
