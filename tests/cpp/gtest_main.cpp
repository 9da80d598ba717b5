// gtest_main.cpp -- what linking `gtest_main` gives the reference's test target (tests/CMakeLists.txt:27).
#include <gtest/gtest.h>

int main(int argc, char** argv) {
    ::testing::InitGoogleTest(&argc, argv);
    return RUN_ALL_TESTS();
}
