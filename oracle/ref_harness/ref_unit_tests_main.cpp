// TEST INFRASTRUCTURE: main() for the reference's own unit tests (see gtest_min/gtest/gtest.h).
#define GTEST_MIN_MAIN
#include <gtest/gtest.h>
