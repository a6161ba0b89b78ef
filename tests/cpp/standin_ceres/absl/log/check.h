// Stand-in for abseil's absl/log/check.h (test only): CHECK aborts with a message.
#pragma once
#include <cstdio>
#include <cstdlib>
#define CHECK(cond) \
  do { if (!(cond)) { std::fprintf(stderr, "CHECK failed: %s\n", #cond); std::abort(); } } while (0)
