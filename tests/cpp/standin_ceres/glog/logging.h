// Stand-in for glog (test only): CHECK aborts with a message.
#pragma once
#include <cstdio>
#include <cstdlib>
#define CHECK(cond) \
  do { if (!(cond)) { std::fprintf(stderr, "CHECK failed: %s\n", #cond); std::abort(); } } while (0)
