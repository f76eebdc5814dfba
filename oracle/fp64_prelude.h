// fp64_prelude.h — TEST INFRASTRUCTURE ONLY.
// Force-included (-include) when oracle/Makefile builds the FP64 variant of the
// UNMODIFIED reference: every standard header the reference uses is pulled in
// first (so libstdc++ itself is compiled with the real `float`), then the
// keyword is redefined so that every float table / temporary of the reference
// (e.g. /root/reference/src/base_pattern.h:117-120, src/peng.cpp:63-82) becomes
// a double.  This is "Oracle B" of SURVEY.md §8c: the FP64 arm that the
// 1e-9-relative checks compare against.
#pragma once
#include <omp.h>
#include <algorithm>
#include <array>
#include <assert.h>
#include <cassert>
#include <chrono>
#include <climits>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctype.h>
#include <dirent.h>
#include <fstream>
#include <iomanip>
#include <iostream>
#include <limits>
#include <map>
#include <memory>
#include <numeric>
#include <set>
#include <sstream>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <string>
#include <sys/stat.h>
#include <sys/time.h>
#include <tuple>
#include <type_traits>
#include <vector>
#define float double
