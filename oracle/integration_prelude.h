/* TEST INFRASTRUCTURE ONLY.  Force-included (-include) when tests/test_integration_patch.py compiles the patched reference
 * driver: CountModel keeps its tally arrays private; a maintainer applying INTEGRATION.md's patch B would add an accessor
 * (or a friend), the compile check opens them the way oracle/ref_harness.cpp does.  Standard headers first, so that the
 * macro only touches the reference's own headers. */
#include <vector>
#include <string>
#include <queue>
#include <deque>
#include <map>
#include <set>
#include <valarray>
#include <sstream>
#include <iostream>
#include <fstream>
#include <memory>
#include <random>
#include <cmath>
#include <cstring>
#include <cstdint>
#include <stdexcept>
#include <algorithm>
#include <unordered_map>
#include <unordered_set>
#include <iomanip>
#include <climits>
#include <cfloat>
#include <cassert>
#include <functional>
#include <list>
#include <ctime>
#include <chrono>
#include <fenv.h>
#include <zlib.h>
#define private public
#define protected public
