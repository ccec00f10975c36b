/* Force-included ahead of the UNMODIFIED reference sources (g++ -include):
 * supplies what MSVC / <windows.h> give the reference for free. */
#ifndef ICX_REF_SHIM_H
#define ICX_REF_SHIM_H
#define UNREFERENCED_PARAMETER(x) (void)(x)
#define GLM_ENABLE_EXPERIMENTAL
#include <cstring>
#include <cmath>
#include <array>
#include <functional>
#include <vector>
#include <set>
#include <unordered_map>
#include <algorithm>
#include <stdexcept>
#include <sstream>
#include <condition_variable>
#include <memory>
#include <string>
#include <cassert>
using std::isnan;
#endif
