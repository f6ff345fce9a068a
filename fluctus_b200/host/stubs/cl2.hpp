// Typedef-only stand-in for the OpenCL C++ header, so that the reference's HOST sources
// (scene/bvh/sbvh/envmap/settings) compile headless. Used by the headless host build (fluctus_b200/host) and by oracle/ref_host.
#pragma once
#include <cstdint>
#include <cstddef>
#include <limits>
#include <cassert>
#include <stdexcept>
#include <sstream>
#include <chrono>
#include <functional>
#include <map>
#include <string>
#include <vector>
#include <cmath>
#include <cstring>
#include <algorithm>
#include <sys/stat.h>
typedef float cl_float;
typedef int32_t cl_int;
typedef uint32_t cl_uint;
typedef unsigned char cl_uchar;
typedef uint64_t cl_ulong;
typedef bool cl_bool;
#define CL_PLATFORM_NAME 0x0902
namespace cl {
struct Platform { template <int N> std::string getInfo() const { return "headless"; } };
struct Device {}; struct Context {}; struct CommandQueue {}; struct Memory {};
struct Buffer : Memory {}; struct BufferGL : Buffer {}; struct Image2D : Memory {}; struct Event {};
}
