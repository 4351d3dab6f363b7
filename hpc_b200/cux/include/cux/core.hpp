// core.hpp -- enums, logging, type tags and small utilities of the cux-compatible host layer.
//
// This layer re-provides the public surface of cjmcv/hpc's cux framework (namespace cux: Executor,
// Operator, OpFactory, OpAssistor, ResultChecker, Array4D, KernelInterface structs, Config1D/2D,
// GpuTimer/CpuTimer, TypeFlag, OpRunMode, DataFetchMode) so that code written against
// 0-frameworks/cux/include/*.h builds unchanged, but it contains no CUDA: every device action goes
// through the C ABI in include/cuxb.h.  Names and semantics follow
// 0-frameworks/cux/include/util/util.h:22-176; the implementation is new.
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include "cuxb.h"

// cux's Config2D/3D use CUDA's dim3.  Host code here never includes a CUDA header, so provide a
// layout-compatible stand-in unless CUDA's own vector types are already visible (user .cu plugins).
#if !defined(__VECTOR_TYPES_H__) && !defined(__CUDACC__) && !defined(CUX_HAVE_DIM3)
#define CUX_HAVE_DIM3
struct dim3 {
  unsigned int x, y, z;
  constexpr dim3(unsigned int vx = 1, unsigned int vy = 1, unsigned int vz = 1) : x(vx), y(vy), z(vz) {}
};
#endif

namespace cux {

// util.h:22-28, plus BFLOAT16 for the bf16 linear path (simd/x86/linear.cpp:137-182).
enum TypeFlag { FLOAT32 = 0, INT32 = 1, FLOAT16 = 2, INT8 = 3, BFLOAT16 = 4, TYPES_NUM = 5 };
enum OpRunMode { ON_HOST, ON_DEVICE };                 // util.h:30-33
enum Shape { NUMBER, CHANNELS, HEIGHT, WIDTH };        // util.h:35-40
enum DataFetchMode { NO_PUSH, PUSH, PUSH_IF_EMPTY };   // util.h:42-46

}  // namespace cux

// ---- logging (util.h:91-98).  CUXLOG_ERR aborts, exactly like the reference.
#define CUXLOG_ERR(...)                                            \
  do {                                                             \
    std::fprintf(stderr, "[ERROR]: (%s: %d)", __FILE__, __LINE__); \
    std::fprintf(stderr, __VA_ARGS__);                             \
    std::fprintf(stderr, "\n");                                    \
    std::abort();                                                  \
  } while (0)
#define CUXLOG_WARN(...) do { std::fprintf(stdout, "[WARN]: "); std::fprintf(stdout, __VA_ARGS__); std::fprintf(stdout, "\n"); } while (0)
#define CUXLOG_INFO(...) do { std::fprintf(stdout, "[INFO]: "); std::fprintf(stdout, __VA_ARGS__); std::fprintf(stdout, "\n"); } while (0)
#define CUXLOG_COUT(...) do { std::fprintf(stdout, "> "); std::fprintf(stdout, __VA_ARGS__); std::fprintf(stdout, "\n"); } while (0)

// The reference wraps CUDA calls in CUDA_CHECK / CUBLAS_CHECK, which print and exit(EXIT_FAILURE)
// (util.h:69-88).  The C ABI returns codes instead; this maps them back onto that behaviour.
#define CUXB_CHECK(call)                                                                        \
  do {                                                                                          \
    const int cuxb_rc__ = (call);                                                               \
    if (cuxb_rc__ != 0) {                                                                       \
      std::fprintf(stderr, "CUXB_CHECK error in line %d of file %s : (%d) %s \n", __LINE__,     \
                   __FILE__, cuxb_rc__, cuxb_last_error_string());                              \
      std::exit(EXIT_FAILURE);                                                                  \
    }                                                                                           \
  } while (0)
#define CUDA_CHECK(call) CUXB_CHECK(call)

#define FLT_PRECISION_ERR_THRESH (0.001)  // util.h:137, the checker's pass threshold

#include "cux/half.hpp"

namespace cux {

// util.h:53-64: "key: value, key: value" parameter strings.
class StrProcessor {
public:
  static std::string FetchSubStr(std::string &src_str, std::string start_str, std::string end_str) {
    const size_t at = src_str.find(start_str);
    const size_t from = (at == std::string::npos ? 0 : at) + start_str.length();
    const size_t to = src_str.find(end_str, from);
    return src_str.substr(from, to == std::string::npos ? std::string::npos : to - from);
  }
};

// What the host needs to know about a device (the reference keeps a whole cudaDeviceProp,
// util.h:142-145; these are the fields cux actually reads).
struct DeviceProp {
  char name[256] = {0};
  int major = 0, minor = 0;
  int multiProcessorCount = 0;
  int maxThreadsPerMultiProcessor = 0;
  int maxThreadsPerBlock = 0;
  int warpSize = 32;
  size_t totalGlobalMem = 0;
  size_t sharedMemPerBlock = 0;
};
struct Device {
  int id = 0;
  DeviceProp prop;
};

// util.h:147-174
template <typename DType> struct DataType;
template <> struct DataType<float> { static const int kFlag = FLOAT32; };
template <> struct DataType<half> { static const int kFlag = FLOAT16; };
template <> struct DataType<bfloat16> { static const int kFlag = BFLOAT16; };
template <> struct DataType<int32_t> { static const int kFlag = INT32; };
template <> struct DataType<uint32_t> { static const int kFlag = INT32; };
template <> struct DataType<int8_t> { static const int kFlag = INT8; };
template <> struct DataType<uint8_t> { static const int kFlag = INT8; };

inline size_t TypeSize(int type_flag) {
  switch (type_flag) {
    case FLOAT32: case INT32: return 4;
    case FLOAT16: case BFLOAT16: return 2;
    case INT8: return 1;
    default: CUXLOG_ERR("Unknown type enum %d", type_flag);
  }
  return 0;
}

}  // namespace cux

// util.h:107-134
#define TYPE_SWITCH(type, DType, ...)                                   \
  switch (type) {                                                       \
    case cux::TypeFlag::FLOAT32: { typedef float DType; {__VA_ARGS__} } break;        \
    case cux::TypeFlag::FLOAT16: { typedef cux::half DType; {__VA_ARGS__} } break;    \
    case cux::TypeFlag::BFLOAT16: { typedef cux::bfloat16 DType; {__VA_ARGS__} } break; \
    case cux::TypeFlag::INT32: { typedef int32_t DType; {__VA_ARGS__} } break;        \
    case cux::TypeFlag::INT8: { typedef int8_t DType; {__VA_ARGS__} } break;          \
    default: CUXLOG_ERR("Unknown type enum %d", (int)(type));            \
  }
