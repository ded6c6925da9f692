// Shim: pcl_macros.h:48-56 only needs boost's names for the fixed-width integer typedefs.
#pragma once
#include <cstdint>
namespace boost {
using ::int8_t; using ::uint8_t; using ::int16_t; using ::uint16_t;
using ::int32_t; using ::uint32_t; using ::int64_t; using ::uint64_t; using ::int_fast16_t;
}
