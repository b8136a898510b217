// TEST INFRASTRUCTURE (oracle/): stand-in for DALI's data type tags, see ../../core/error_handling.h.
#pragma once
enum daliDataType_t { DALI_NO_TYPE = 0, DALI_FLOAT, DALI_BOOL, DALI_INT32, DALI_FLOAT_VEC };
namespace dali {
using DALIDataType = ::daliDataType_t;
struct CPUBackend {};
struct CPU {};
}  // namespace dali
