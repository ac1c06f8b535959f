// Internal umbrella header of libg16b200 (not part of the C ABI).
#pragma once
#include "common.cuh"
#include "curve.cuh"
#include "field.cuh"
#include "hostfield.hpp"
#include "msm.cuh"
#include "ntt.cuh"

namespace g16 {
extern thread_local std::string g_last_error;
int fail(int code, const std::string& msg);
}  // namespace g16
