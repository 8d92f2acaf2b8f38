// capi_common.h -- error plumbing shared by the C-ABI translation units.
#pragma once
#include <string>
#include <cuda_runtime.h>

namespace mfhe {
extern thread_local std::string g_last_error;
int fail(int code, const std::string& msg);
int cuda_fail(cudaError_t e, const char* where);
}  // namespace mfhe
