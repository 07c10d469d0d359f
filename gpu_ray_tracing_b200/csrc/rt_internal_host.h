// rt_internal_host.h — error plumbing shared with the CUDA-free translation units.
#pragma once
namespace rti {
int fail(int code, const char* fmt, ...);
}
