// api_internal.h — what the translation units that carry their own extern "C" entry points share with api.cu.
#pragma once

namespace rnnt {
// records the thread-local message rnnt_last_error() returns and hands back `code`
int api_fail(int code, const char* fmt, ...);
}  // namespace rnnt
