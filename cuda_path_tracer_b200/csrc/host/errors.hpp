// errors.hpp -- per-thread last-error message behind ptb_last_error().
#pragma once
#include <string>
namespace ptb {
extern thread_local std::string g_last_error;
int fail(int code, const std::string& msg);   // stores msg, returns code
}
