// Error plumbing shared by the C-ABI translation units.
#pragma once
#include <string>

namespace j2k {
std::string& last_error();                       // thread-local
int set_error(int code, const std::string& msg); // stores msg, returns code
}  // namespace j2k
