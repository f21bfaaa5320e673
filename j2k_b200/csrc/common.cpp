#include "common.h"
namespace j2k {
std::string& last_error() { static thread_local std::string e; return e; }
int set_error(int code, const std::string& msg) { last_error() = msg; return code; }
}  // namespace j2k
