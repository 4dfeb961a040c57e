#pragma once
namespace pcb {
// records a thread-local message for pcb_last_error() and returns `code`
int fail(int code, const char* fmt, ...);
}  // namespace pcb
