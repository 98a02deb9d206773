// Shared between libheat3d's translation units; not part of the public ABI.
#pragma once
namespace h3 {
// printf-style: stores the calling thread's heat3d_last_error() text and returns `code`.
int set_last_error(int code, const char* fmt, ...);
}
