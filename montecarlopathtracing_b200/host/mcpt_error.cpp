// Thread-local last-error string shared by the host shell and the CUDA layer.
#include <string>
#include "../../include/mcpt.h"
static thread_local std::string g_last_error;
void mcpt_set_error(const char* msg) { g_last_error = msg ? msg : ""; }
extern "C" MCPT_API const char* mcpt_last_error(void) { return g_last_error.c_str(); }
