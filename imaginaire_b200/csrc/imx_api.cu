// Library-wide entry points: version and the per-thread error text.
#include "imx_common.cuh"
#include "../../include/imx.h"

namespace imx {
char* last_error_buf() {
    static thread_local char buf[512] = {0};
    return buf;
}
int set_error(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(last_error_buf(), 512, fmt, ap);
    va_end(ap);
    return code;
}
}  // namespace imx

extern "C" int imx_version(void) { return 100; }
extern "C" const char* imx_last_error(void) { return imx::last_error_buf(); }
