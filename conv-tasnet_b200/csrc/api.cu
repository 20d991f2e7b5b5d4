// Library-level plumbing of libctn_b200.so: error reporting, version, device query.
#include <stdarg.h>
#include <string.h>

#include <stdlib.h>

#include "common.cuh"

namespace ctn {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what) {
    set_error("%s: %s (%s)", what, cudaGetErrorName(e), cudaGetErrorString(e));
    return CTN_ERR_CUDA;
}

int pdl_enabled() {
    static int cached = -1;   // racing writers store the same value
    if (cached < 0) {
        const char* e = getenv("CTN_PDL");
        cached = (e != nullptr && e[0] == '0') ? 0 : 1;
    }
    return cached;
}

int sm_count() {
    static int cached[64] = {0};  // per device ordinal; racing writers store the same value
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (cached[dev] == 0) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cached[dev] = n;
    }
    return cached[dev];
}

}  // namespace ctn

extern "C" int ctn_version(void) { return 100; }
extern "C" int ctn_operand_format(void) { return CTN_MMA_TERMS + (CTN_OPERAND_FP16 ? 100 : 0); }
extern "C" const char* ctn_last_error(void) { return ctn::g_err; }
