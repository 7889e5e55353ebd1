// Error reporting and device gate of libmdt_b200 (SURVEY.md section 8b "Error convention").
#include "common.cuh"

namespace mdt {

char *last_error_buffer() {
    static thread_local char buf[512] = "";
    return buf;
}

int fail(int code, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(last_error_buffer(), 512, fmt, ap);
    va_end(ap);
    return code;
}

int require_sm100() {
    static int ok = -1;
    if (ok < 0) {
        int dev = 0, major = 0;
        if (cudaGetDevice(&dev) != cudaSuccess ||
            cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess) {
            cudaGetLastError();
            return fail(MDT_ECUDA, "no CUDA device: libmdt_b200 has no CPU path");
        }
        ok = (major == 10) ? 1 : 0;
    }
    if (!ok) return fail(MDT_ENOSUP, "libmdt_b200 is built for sm_100a (B200) only");
    return MDT_OK;
}

}  // namespace mdt

extern "C" int mdt_abi_version(void) { return MDT_ABI_VERSION; }
extern "C" const char *mdt_last_error_string(void) { return mdt::last_error_buffer(); }
extern "C" int mdt_device_ok(void) { return mdt::require_sm100() == MDT_OK ? 1 : 0; }
