// Error plumbing and bookkeeping of the C ABI (include/tss_b200.h).
#include "common.cuh"

#include <atomic>

namespace tss {

static thread_local char g_err[512] = "";
static std::atomic<uint64_t> g_launches{0};

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int cuda_status(cudaError_t e, const char* what) {
    set_error("%s: CUDA error %d (%s)", what, static_cast<int>(e), cudaGetErrorString(e));
    return TSS_E_CUDA + static_cast<int>(e);
}

void count_launch(int n) { g_launches.fetch_add(static_cast<uint64_t>(n), std::memory_order_relaxed); }

}  // namespace tss

extern "C" {

int tss_version(void) { return TSS_VERSION; }

const char* tss_last_error_string(void) { return tss::g_err; }

uint64_t tss_launch_count(void) { return tss::g_launches.load(std::memory_order_relaxed); }

#ifndef TSS_SOURCE_DIGEST
#define TSS_SOURCE_DIGEST ""
#endif
// the marker lets the host mirror read the digest from the FILE (a dlopen of a stale library would pin it in the process)
static const char kSourceDigest[] = "TSS_SOURCE_DIGEST=" TSS_SOURCE_DIGEST;
const char* tss_source_digest(void) { return kSourceDigest + 18; }

}  // extern "C"
