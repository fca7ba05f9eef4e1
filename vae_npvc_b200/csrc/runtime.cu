// Error string, launch counter and backend switch of the C ABI.
#include <stdarg.h>
#include <stdlib.h>

#include <atomic>

#include "common.cuh"

namespace vnb {

static thread_local char g_err[512] = "";
static std::atomic<int64_t> g_launches{0};
static std::atomic<int> g_backend{VNB_BACKEND_AUTO};

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
int backend() { return g_backend.load(std::memory_order_relaxed); }
// Runtime options (vnb_set_option): initial values from the environment, else the compiled defaults.
static int env_or(const char* name, int dflt) {
    const char* e = getenv(name);
    return e ? atoi(e) : dflt;
}
static std::atomic<int> g_opts[VNB_OPT_COUNT_] = {};
static std::atomic<bool> g_opts_init{false};
static void opts_init() {
    if (g_opts_init.load(std::memory_order_acquire)) return;
    g_opts[VNB_OPT_EW_V2].store(env_or("VNB_EW_V2", VNB_EW_V2_DEFAULT));
    g_opts[VNB_OPT_FAST_GATE].store(env_or("VNB_FAST_GATE", VNB_FAST_GATE_DEFAULT));
    g_opts[VNB_OPT_WN_V2].store(env_or("VNB_WN_V2", VNB_WN_V2_DEFAULT));
    g_opts[VNB_OPT_GEMM_V3].store(env_or("VNB_GEMM_V3", 1));
    g_opts[VNB_OPT_TC3_BN].store(env_or("VNB_TC3_BN", 0));
    g_opts[VNB_OPT_TC3_PAIR].store(env_or("VNB_TC3_PAIR", -1));
    g_opts[VNB_OPT_TC3_STAGES].store(env_or("VNB_TC3_STAGES", 0));
    g_opts[VNB_OPT_TC_DBG].store(env_or("VNB_TC_DBG", 0));
    g_opts[VNB_OPT_TC3_SMEM_KB].store(env_or("VNB_TC3_SMEM_KB", 0));
    g_opts[VNB_OPT_FAST_EPI].store(env_or("VNB_FAST_EPI", 0));
    g_opts[VNB_OPT_VQ_FUSED].store(env_or("VNB_VQ_FUSED", 1));
    g_opts[VNB_OPT_WG_FUSE].store(env_or("VNB_WG_FUSE", 0));
    g_opts_init.store(true, std::memory_order_release);
}
int option(int which) {
    opts_init();
    return (which >= 0 && which < VNB_OPT_COUNT_) ? g_opts[which].load(std::memory_order_relaxed) : 0;
}
void set_option(int which, int value) {
    opts_init();
    if (which >= 0 && which < VNB_OPT_COUNT_) g_opts[which].store(value);
}
bool pdl_enabled() {
    static const bool on = !(getenv("VNB_PDL") && atoi(getenv("VNB_PDL")) == 0);
    return on;
}

}  // namespace vnb

namespace vnb {
__global__ void debug_stamp_kernel(long long* buf, int slot) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    buf[slot] = (long long)t;
}
}  // namespace vnb

// diagnostics: write the GPU's nanosecond timer into buf[slot] when the stream reaches this point (works inside a
// captured graph, where CUDA events cannot be read back) - tools/diag_ddp_timeline.py
extern "C" int vnb_debug_stamp(int64_t* buf, int slot, void* stream) {
    VNB_REQUIRE(buf && slot >= 0, "debug_stamp: bad arguments");
    vnb::debug_stamp_kernel<<<1, 1, 0, (cudaStream_t)stream>>>((long long*)buf, slot);
    VNB_LAUNCH_CHECK("debug_stamp");
    return VNB_OK;
}

extern "C" {
const char* vnb_last_error(void) { return vnb::g_err; }
int vnb_version(void) { return 100; }
int64_t vnb_launch_count(void) { return vnb::g_launches.load(); }
void vnb_launch_count_reset(void) { vnb::g_launches.store(0); }
void vnb_set_backend(int b) { vnb::g_backend.store(b); }
int vnb_get_backend(void) { return vnb::g_backend.load(); }
void vnb_set_option(int which, int value) { vnb::set_option(which, value); }
int vnb_get_option(int which) { return vnb::option(which); }
}
