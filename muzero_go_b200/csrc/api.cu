// Error plumbing, ABI version, device probe and launch accounting for libmzgo.so.
#include <stdlib.h>
#include <mutex>
#include <unordered_map>

#include "common.cuh"

static thread_local char g_err[512] = "";
static thread_local int64_t g_launches = 0;

void mzgo_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

void mzgo_count_launch() { g_launches++; }
void mzgo_count_launches(int k) { g_launches += k; }
int64_t mzgo_launch_count_peek() { return g_launches; }

static int g_pdl = -1;
bool mzgo_pdl_enabled() {
    if (g_pdl < 0) g_pdl = getenv("MZGO_NO_PDL") ? 0 : 1;
    return g_pdl == 1;
}
void mzgo_pdl_disable() { g_pdl = 0; }

// weight generation per network workspace (the unit mzgo_net_prepare works on): preparing one network's weights
// must not invalidate what was captured for another network (ArenaEvaluator's best_net) or another stream's workspace
static std::mutex g_gen_mu;
static std::unordered_map<const void*, unsigned long long> g_net_generation;
unsigned long long mzgo_net_generation(const void* workspace) {
    std::lock_guard<std::mutex> lk(g_gen_mu);
    auto it = g_net_generation.find(workspace);
    return it == g_net_generation.end() ? 0ull : it->second;
}
void mzgo_net_bump_generation(const void* workspace) {
    std::lock_guard<std::mutex> lk(g_gen_mu);
    g_net_generation[workspace]++;
}

extern "C" const char* mzgo_last_error(void) { return g_err; }

extern "C" int mzgo_abi_version(void) { return MZGO_ABI_VERSION; }

extern "C" int mzgo_device_ok(void) {
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) {
        cudaGetLastError();
        return 0;
    }
    int dev = 0, major = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 0;
    if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess) return 0;
    return major == 10;
}

extern "C" int64_t mzgo_launch_count(int reset) {
    int64_t v = g_launches;
    if (reset) g_launches = 0;
    return v;
}

extern "C" int64_t mzgo_sizeof(int which) {
    switch (which) {
        case 0: return sizeof(mzgo_tree_t);
        case 1: return sizeof(mzgo_search_params_t);
        case 2: return sizeof(mzgo_net_t);
        case 3: return sizeof(mzgo_selfplay_t);
    }
    return -1;
}
