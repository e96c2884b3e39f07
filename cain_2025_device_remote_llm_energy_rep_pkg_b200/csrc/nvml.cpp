// nvml.cpp — NVML energy / power / utilisation readings for the runner's energy-profiler plugin
// (BASELINE.json north_star: nvmlDeviceGetTotalEnergyConsumption written next to the existing
// run_table.csv columns; replaces the macOS `powermetrics` sampler of experiment/RunnerConfig.py:140).
// libnvidia-ml is opened at run time so the library still loads on a machine without a driver.
#include "../../include/b2llm.h"
#include <dlfcn.h>
#include <mutex>
#include <stdint.h>

void b2_set_error(const char *fmt, ...);

typedef int (*fn_init_t)(void);
typedef int (*fn_handle_t)(unsigned, void **);
typedef int (*fn_energy_t)(void *, unsigned long long *);
typedef int (*fn_power_t)(void *, unsigned *);
typedef struct { unsigned gpu, memory; } util_t;
typedef int (*fn_util_t)(void *, util_t *);
typedef int (*fn_clock_t)(void *, int, unsigned *);
typedef int (*fn_reasons_t)(void *, unsigned long long *);
typedef const char *(*fn_errstr_t)(int);

static struct {
    void *lib = nullptr;
    bool ok = false, tried = false;
    fn_handle_t handle = nullptr;
    fn_energy_t energy = nullptr;
    fn_power_t power = nullptr;
    fn_util_t util = nullptr;
    fn_clock_t clock = nullptr, max_clock = nullptr;
    fn_reasons_t reasons = nullptr;
    fn_errstr_t errstr = nullptr;
    std::mutex mu;
} g;

static int nvml_open() {
    std::lock_guard<std::mutex> lk(g.mu);
    if (g.tried) {
        if (!g.ok) b2_set_error("NVML unavailable (libnvidia-ml.so.1 not loadable or nvmlInit failed)");
        return g.ok ? B2_OK : B2_ENVML;
    }
    g.tried = true;
    g.lib = dlopen("libnvidia-ml.so.1", RTLD_NOW | RTLD_LOCAL);
    if (!g.lib) { b2_set_error("NVML unavailable: %s", dlerror()); return B2_ENVML; }
    fn_init_t init = (fn_init_t)dlsym(g.lib, "nvmlInit_v2");
    g.handle = (fn_handle_t)dlsym(g.lib, "nvmlDeviceGetHandleByIndex_v2");
    g.energy = (fn_energy_t)dlsym(g.lib, "nvmlDeviceGetTotalEnergyConsumption");
    g.power = (fn_power_t)dlsym(g.lib, "nvmlDeviceGetPowerUsage");
    g.util = (fn_util_t)dlsym(g.lib, "nvmlDeviceGetUtilizationRates");
    g.clock = (fn_clock_t)dlsym(g.lib, "nvmlDeviceGetClockInfo");
    g.max_clock = (fn_clock_t)dlsym(g.lib, "nvmlDeviceGetMaxClockInfo");
    g.reasons = (fn_reasons_t)dlsym(g.lib, "nvmlDeviceGetCurrentClocksEventReasons");
    if (!g.reasons) g.reasons = (fn_reasons_t)dlsym(g.lib, "nvmlDeviceGetCurrentClocksThrottleReasons");
    g.errstr = (fn_errstr_t)dlsym(g.lib, "nvmlErrorString");
    if (!init || !g.handle || !g.energy) { b2_set_error("NVML unavailable: missing symbols"); return B2_ENVML; }
    int r = init();
    if (r != 0) { b2_set_error("nvmlInit failed: %s", g.errstr ? g.errstr(r) : "?"); return B2_ENVML; }
    g.ok = true;
    return B2_OK;
}
static int nvml_dev(int ordinal, void **h) {
    int r = nvml_open();
    if (r) return r;
    int e = g.handle((unsigned)ordinal, h);
    if (e != 0) { b2_set_error("nvmlDeviceGetHandleByIndex(%d): %s", ordinal, g.errstr ? g.errstr(e) : "?"); return B2_ENVML; }
    return B2_OK;
}
#define NVML_CALL(fn, ...)                                                                    \
    do {                                                                                      \
        if (!(fn)) { b2_set_error("NVML symbol missing: " #fn); return B2_ENVML; }            \
        int _e = (fn)(__VA_ARGS__);                                                           \
        if (_e != 0) { b2_set_error(#fn ": %s", g.errstr ? g.errstr(_e) : "?"); return B2_ENVML; } \
    } while (0)

extern "C" int b2_nvml_energy_mj(int ordinal, uint64_t *out) {
    if (!out) { b2_set_error("b2_nvml_energy_mj: NULL"); return B2_EINVAL; }
    void *h;
    int r = nvml_dev(ordinal, &h);
    if (r) return r;
    unsigned long long mj = 0;
    NVML_CALL(g.energy, h, &mj);
    *out = mj;
    return B2_OK;
}
extern "C" int b2_nvml_power_mw(int ordinal, uint32_t *out) {
    if (!out) { b2_set_error("b2_nvml_power_mw: NULL"); return B2_EINVAL; }
    void *h;
    int r = nvml_dev(ordinal, &h);
    if (r) return r;
    unsigned mw = 0;
    NVML_CALL(g.power, h, &mw);
    *out = mw;
    return B2_OK;
}
extern "C" int b2_nvml_util(int ordinal, uint32_t *gpu_pct, uint32_t *mem_pct) {
    void *h;
    int r = nvml_dev(ordinal, &h);
    if (r) return r;
    util_t u{};
    NVML_CALL(g.util, h, &u);
    if (gpu_pct) *gpu_pct = u.gpu;
    if (mem_pct) *mem_pct = u.memory;
    return B2_OK;
}
extern "C" int b2_nvml_clocks(int ordinal, uint32_t *sm_mhz, uint32_t *sm_max_mhz, uint64_t *reasons) {
    void *h;
    int r = nvml_dev(ordinal, &h);
    if (r) return r;
    unsigned c = 0, m = 0;
    unsigned long long rs = 0;
    NVML_CALL(g.clock, h, 1 /* NVML_CLOCK_SM */, &c);
    NVML_CALL(g.max_clock, h, 1, &m);
    NVML_CALL(g.reasons, h, &rs);
    if (sm_mhz) *sm_mhz = c;
    if (sm_max_mhz) *sm_max_mhz = m;
    if (reasons) *reasons = rs;
    return B2_OK;
}
