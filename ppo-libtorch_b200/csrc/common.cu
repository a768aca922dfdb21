// common.cu — error plumbing and small queries of the C ABI (include/ppo_b200.h).
#include "common.cuh"

#include <stdarg.h>

namespace ppo {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what, const char* file, int line) {
    set_error("CUDA error %d (%s) at %s:%d: %s", (int)e, cudaGetErrorString(e), file, line, what);
    return PPO_ERR_CUDA;
}

static int g_math_mode = PPO_MATH_FP16;
int math_mode() { return g_math_mode; }

int sm_count() {
    static int cached_dev = -1, cached = 0;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    if (dev != cached_dev) {
        int v = 0;
        if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) {
            cudaGetLastError();
            return 0;
        }
        cached = v;
        cached_dev = dev;
    }
    return cached;
}

int make_dims(const ppo_dims* d, Dims* out) {
    PPO_REQUIRE(d != nullptr, "dims is NULL");
    PPO_REQUIRE(d->obs_dim >= 1 && d->obs_dim <= 64, "obs_dim=%d out of [1,64]", d->obs_dim);
    PPO_REQUIRE(d->hidden == 64 || d->hidden == 128 || d->hidden == 256,
                "hidden=%d unsupported (64, 128 or 256)", d->hidden);
    PPO_REQUIRE(d->heads >= 1 && d->heads <= PPO_MAX_HEADS, "heads=%d out of [1,%d]", d->heads, PPO_MAX_HEADS);
    out->O = d->obs_dim;
    out->H = d->hidden;
    out->heads = d->heads;
    out->K = 0;
    for (int h = 0; h < PPO_MAX_HEADS; h++) out->nvec[h] = 0;
    for (int h = 0; h < d->heads; h++) {
        PPO_REQUIRE(d->nvec[h] >= 1, "nvec[%d]=%d < 1", h, d->nvec[h]);
        out->nvec[h] = d->nvec[h];
        out->K += d->nvec[h];
    }
    PPO_REQUIRE(out->K <= PPO_MAX_LOGITS, "sum(nvec)=%d exceeds %d", out->K, PPO_MAX_LOGITS);
    return PPO_OK;
}

}  // namespace ppo

using namespace ppo;

extern "C" const char* ppo_last_error(void) { return g_err; }
extern "C" int ppo_abi_version(void) { return PPO_B200_ABI_VERSION; }

namespace ppo {
int g_pdl = 1;              // bit 0: programmatic dependent launch on the single-GPU minibatch chain, bit 1: on the peer chain
long long g_spin_cycles = 0;   // 0 = the kernels' default (4e10 cycles, ~20 s)
}
extern "C" int ppo_debug_pdl(int on) {
    ppo::g_pdl = on;
    return PPO_OK;
}
// Upper bound of every in-kernel wait for another CTA / another rank (gradient exchange over peer memory, grid tickets).
// A wait that runs out sets the trainer's error flag, SKIPS the optimiser step (no rank steps on a partial gradient sum)
// and the next synchronised ppo_trainer_iterate returns PPO_ERR_COMM.
extern "C" int ppo_set_spin_timeout_ms(int64_t ms) {
    PPO_REQUIRE(ms >= 0, "ppo_set_spin_timeout_ms: negative");
    ppo::g_spin_cycles = ms * 1965000ll;   // SM clock ceiling of B200 (1 965 MHz)
    return PPO_OK;
}
extern "C" int ppo_get_math_mode(void) { return g_math_mode; }
extern "C" int ppo_set_math_mode(int mode) {
    PPO_REQUIRE(mode == PPO_MATH_FP32 || mode == PPO_MATH_BF16 || mode == PPO_MATH_FP16, "ppo_set_math_mode: mode=%d", mode);
    g_math_mode = mode;
    return PPO_OK;
}

extern "C" int ppo_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

extern "C" int ppo_device_info(int* sm, int* cc_major, int* cc_minor) {
    int dev = 0;
    PPO_CUDA(cudaGetDevice(&dev));
    cudaDeviceProp p;
    PPO_CUDA(cudaGetDeviceProperties(&p, dev));
    if (sm) *sm = p.multiProcessorCount;
    if (cc_major) *cc_major = p.major;
    if (cc_minor) *cc_minor = p.minor;
    return PPO_OK;
}

extern "C" int64_t ppo_num_params(const ppo_dims* dims) {
    Dims d;
    if (make_dims(dims, &d)) return -1;
    return num_params(d.O, d.H, d.K);
}
