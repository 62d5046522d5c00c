// Shared host/device helpers of the native library (sm_100a only).
//
// There is no native code in the reference; everything here is new.  The PTX
// wrappers are the Blackwell primitives the kernels are built from: mbarrier + TMA
// (cp.async.bulk.tensor) for tile staging, and system-scope acquire/release accesses for the
// cross-GPU flag protocol over NVLink peer memory.
#pragma once

#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../include/igg_b200.h"

// ---------------------------------------------------------------------------------------------
// host side: error plumbing + launch accounting
// ---------------------------------------------------------------------------------------------
namespace igg {

void set_error(const char* fmt, ...);          // stores a message for igg_last_error()
void count_launch(int n = 1);                  // bumps the counter behind igg_launch_count()

#define IGG_CUDA_OK(expr)                                                                      \
    do {                                                                                       \
        cudaError_t _e = (expr);                                                               \
        if (_e != cudaSuccess) {                                                               \
            igg::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__,   \
                           __LINE__);                                                          \
            return (int)_e;                                                                    \
        }                                                                                      \
    } while (0)

#define IGG_REQUIRE(cond, ...)                                                                 \
    do {                                                                                       \
        if (!(cond)) {                                                                         \
            igg::set_error(__VA_ARGS__);                                                       \
            return IGG_ERR_INVALID;                                                            \
        }                                                                                      \
    } while (0)

// returns cudaPeekAtLastError-style status after a launch and counts it
#define IGG_LAUNCH_CHECK()                                                                     \
    do {                                                                                       \
        cudaError_t _e = cudaGetLastError();                                                   \
        if (_e != cudaSuccess) {                                                               \
            igg::set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(_e),         \
                           __FILE__, __LINE__);                                                \
            return (int)_e;                                                                    \
        }                                                                                      \
        igg::count_launch();                                                                   \
    } while (0)

int sm_count();                                  // SMs of the current device (cached)

// Encode (and cache) a 2-D/3-D tiled fp64 tensor map.  dims/box are innermost-first.
int get_tensor_map_f64(const void* base, int rank, const uint64_t* dims, const uint32_t* box,
                       CUtensorMap* out);

static inline int ceil_div(int a, int b) { return (a + b - 1) / b; }

}  // namespace igg

// ---------------------------------------------------------------------------------------------
// device side
// ---------------------------------------------------------------------------------------------
#ifdef __CUDACC__
namespace igg {

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}

// ---- mbarrier -------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_mbar_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    while (!mbar_try_wait(bar, parity)) {
    }
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// barrier among the consumer threads of a warp-specialised kernel only (named barrier 1; the producer warp never
// joins it)
__device__ __forceinline__ void consumer_sync(int nthreads) {
    asm volatile("bar.sync 1, %0;" ::"r"(nthreads) : "memory");
}

// ---- TMA (cp.async.bulk.tensor) --------------------------------------------------------------
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* m) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(m) : "memory");
}
// global -> shared, 2-D tile; c0 = innermost coordinate.  Out-of-bounds elements are zero filled.
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* m, uint64_t* bar,
                                            int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4}], [%2];"
        ::"r"(smem_u32(smem_dst)), "l"(m), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
        : "memory");
}
// same with an L2 eviction-priority hint (policy words as used by CUTLASS: evict-first for data that is
// read exactly once, so it does not push reusable halo lines out of L2)
constexpr uint64_t L2_EVICT_NORMAL = 0x1000000000000000ull;
constexpr uint64_t L2_EVICT_FIRST = 0x12F0000000000000ull;
constexpr uint64_t L2_EVICT_LAST = 0x14F0000000000000ull;
__device__ __forceinline__ void tma_load_2d_hint(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c0,
                                                 int c1, uint64_t policy) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes.L2::cache_hint"
        " [%0], [%1, {%3, %4}], [%2], %5;"
        ::"r"(smem_u32(smem_dst)), "l"(m), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "l"(policy)
        : "memory");
}
__device__ __forceinline__ void tma_load_3d_hint(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c0,
                                                 int c1, int c2, uint64_t policy) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes.L2::cache_hint"
        " [%0], [%1, {%3, %4, %5}], [%2], %6;"
        ::"r"(smem_u32(smem_dst)), "l"(m), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "l"(policy)
        : "memory");
}
__device__ __forceinline__ void tma_load_3d(void* smem_dst, const CUtensorMap* m, uint64_t* bar,
                                            int c0, int c1, int c2) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4, %5}], [%2];"
        ::"r"(smem_u32(smem_dst)), "l"(m), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
        : "memory");
}
// Order earlier generic-proxy observations (e.g. an acquired flag) before later async-proxy
// (TMA) reads of global memory.
__device__ __forceinline__ void fence_proxy_async() {
    asm volatile("fence.proxy.async;" ::: "memory");
}

// ---- cross-GPU flags ---------------------------------------------------------------------------
__device__ __forceinline__ uint64_t ld_acquire_sys(const uint64_t* p) {
    uint64_t v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ uint64_t ld_relaxed_sys(const uint64_t* p) {
    uint64_t v;
    asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(uint64_t* p, uint64_t v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
// Spin until *p >= want.  A bounded number of polls keeps a protocol bug from hanging the GPU:
// on timeout the error word is set and the caller proceeds (the host checks it afterwards).
__device__ __forceinline__ bool wait_flag_ge(const uint64_t* p, uint64_t want, uint64_t* err_word) {
    uint64_t spins = 0;
    while (ld_acquire_sys(p) < want) {
        __nanosleep(64);
        if (++spins > (1ull << 26)) {   // ~ seconds
            if (err_word) atomicExch((unsigned long long*)err_word, (unsigned long long)want | (1ull << 63));
            return false;
        }
    }
    return true;
}

// Variant of wait_flag_ge that polls with relaxed system-scope loads and issues a single acquire fence
// once the value is there (mode 1), or no fence at all (mode 2, timing experiments only).
__device__ __forceinline__ bool wait_flag_ge_relaxed(const uint64_t* p, uint64_t want, uint64_t* err_word, int mode) {
    uint64_t spins = 0;
    while (ld_relaxed_sys(p) < want) {
        __nanosleep(64);
        if (++spins > (1ull << 26)) {
            if (err_word) atomicExch((unsigned long long*)err_word, (unsigned long long)want | (1ull << 63));
            return false;
        }
    }
    if (mode == 1) asm volatile("fence.acq_rel.sys;" ::: "memory");
    return true;
}

// max that propagates NaN (fmax drops it): a blown-up field must show up in the max diagnostic the way it does in
// the reference's maximum(T_v) (scripts/diffusion_2D_perf_hide.jl:115)
__device__ __forceinline__ double max_nan(double a, double b) { return (a > b || a != a) ? a : b; }

// One cell of the reference's THREE-kernel form fused into one expression with its roundings kept
// (scripts/diffusion_2D_kp.jl: Flux! :20,:23 -> Residual! :37 -> Update! :51): the fluxes are rounded products
// (they are stored to qx / qy there, so no multiply may be contracted into the following subtraction),
// dTdt = 1.0/Cp * -((qxR - qxL)*_dx + (qyT - qyB)*_dy) -- a DIVISION by Cp where the perf form multiplies --
// and T + dt*dTdt.  Used by the single-launch multi-step kernels when they serve the kp application.
__device__ __forceinline__ double update_cell_kp(double c, double xm, double xp, double ym, double yp, double cp,
                                                 double lam, double dt, double _dx, double _dy) {
    const double qxr = __dmul_rn(__dmul_rn(-lam, xp - c), _dx);
    const double qxl = __dmul_rn(__dmul_rn(-lam, c - xm), _dx);
    const double qyt = __dmul_rn(__dmul_rn(-lam, yp - c), _dy);
    const double qyb = __dmul_rn(__dmul_rn(-lam, c - ym), _dy);
    const double dqx = qxr - qxl, dqy = qyt - qyb;
    const double dTdt = 1.0 / cp * -(dqx * _dx + dqy * _dy);
    return c + dt * dTdt;
}

// max for the hot loops of the in-kernel diagnostics: one compare + select (fmax() expands to a NaN-aware sequence
// of ~5 instructions on sm_100a).  It drops NaNs like fmax; the kernels restore them from the sum at the final fold.
__device__ __forceinline__ double max_fast(double a, double b) { return a > b ? a : b; }

// streaming (evict-first) 16-byte / 8-byte global stores and read-only loads
__device__ __forceinline__ void st_cs_f64x2(double* p, double a, double b) {
    asm volatile("st.global.cs.v2.f64 [%0], {%1, %2};" ::"l"(p), "d"(a), "d"(b) : "memory");
}
__device__ __forceinline__ void st_cs_f64(double* p, double a) {
    asm volatile("st.global.cs.f64 [%0], %1;" ::"l"(p), "d"(a) : "memory");
}

}  // namespace igg
#endif  // __CUDACC__
