// Peer-memory fabric: allocation, CUDA-IPC export/import, and the communication kernels that use
// peer-mapped pointers (one-launch update_halo, neighbour barrier, epoch wait, peer copy).
//
// This layer stands where the reference has GPU-aware MPI: device buffers handed directly to
// MPI.Sendrecv! (scripts/rocmaware_test_selectdevice.jl:15-22) and, through ImplicitGlobalGrid, to
// Isend/Irecv inside update_halo! (call site scripts/diffusion_2D_ap.jl:42).  On an NVSwitch node every
// GPU can address every other GPU's memory, so "send" is a store and "receive complete" is a flag.
#include <string.h>

#include <mutex>

#include "common.cuh"

namespace {
using namespace igg;

typedef CUresult (*GetAddressRangeFn)(CUdeviceptr*, size_t*, CUdeviceptr);
static GetAddressRangeFn address_range_fn() {
    static GetAddressRangeFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuMemGetAddressRange", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = (GetAddressRangeFn)p;
    });
    return fn;
}

// ---- one-launch update_halo ---------------------------------------------------------------------
constexpr int HX_THREADS = 256;

__global__ void __launch_bounds__(HX_THREADS) halo_exchange_kernel(const igg_halo_xfer* __restrict__ xf, int nx,
                                                                   const __grid_constant__ igg_halo_sync sy) {
    const int tid = threadIdx.x;
    uint64_t* err = sy.my_bank + IGG_SLOT_ERR;
    // (1) tell every neighbour that my halo cells may be overwritten (all earlier work on this
    //     stream has finished), then wait for the same promise from them.  The promise I make to the
    //     neighbour at offset o shows up in ITS bank under the opposite offset 26-o.
    if (blockIdx.x == 0 && tid < 27 && sy.peer_bank[tid])
        st_release_sys(sy.peer_bank[tid] + IGG_SLOT_READY + (26 - tid), sy.epoch);
    if (tid < 27 && sy.peer_bank[tid]) wait_flag_ge(sy.my_bank + IGG_SLOT_READY + tid, sy.epoch, err);
    __syncthreads();
    // (2) store every box into the neighbours' arrays
    for (int k = 0; k < nx; ++k) {
        const igg_halo_xfer t = xf[k];
        const size_t e2 = (size_t)t.extent[2], e1 = (size_t)t.extent[1];
        const size_t count = (size_t)t.extent[0] * e1 * e2;
        for (size_t e = (size_t)blockIdx.x * HX_THREADS + tid; e < count; e += (size_t)gridDim.x * HX_THREADS) {
            const size_t c = e % e2, b = (e / e2) % e1, a = e / (e1 * e2);
            const size_t si = ((a + t.src_start[0]) * t.n[1] + (b + t.src_start[1])) * t.n[2] + (c + t.src_start[2]);
            const size_t di = ((a + t.dst_start[0]) * t.n[1] + (b + t.dst_start[1])) * t.n[2] + (c + t.dst_start[2]);
            switch (t.elem_bytes) {
                case 8: ((uint64_t*)t.dst)[di] = ((const uint64_t*)t.src)[si]; break;
                case 4: ((uint32_t*)t.dst)[di] = ((const uint32_t*)t.src)[si]; break;
                case 2: ((uint16_t*)t.dst)[di] = ((const uint16_t*)t.src)[si]; break;
                default: ((uint8_t*)t.dst)[di] = ((const uint8_t*)t.src)[si]; break;
            }
        }
    }
    // (3) last CTA to finish releases the epoch to every neighbour
    __syncthreads();
    __shared__ bool last;
    if (tid == 0) {
        __threadfence_system();
        const unsigned long long old = atomicAdd((unsigned long long*)(sy.my_count_bank + IGG_SLOT_CTA_COUNT), 1ull);
        last = ((old + 1ull) % gridDim.x) == 0ull;
        if (last) __threadfence_system();
    }
    __syncthreads();
    if (last && tid < 27 && sy.peer_bank[tid])
        st_release_sys(sy.peer_bank[tid] + IGG_SLOT_DATA + (26 - tid), sy.epoch);
    // (4) the launch is complete only when my own halo cells have been filled
    if (tid < 27 && sy.peer_bank[tid]) wait_flag_ge(sy.my_bank + IGG_SLOT_DATA + tid, sy.epoch, err);
}

__global__ void neighbor_barrier_kernel(const __grid_constant__ igg_halo_sync sy) {
    const int tid = threadIdx.x;
    if (tid < 27 && sy.peer_bank[tid]) {
        __threadfence_system();
        st_release_sys(sy.peer_bank[tid] + IGG_SLOT_DATA + (26 - tid), sy.epoch);
        wait_flag_ge(sy.my_bank + IGG_SLOT_DATA + tid, sy.epoch, sy.my_bank + IGG_SLOT_ERR);
    }
}

__global__ void wait_epoch_kernel(const __grid_constant__ igg_halo_sync sy) {
    const int tid = threadIdx.x;
    if (tid < 27 && sy.peer_bank[tid])
        wait_flag_ge(sy.my_bank + IGG_SLOT_DATA + tid, sy.epoch, sy.my_bank + IGG_SLOT_ERR);
}

__global__ void wait_value_kernel(uint64_t* flag, uint64_t value, uint64_t* err) {
    wait_flag_ge(flag, value, err);
}

// ---- peer copy ------------------------------------------------------------------------------------
constexpr int PC_THREADS = 512;
__global__ void __launch_bounds__(PC_THREADS) peer_copy_kernel(uint4* __restrict__ dst, const uint4* __restrict__ src,
                                                               size_t n16, unsigned char* dst_tail,
                                                               const unsigned char* src_tail, int tail,
                                                               uint64_t* peer_flag, uint64_t value, uint64_t* my_count) {
    const size_t stride = (size_t)gridDim.x * PC_THREADS;
    size_t e = (size_t)blockIdx.x * PC_THREADS + threadIdx.x;
    // 4 independent 16-byte loads in flight per thread
    for (; e + 3 * stride < n16; e += 4 * stride) {
        const uint4 v0 = src[e], v1 = src[e + stride], v2 = src[e + 2 * stride], v3 = src[e + 3 * stride];
        dst[e] = v0;
        dst[e + stride] = v1;
        dst[e + 2 * stride] = v2;
        dst[e + 3 * stride] = v3;
    }
    for (; e < n16; e += stride) dst[e] = src[e];
    if (blockIdx.x == 0 && (int)threadIdx.x < tail) dst_tail[threadIdx.x] = src_tail[threadIdx.x];
    if (peer_flag) {
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence_system();
            const unsigned long long old = atomicAdd((unsigned long long*)my_count, 1ull);
            if (((old + 1ull) % gridDim.x) == 0ull) {
                __threadfence_system();
                st_release_sys(peer_flag, value);
            }
        }
    }
}

}  // namespace

extern "C" {

int igg_fabric_alloc(size_t bytes, void** ptr) {
    IGG_REQUIRE(ptr && bytes > 0, "igg_fabric_alloc: bad arguments");
    void* p = nullptr;
    IGG_CUDA_OK(cudaMalloc(&p, bytes));
    cudaError_t e = cudaMemset(p, 0, bytes);
    if (e != cudaSuccess) {
        cudaFree(p);
        igg::set_error("igg_fabric_alloc: cudaMemset failed: %s", cudaGetErrorString(e));
        return (int)e;
    }
    IGG_CUDA_OK(cudaDeviceSynchronize());
    *ptr = p;
    return 0;
}

int igg_fabric_free(void* ptr) {
    if (ptr) IGG_CUDA_OK(cudaFree(ptr));
    return 0;
}

int igg_ipc_export(const void* ptr, void* handle64, uint64_t* offset) {
    IGG_REQUIRE(ptr && handle64 && offset, "igg_ipc_export: bad arguments");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    cudaIpcMemHandle_t h;
    IGG_CUDA_OK(cudaIpcGetMemHandle(&h, const_cast<void*>(ptr)));
    memcpy(handle64, &h, 64);
    *offset = 0;
    GetAddressRangeFn fn = address_range_fn();
    if (fn) {
        CUdeviceptr base = 0;
        size_t size = 0;
        if (fn(&base, &size, (CUdeviceptr)(uintptr_t)ptr) == CUDA_SUCCESS) *offset = (uint64_t)((uintptr_t)ptr - (uintptr_t)base);
    }
    return 0;
}

int igg_ipc_open(const void* handle64, void** base) {
    IGG_REQUIRE(handle64 && base, "igg_ipc_open: bad arguments");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, 64);
    IGG_CUDA_OK(cudaIpcOpenMemHandle(base, h, cudaIpcMemLazyEnablePeerAccess));
    return 0;
}

int igg_ipc_close(void* base) {
    if (base) IGG_CUDA_OK(cudaIpcCloseMemHandle(base));
    return 0;
}

int igg_can_access_peer(int dev, int peer_dev, int* can) {
    IGG_REQUIRE(can, "igg_can_access_peer: bad arguments");
    IGG_CUDA_OK(cudaDeviceCanAccessPeer(can, dev, peer_dev));
    return 0;
}

int igg_halo_exchange_peer(const igg_halo_xfer* xfers_dev, int nxfers, const igg_halo_sync* sync, void* stream) {
    IGG_REQUIRE(sync && sync->my_bank && sync->my_count_bank && nxfers >= 0 && (nxfers == 0 || xfers_dev),
                "igg_halo_exchange_peer: bad arguments");
    IGG_REQUIRE(sync->epoch >= 1, "igg_halo_exchange_peer: epoch must be >= 1");
    int grid = igg::sm_count() / 4;   // a quarter of the SMs saturates NVLink for plane-sized messages
    if (grid < 1) grid = 1;
    halo_exchange_kernel<<<grid, HX_THREADS, 0, (cudaStream_t)stream>>>(xfers_dev, nxfers, *sync);
    IGG_LAUNCH_CHECK();
    return 0;
}

int igg_neighbor_barrier(const igg_halo_sync* sync, void* stream) {
    IGG_REQUIRE(sync && sync->my_bank && sync->epoch >= 1, "igg_neighbor_barrier: bad arguments");
    neighbor_barrier_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(*sync);
    IGG_LAUNCH_CHECK();
    return 0;
}

int igg_wait_epoch(const igg_halo_sync* sync, void* stream) {
    IGG_REQUIRE(sync && sync->my_bank, "igg_wait_epoch: bad arguments");
    wait_epoch_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(*sync);
    IGG_LAUNCH_CHECK();
    return 0;
}

int igg_wait_value(uint64_t* my_flag, uint64_t value, uint64_t* err_word, void* stream) {
    IGG_REQUIRE(my_flag, "igg_wait_value: bad arguments");
    wait_value_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(my_flag, value, err_word);
    IGG_LAUNCH_CHECK();
    return 0;
}

int igg_peer_copy(void* dst, const void* src, size_t bytes, uint64_t* peer_flag, uint64_t flag_value,
                  uint64_t* my_count, void* stream) {
    IGG_REQUIRE(dst && src && bytes > 0, "igg_peer_copy: bad arguments");
    IGG_REQUIRE((((uintptr_t)dst | (uintptr_t)src) & 15) == 0, "igg_peer_copy: pointers must be 16-byte aligned");
    IGG_REQUIRE(!peer_flag || my_count, "igg_peer_copy: a flag needs a counter");
    const size_t n16 = bytes / 16;
    const int tail = (int)(bytes % 16);
    size_t blocks = (n16 + (size_t)PC_THREADS * 4 - 1) / ((size_t)PC_THREADS * 4);
    const size_t cap = (size_t)igg::sm_count() * 2;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    peer_copy_kernel<<<(int)blocks, PC_THREADS, 0, (cudaStream_t)stream>>>(
        (uint4*)dst, (const uint4*)src, n16, (unsigned char*)dst + n16 * 16, (const unsigned char*)src + n16 * 16, tail,
        peer_flag, flag_value, my_count);
    IGG_LAUNCH_CHECK();
    return 0;
}

}  // extern "C"
