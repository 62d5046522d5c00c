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
// All boxes of all fields are flattened into one list of "units" that the CTAs share evenly:
//   * a box whose rows are long contiguous runs with equal 16-byte phase on both sides (y-planes of a 2-D field,
//     faces of a 3-D field that keep the last dimension) moves 16-byte vectors: unit 0 of a row is the scalar
//     head + tail, units 1.. are the vectors;
//   * every other box (x-planes of a row-major field: one element per row, strided) moves single elements.
// A thread keeps four units in flight (independent loads, then the four stores).  The strided local loads do
// not need the neighbours' permission, so the first batch is fetched BEFORE the ready handshake is awaited:
// the handshake's NVLink round trip overlaps the load latency instead of preceding it.
constexpr int HX_THREADS = 256;
constexpr int HX_MAX_BOXES = 27 * IGG_MAX_HALO_FIELDS;
constexpr int HX_BATCH = 4;

struct HxBox {
    const unsigned char* src;       // address of the box's first element in my array
    unsigned char* dst;             // address of its destination in the neighbour's array
    unsigned long long pitch_a, pitch_b;   // bytes between consecutive a / b indices (same on both sides)
    unsigned int e1;                // extent[1]
    unsigned int upr;               // units per row
    unsigned int run;               // elements per row
    unsigned int eb;                // element bytes
    unsigned int head;              // scalar elements before the first 16-byte vector (vector boxes)
    unsigned int nvec;              // 16-byte vectors per row (0 = scalar box)
};

__device__ __forceinline__ void hx_setup_box(const igg_halo_xfer& t, HxBox& b, unsigned long long& units) {
    const unsigned long long eb = (unsigned long long)t.elem_bytes;
    b.eb = (unsigned int)eb;
    b.pitch_b = (unsigned long long)t.n[2] * eb;
    b.pitch_a = (unsigned long long)t.n[1] * b.pitch_b;
    const unsigned long long so = ((unsigned long long)t.src_start[0] * t.n[1] + t.src_start[1]) * t.n[2] + t.src_start[2];
    const unsigned long long dof = ((unsigned long long)t.dst_start[0] * t.n[1] + t.dst_start[1]) * t.n[2] + t.dst_start[2];
    b.src = (const unsigned char*)t.src + so * eb;
    b.dst = (unsigned char*)t.dst + dof * eb;
    b.e1 = (unsigned int)t.extent[1];
    b.run = (unsigned int)t.extent[2];
    const unsigned long long rows = (unsigned long long)t.extent[0] * t.extent[1];
    const unsigned long long run_bytes = (unsigned long long)b.run * eb;
    const unsigned int ps = (unsigned int)((uintptr_t)b.src & 15), pd = (unsigned int)((uintptr_t)b.dst & 15);
    const bool vec = run_bytes >= 64 && ps == pd && (b.pitch_b & 15) == 0 && (ps % eb) == 0;
    if (vec) {
        const unsigned int head_bytes = (16 - ps) & 15;
        b.head = head_bytes / (unsigned int)eb;
        b.nvec = (unsigned int)((run_bytes - head_bytes) / 16);
        b.upr = 1 + b.nvec;
    } else {
        b.head = 0;
        b.nvec = 0;
        b.upr = b.run;
    }
    units = rows * b.upr;
}

__device__ __forceinline__ void hx_copy_elem(unsigned char* d, const unsigned char* s, unsigned int eb) {
    switch (eb) {
        case 8: *(uint64_t*)d = *(const uint64_t*)s; break;
        case 4: *(uint32_t*)d = *(const uint32_t*)s; break;
        case 2: *(uint16_t*)d = *(const uint16_t*)s; break;
        default: *d = *s; break;
    }
}

__global__ void __launch_bounds__(HX_THREADS) halo_exchange_kernel(const igg_halo_xfer* __restrict__ xf, int nx,
                                                                   const __grid_constant__ igg_halo_sync sy) {
    __shared__ HxBox box[HX_MAX_BOXES];
    __shared__ unsigned long long first_unit[HX_MAX_BOXES + 1];
    __shared__ int abort_flag;
    const int tid = threadIdx.x;
    uint64_t* err = sy.my_bank + IGG_SLOT_ERR;
    // (1) tell every neighbour that my halo cells may be overwritten (all earlier work on this stream has
    //     finished).  The promise I make to the neighbour at offset o shows up in ITS bank under 26-o.
    if (blockIdx.x == 0 && tid < 27 && sy.peer_bank[tid])
        st_release_sys(sy.peer_bank[tid] + IGG_SLOT_READY + (26 - tid), sy.epoch);
    if (tid == 0) abort_flag = 0;
    for (int k = tid; k < nx; k += HX_THREADS) {
        unsigned long long u;
        hx_setup_box(xf[k], box[k], u);
        first_unit[k + 1] = u;
    }
    __syncthreads();
    if (tid == 0) {
        first_unit[0] = 0;
        for (int k = 0; k < nx; ++k) first_unit[k + 1] += first_unit[k];
    }
    __syncthreads();
    const unsigned long long total = first_unit[nx];
    const unsigned long long per_cta = (total + gridDim.x - 1) / gridDim.x;
    const unsigned long long u_begin = (unsigned long long)blockIdx.x * per_cta;
    const unsigned long long u_end = u_begin + per_cta < total ? u_begin + per_cta : total;

    // unit -> addresses.  kind: 0 = nothing, 1 = one element, 2 = 16-byte vector, 3 = head + tail of a row
    auto locate = [&](unsigned long long u, const unsigned char*& s, unsigned char*& d, unsigned int& eb,
                      unsigned int& aux) -> int {
        if (u >= u_end) return 0;
        int lo = 0, hi = nx - 1;
        while (lo < hi) {
            const int mid = (lo + hi + 1) >> 1;
            if (first_unit[mid] <= u) lo = mid; else hi = mid - 1;
        }
        const HxBox& b = box[lo];
        const unsigned int ul = (unsigned int)(u - first_unit[lo]);
        const unsigned int row = ul / b.upr, cu = ul - row * b.upr;
        const unsigned int a = row / b.e1, bb = row - a * b.e1;
        const unsigned long long off = (unsigned long long)a * b.pitch_a + (unsigned long long)bb * b.pitch_b;
        s = b.src + off;
        d = b.dst + off;
        eb = b.eb;
        if (b.nvec == 0) {
            s += (size_t)cu * eb;
            d += (size_t)cu * eb;
            return 1;
        }
        if (cu == 0) {
            aux = (unsigned int)lo;
            return 3;
        }
        const size_t vo = (size_t)b.head * eb + (size_t)(cu - 1) * 16;
        s += vo;
        d += vo;
        return 2;
    };

    bool ready_checked = false;
    for (unsigned long long cbase = u_begin; cbase < u_end; cbase += (unsigned long long)HX_BATCH * HX_THREADS) {
        // the loop bound is CTA-uniform, so the barrier below is reached by every thread
        const unsigned char* s[HX_BATCH];
        unsigned char* d[HX_BATCH];
        unsigned int eb[HX_BATCH], aux[HX_BATCH];
        int kind[HX_BATCH];
        uint4 v[HX_BATCH];
#pragma unroll
        for (int q = 0; q < HX_BATCH; ++q) {
            kind[q] = locate(cbase + (unsigned long long)q * HX_THREADS + tid, s[q], d[q], eb[q], aux[q]);
            if (kind[q] == 2) {
                v[q] = *(const uint4*)s[q];
            } else if (kind[q] == 1) {
                switch (eb[q]) {
                    case 8: { const uint64_t w = *(const uint64_t*)s[q]; v[q].x = (uint32_t)w; v[q].y = (uint32_t)(w >> 32); } break;
                    case 4: v[q].x = *(const uint32_t*)s[q]; break;
                    case 2: v[q].x = *(const uint16_t*)s[q]; break;
                    default: v[q].x = *s[q]; break;
                }
            }
        }
        if (!ready_checked) {
            // (2) the neighbours' promise, awaited with the first batch of loads already in flight
            if (tid < 27 && sy.peer_bank[tid] && !wait_flag_ge(sy.my_bank + IGG_SLOT_READY + tid, sy.epoch, err))
                abort_flag = 1;
            __syncthreads();
            ready_checked = true;
        }
        if (abort_flag) break;      // a neighbour never became ready: store nothing (the error word is set)
#pragma unroll
        for (int q = 0; q < HX_BATCH; ++q) {
            if (kind[q] == 2) {
                *(uint4*)d[q] = v[q];
            } else if (kind[q] == 1) {
                switch (eb[q]) {
                    case 8: *(uint64_t*)d[q] = (uint64_t)v[q].x | ((uint64_t)v[q].y << 32); break;
                    case 4: *(uint32_t*)d[q] = v[q].x; break;
                    case 2: *(uint16_t*)d[q] = (uint16_t)v[q].x; break;
                    default: *d[q] = (unsigned char)v[q].x; break;
                }
            } else if (kind[q] == 3) {
                const HxBox& b = box[aux[q]];
                for (unsigned int e = 0; e < b.head; ++e) hx_copy_elem(d[q] + (size_t)e * b.eb, s[q] + (size_t)e * b.eb, b.eb);
                const unsigned int done = b.head + b.nvec * (16u / b.eb);
                for (unsigned int e = done; e < b.run; ++e) hx_copy_elem(d[q] + (size_t)e * b.eb, s[q] + (size_t)e * b.eb, b.eb);
            }
        }
    }
    if (!ready_checked) {           // CTA without units (more CTAs than work): still part of the protocol
        if (tid < 27 && sy.peer_bank[tid] && !wait_flag_ge(sy.my_bank + IGG_SLOT_READY + tid, sy.epoch, err))
            abort_flag = 1;
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
                                                               uint64_t* peer_flag, uint64_t value, uint64_t* my_count,
                                                               const uint64_t* my_flag, uint64_t* err_word) {
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
    // send + receive in ONE launch: the launch is complete when the matching message has landed here too
    if (my_flag && blockIdx.x == 0 && threadIdx.x == 0) wait_flag_ge(my_flag, value, err_word);
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
    // The handle names the whole allocation; the tensor may start anywhere inside it (torch's caching
    // allocator sub-allocates).  Without the allocation base the peers would be pointed at the wrong bytes,
    // so failing to obtain it is an error, never "offset 0".
    *offset = 0;
    GetAddressRangeFn fn = address_range_fn();
    CUdeviceptr base = 0;
    size_t size = 0;
    if (!fn || fn(&base, &size, (CUdeviceptr)(uintptr_t)ptr) != CUDA_SUCCESS) {
        igg::set_error("igg_ipc_export: cannot resolve the allocation base of %p (cuMemGetAddressRange %s)", ptr,
                       fn ? "failed" : "is unavailable");
        return IGG_ERR_UNSUPPORTED;
    }
    *offset = (uint64_t)((uintptr_t)ptr - (uintptr_t)base);
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

int igg_halo_exchange_peer_ex(const igg_halo_xfer* xfers_dev, int nxfers, const igg_halo_sync* sync, int ctas,
                              void* stream) {
    IGG_REQUIRE(sync && sync->my_bank && sync->my_count_bank && nxfers >= 0 && (nxfers == 0 || xfers_dev),
                "igg_halo_exchange_peer: bad arguments");
    IGG_REQUIRE(sync->epoch >= 1, "igg_halo_exchange_peer: epoch must be >= 1");
    IGG_REQUIRE(nxfers <= HX_MAX_BOXES, "igg_halo_exchange_peer: at most %d boxes per call (got %d)", HX_MAX_BOXES,
                nxfers);
    // `ctas` = what the caller derived from the bytes it moves (one CTA per 16 KB is plenty); 0 = a quarter
    // of the SMs, which saturates NVLink for plane-sized messages
    int grid = ctas > 0 ? ctas : igg::sm_count() / 4;
    const int cap = igg::sm_count() * 2;         // all CTAs wait on flags: they must be co-resident
    if (grid > cap) grid = cap;
    if (grid < 1) grid = 1;
    halo_exchange_kernel<<<grid, HX_THREADS, 0, (cudaStream_t)stream>>>(xfers_dev, nxfers, *sync);
    IGG_LAUNCH_CHECK();
    return 0;
}

int igg_halo_exchange_peer(const igg_halo_xfer* xfers_dev, int nxfers, const igg_halo_sync* sync, void* stream) {
    return igg_halo_exchange_peer_ex(xfers_dev, nxfers, sync, 0, stream);
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
        peer_flag, flag_value, my_count, nullptr, nullptr);
    IGG_LAUNCH_CHECK();
    return 0;
}

int igg_peer_sendrecv(void* dst, const void* src, size_t bytes, uint64_t* peer_flag, const uint64_t* my_flag,
                      uint64_t flag_value, uint64_t* my_count, uint64_t* err_word, void* stream) {
    IGG_REQUIRE(dst && src && bytes > 0 && peer_flag && my_flag && my_count, "igg_peer_sendrecv: bad arguments");
    IGG_REQUIRE((((uintptr_t)dst | (uintptr_t)src) & 15) == 0, "igg_peer_sendrecv: pointers must be 16-byte aligned");
    const size_t n16 = bytes / 16;
    const int tail = (int)(bytes % 16);
    size_t blocks = (n16 + (size_t)PC_THREADS * 4 - 1) / ((size_t)PC_THREADS * 4);
    const size_t cap = (size_t)igg::sm_count() * 2;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    peer_copy_kernel<<<(int)blocks, PC_THREADS, 0, (cudaStream_t)stream>>>(
        (uint4*)dst, (const uint4*)src, n16, (unsigned char*)dst + n16 * 16, (const unsigned char*)src + n16 * 16, tail,
        peer_flag, flag_value, my_count, my_flag, err_word);
    IGG_LAUNCH_CHECK();
    return 0;
}

}  // extern "C"
