// Small device-side utilities: initial condition, plane pack/unpack, interior extraction,
// max/sum reduction, colour mapping.
#include <mutex>

#include "common.cuh"

namespace {
using namespace igg;

// ---- initial condition (reference scripts/diffusion_2D_ap.jl:28: host comprehension + upload) ---------
__global__ void gaussian_ic2d_kernel(double* __restrict__ T, int nx, int ny, double x0, double y0, double dx,
                                     double dy, double lx, double ly) {
    const int iy = blockIdx.x * blockDim.x + threadIdx.x;
    const int ix = blockIdx.y * blockDim.y + threadIdx.y;
    if (ix < nx && iy < ny) {
        const double x = (x0 + ix * dx) + dx / 2 - lx / 2;
        const double y = (y0 + iy * dy) + dy / 2 - ly / 2;
        T[(size_t)ix * ny + iy] = exp(-(x * x) - (y * y));
    }
}

__global__ void gaussian_ic3d_kernel(double* __restrict__ T, int nx, int ny, int nz, double x0, double y0,
                                     double z0, double dx, double dy, double dz, double lx, double ly, double lz) {
    const int iz = blockIdx.x * blockDim.x + threadIdx.x;
    const int iy = blockIdx.y * blockDim.y + threadIdx.y;
    const int ix = blockIdx.z;
    if (ix < nx && iy < ny && iz < nz) {
        const double x = (x0 + ix * dx) + dx / 2 - lx / 2;
        const double y = (y0 + iy * dy) + dy / 2 - ly / 2;
        const double z = (z0 + iz * dz) + dz / 2 - lz / 2;
        T[((size_t)ix * ny + iy) * nz + iz] = exp(-(x * x) - (y * y) - (z * z));
    }
}

// ---- plane pack / unpack ----------------------------------------------------------------------------
// plane of dimension `dim` at `idx`: elements (a, b) over the two remaining dimensions, enumerated in
// row-major order of the remaining dimensions.
__device__ __forceinline__ size_t plane_elem(int n0, int n1, int n2, int dim, int idx, size_t e) {
    if (dim == 0) return (size_t)idx * n1 * n2 + e;                       // contiguous block
    if (dim == 1) { const size_t a = e / n2, c = e % n2; return (a * n1 + idx) * n2 + c; }
    return e * n2 + idx;                                                  // dim == 2: stride n2
}

__global__ void pack_plane_kernel(const double* __restrict__ A, int n0, int n1, int n2, int dim, int idx,
                                  double* __restrict__ buf, size_t count) {
    for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < count; e += (size_t)gridDim.x * blockDim.x)
        buf[e] = A[plane_elem(n0, n1, n2, dim, idx, e)];
}

__global__ void unpack_plane_kernel(double* __restrict__ A, int n0, int n1, int n2, int dim, int idx,
                                    const double* __restrict__ buf, size_t count) {
    for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < count; e += (size_t)gridDim.x * blockDim.x)
        A[plane_elem(n0, n1, n2, dim, idx, e)] = buf[e];
}

// Halo-free interior of A (n0 x n1 x n2, last index contiguous) copied row by row into a block of a larger array G
// (row pitches G1*G2 and G2, block origin g0, g1, g2): the dense buffer of extract_inner when G is the interior
// itself, or -- with G peer-mapped -- this rank's tile of the gathered field in the ROOT's memory (gather without
// host staging).  A CTA takes whole rows (one division per row, none per element); a thread moves two doubles per
// unit: two 8-byte loads (the interior starts one element into the row, so the source is only 8-byte aligned) and one
// 16-byte store when the destination is 16-byte aligned -- the width NVLink stores want -- else two 8-byte stores.
constexpr int CIR_THREADS = 256;
constexpr int CIR_SEG_UNITS = 4 * CIR_THREADS;        // one work item = 1024 two-element units (16 KB) of one row
__global__ void __launch_bounds__(CIR_THREADS) copy_inner_rows_kernel(const double* __restrict__ A, int n1, int n2,
                                                                      int m1, int m2, int o0, int o1, int o2,
                                                                      double* __restrict__ G, long long G1, long long G2,
                                                                      long long g0, long long g1, long long g2,
                                                                      long long rows) {
    // Work items are (row, segment) pairs with the segment running fastest, handed to the CTAs round-robin: at any
    // moment the grid works on a window of neighbouring 16 KB segments, like a plain copy does (one whole row per
    // CTA scattered 1184 concurrent streams over 100 MB and reached 54 % of the copy bandwidth).
    const int upr = m2 >> 1;                          // two-element units per row (+ one tail element if m2 is odd)
    const int segs = upr > 0 ? (upr + CIR_SEG_UNITS - 1) / CIR_SEG_UNITS : 1;
    const long long items = rows * segs;
    for (long long item = blockIdx.x; item < items; item += gridDim.x) {
        const long long row = item / segs;
        const int seg = (int)(item - row * segs);
        const long long a = row / m1, b = row - a * m1;
        const double* __restrict__ src = A + ((a + o0) * (long long)n1 + (b + o1)) * (long long)n2 + o2;
        double* __restrict__ dst = G + ((a + g0) * G1 + (b + g1)) * G2 + g2;
        const bool al16 = (((uintptr_t)dst) & 15) == 0;
        const int u0 = seg * CIR_SEG_UNITS + threadIdx.x;
        double x[4], y[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {                 // four units in flight per thread
            const int u = u0 + q * CIR_THREADS;
            if (u < upr) {
                x[q] = __ldcs(src + 2 * u);
                y[q] = __ldcs(src + 2 * u + 1);
            }
        }
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int u = u0 + q * CIR_THREADS;
            if (u < upr) {
                double* d = dst + 2 * u;
                if (al16) {
                    __stcs(reinterpret_cast<double2*>(d), make_double2(x[q], y[q]));
                } else {
                    __stcs(d, x[q]);
                    __stcs(d + 1, y[q]);
                }
            }
        }
        if ((m2 & 1) && seg == segs - 1 && threadIdx.x == 0) dst[m2 - 1] = src[m2 - 1];
    }
}

// shape with trailing singleton dimensions moved to the front, so that the LAST entry is the contiguous one
static void canonical_shape(int& n0, int& n1, int& n2) {
    for (int k = 0; k < 2 && n2 == 1; ++k) {
        n2 = n1;
        n1 = n0;
        n0 = 1;
    }
}

static int launch_copy_inner(const double* A, int n0, int n1, int n2, double* G, long long G1, long long G2,
                             long long c0, long long c1, long long c2, cudaStream_t st) {
    const int o0 = n0 > 2, o1 = n1 > 2, o2 = n2 > 2;
    const int m0 = n0 - 2 * o0, m1 = n1 - 2 * o1, m2 = n2 - 2 * o2;
    const long long rows = (long long)m0 * m1;
    const int upr = m2 >> 1;
    long long grid = rows * (upr > 0 ? (upr + CIR_SEG_UNITS - 1) / CIR_SEG_UNITS : 1);
    const long long cap = (long long)igg::sm_count() * 8;
    if (grid > cap) grid = cap;
    if (grid < 1) grid = 1;
    copy_inner_rows_kernel<<<(int)grid, CIR_THREADS, 0, st>>>(A, n1, n2, m1, m2, o0, o1, o2, G, G1, G2, c0 * m0, c1 * m1,
                                                           c2 * m2, rows);
    return 0;
}

// ---- max / sum reduction ----------------------------------------------------------------------------
// Two-level: every block writes its partial {max, sum} to scratch and takes a ticket; the block that
// draws the last ticket folds the partials in a fixed order (deterministic result) and resets the
// ticket for the next call.
constexpr int RED_THREADS = 256;
constexpr int RED_MAX_BLOCKS = 1024;

struct RedScratch {
    unsigned int ticket;
    unsigned int pad[3];
    double pmax[RED_MAX_BLOCKS];
    double psum[RED_MAX_BLOCKS];
};

__device__ __forceinline__ void block_reduce(double& mx, double& sm, double* sh_max, double* sh_sum) {
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        mx = igg::max_nan(mx, __shfl_xor_sync(0xffffffffu, mx, off));
        sm += __shfl_xor_sync(0xffffffffu, sm, off);
    }
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    if (l == 0) { sh_max[w] = mx; sh_sum[w] = sm; }
    __syncthreads();
    if (w == 0) {
        mx = l < RED_THREADS / 32 ? sh_max[l] : -INFINITY;
        sm = l < RED_THREADS / 32 ? sh_sum[l] : 0.0;
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            mx = igg::max_nan(mx, __shfl_xor_sync(0xffffffffu, mx, off));
            sm += __shfl_xor_sync(0xffffffffu, sm, off);
        }
    }
    __syncthreads();
}

__global__ void __launch_bounds__(RED_THREADS) reduce_max_sum_kernel(const double* __restrict__ x, size_t n,
                                                                     RedScratch* sc, double* __restrict__ out) {
    __shared__ double sh_max[RED_THREADS / 32], sh_sum[RED_THREADS / 32];
    __shared__ bool is_last;
    double mx = -INFINITY, sm = 0.0;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if ((((uintptr_t)x) & 15) == 0) {
        // 16-byte loads, four per thread in flight (a scalar 8-byte loop reached 31 % of the DRAM bandwidth)
        const double2* __restrict__ x2 = reinterpret_cast<const double2*>(x);
        const size_t n2 = n >> 1;
        for (; e + 3 * stride < n2; e += 4 * stride) {
            const double2 a = x2[e], b = x2[e + stride], c = x2[e + 2 * stride], d = x2[e + 3 * stride];
            mx = igg::max_nan(mx, igg::max_nan(igg::max_nan(a.x, a.y), igg::max_nan(b.x, b.y)));
            mx = igg::max_nan(mx, igg::max_nan(igg::max_nan(c.x, c.y), igg::max_nan(d.x, d.y)));
            sm += ((a.x + a.y) + (b.x + b.y)) + ((c.x + c.y) + (d.x + d.y));
        }
        for (; e < n2; e += stride) {
            const double2 a = x2[e];
            mx = igg::max_nan(mx, igg::max_nan(a.x, a.y));
            sm += a.x + a.y;
        }
        if ((n & 1) && blockIdx.x == 0 && threadIdx.x == 0) {
            mx = igg::max_nan(mx, x[n - 1]);
            sm += x[n - 1];
        }
    } else {
        for (; e < n; e += stride) {
            const double v = x[e];
            mx = igg::max_nan(mx, v);
            sm += v;
        }
    }
    block_reduce(mx, sm, sh_max, sh_sum);
    if (threadIdx.x == 0) {
        sc->pmax[blockIdx.x] = mx;
        sc->psum[blockIdx.x] = sm;
        __threadfence();
        const unsigned int t = atomicAdd(&sc->ticket, 1u);
        is_last = (t == gridDim.x - 1);
    }
    __syncthreads();
    if (is_last) {
        __threadfence();
        mx = -INFINITY;
        sm = 0.0;
        for (unsigned int b = threadIdx.x; b < gridDim.x; b += blockDim.x) {
            mx = igg::max_nan(mx, sc->pmax[b]);
            sm += sc->psum[b];
        }
        block_reduce(mx, sm, sh_max, sh_sum);
        if (threadIdx.x == 0) {
            out[0] = mx;
            out[1] = sm;
            sc->ticket = 0;   // ready for the next call on this scratch
        }
    }
}

// ---- colour map (reference scripts/diffusion_2D_ap.jl:47 heatmap rasterisation, K15) -----------------
__global__ void colormap_kernel(const double* __restrict__ f, size_t n, double vmin, double scale,
                                const uint8_t* __restrict__ lut, uint8_t* __restrict__ rgb) {
    __shared__ uint8_t s_lut[768];
    for (int k = threadIdx.x; k < 768; k += blockDim.x) s_lut[k] = lut[k];
    __syncthreads();
    for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += (size_t)gridDim.x * blockDim.x) {
        double t = floor((f[e] - vmin) * scale + 0.5);
        t = t < 0.0 ? 0.0 : (t > 255.0 ? 255.0 : t);
        const int k = (int)t * 3;
        rgb[3 * e + 0] = s_lut[k + 0];
        rgb[3 * e + 1] = s_lut[k + 1];
        rgb[3 * e + 2] = s_lut[k + 2];
    }
}

static int grid_for(size_t count, int threads, int cap_mult = 8) {
    size_t b = (count + threads - 1) / threads;
    const size_t cap = (size_t)igg::sm_count() * cap_mult;
    if (b > cap) b = cap;
    if (b < 1) b = 1;
    return (int)b;
}

}  // namespace

extern "C" {

int igg_gaussian_ic2d(double* T, int nx, int ny, double x0, double y0, double dx, double dy, double lx, double ly,
                      void* stream) {
    IGG_REQUIRE(T && nx > 0 && ny > 0, "igg_gaussian_ic2d: bad arguments");
    dim3 blk(32, 8), grd(igg::ceil_div(ny, 32), igg::ceil_div(nx, 8));
    IGG_REQUIRE(grd.y <= 65535, "igg_gaussian_ic2d: nx too large for this launch shape");
    gaussian_ic2d_kernel<<<grd, blk, 0, (cudaStream_t)stream>>>(T, nx, ny, x0, y0, dx, dy, lx, ly);
    IGG_LAUNCH_CHECK();
    return 0;
}

int igg_gaussian_ic3d(double* T, int nx, int ny, int nz, double x0, double y0, double z0, double dx, double dy,
                      double dz, double lx, double ly, double lz, void* stream) {
    IGG_REQUIRE(T && nx > 0 && ny > 0 && nz > 0, "igg_gaussian_ic3d: bad arguments");
    dim3 blk(32, 8), grd(igg::ceil_div(nz, 32), igg::ceil_div(ny, 8), nx);
    IGG_REQUIRE(grd.y <= 65535 && grd.z <= 65535, "igg_gaussian_ic3d: grid too large");
    gaussian_ic3d_kernel<<<grd, blk, 0, (cudaStream_t)stream>>>(T, nx, ny, nz, x0, y0, z0, dx, dy, dz, lx, ly, lz);
    IGG_LAUNCH_CHECK();
    return 0;
}

int igg_pack_plane(const double* A, int n0, int n1, int n2, int dim, int idx, double* buf, void* stream) {
    IGG_REQUIRE(A && buf && n0 > 0 && n1 > 0 && n2 > 0 && dim >= 0 && dim < 3, "igg_pack_plane: bad arguments");
    const int nd[3] = {n0, n1, n2};
    IGG_REQUIRE(idx >= 0 && idx < nd[dim], "igg_pack_plane: plane index out of range");
    const size_t count = (size_t)n0 * n1 * n2 / nd[dim];
    pack_plane_kernel<<<grid_for(count, 256), 256, 0, (cudaStream_t)stream>>>(A, n0, n1, n2, dim, idx, buf, count);
    IGG_LAUNCH_CHECK();
    return 0;
}

int igg_unpack_plane(double* A, int n0, int n1, int n2, int dim, int idx, const double* buf, void* stream) {
    IGG_REQUIRE(A && buf && n0 > 0 && n1 > 0 && n2 > 0 && dim >= 0 && dim < 3, "igg_unpack_plane: bad arguments");
    const int nd[3] = {n0, n1, n2};
    IGG_REQUIRE(idx >= 0 && idx < nd[dim], "igg_unpack_plane: plane index out of range");
    const size_t count = (size_t)n0 * n1 * n2 / nd[dim];
    unpack_plane_kernel<<<grid_for(count, 256), 256, 0, (cudaStream_t)stream>>>(A, n0, n1, n2, dim, idx, buf, count);
    IGG_LAUNCH_CHECK();
    return 0;
}

int igg_extract_inner(const double* A, int n0, int n1, int n2, double* out, void* stream) {
    IGG_REQUIRE(A && out && n0 > 0 && n1 > 0 && n2 > 0, "igg_extract_inner: bad arguments");
    canonical_shape(n0, n1, n2);
    const int m1 = n1 - 2 * (n1 > 2), m2 = n2 - 2 * (n2 > 2);
    launch_copy_inner(A, n0, n1, n2, out, m1, m2, 0, 0, 0, (cudaStream_t)stream);
    IGG_LAUNCH_CHECK();
    return 0;
}

int igg_gather_inner(const double* A, int n0, int n1, int n2, double* G, int d0, int d1, int d2, int c0, int c1,
                     int c2, void* stream) {
    IGG_REQUIRE(A && G && n0 > 0 && n1 > 0 && n2 > 0, "igg_gather_inner: bad arguments");
    IGG_REQUIRE(d0 >= 1 && d1 >= 1 && d2 >= 1 && c0 >= 0 && c0 < d0 && c1 >= 0 && c1 < d1 && c2 >= 0 && c2 < d2,
                "igg_gather_inner: block coordinates (%d,%d,%d) outside the process grid %d x %d x %d", c0, c1, c2, d0,
                d1, d2);
    // singleton dimensions travel with their process-grid entries
    for (int k = 0; k < 2 && n2 == 1; ++k) {
        IGG_REQUIRE(d2 == 1, "igg_gather_inner: a dimension of size 1 cannot be decomposed");
        n2 = n1; n1 = n0; n0 = 1;
        d2 = d1; d1 = d0; d0 = 1;
        c2 = c1; c1 = c0; c0 = 0;
    }
    const long long m1 = n1 - 2 * (n1 > 2), m2 = n2 - 2 * (n2 > 2);
    launch_copy_inner(A, n0, n1, n2, G, m1 * d1, m2 * d2, c0, c1, c2, (cudaStream_t)stream);
    IGG_LAUNCH_CHECK();
    return 0;
}

size_t igg_reduce_scratch_bytes(size_t n) {
    (void)n;
    return sizeof(RedScratch);
}

int igg_reduce_max_sum(const double* x, size_t n, void* scratch, double* out, void* stream) {
    IGG_REQUIRE(x && scratch && out && n > 0, "igg_reduce_max_sum: bad arguments");
    IGG_REQUIRE(((uintptr_t)scratch & 15) == 0, "igg_reduce_max_sum: scratch must be 16-byte aligned");
    int blocks = grid_for(n, RED_THREADS, 4);
    if (blocks > RED_MAX_BLOCKS) blocks = RED_MAX_BLOCKS;
    reduce_max_sum_kernel<<<blocks, RED_THREADS, 0, (cudaStream_t)stream>>>(x, n, (RedScratch*)scratch, out);
    IGG_LAUNCH_CHECK();
    return 0;
}

int igg_colormap(const double* field, size_t n, double vmin, double vmax, const uint8_t* lut_dev, uint8_t* rgb,
                 void* stream) {
    IGG_REQUIRE(field && lut_dev && rgb && n > 0, "igg_colormap: bad arguments");
    const double scale = vmax > vmin ? 255.0 / (vmax - vmin) : 0.0;
    colormap_kernel<<<grid_for(n, 256), 256, 0, (cudaStream_t)stream>>>(field, n, vmin, scale, lut_dev, rgb);
    IGG_LAUNCH_CHECK();
    return 0;
}

}  // extern "C"

// =====================================================================================================
// igg_reduce_max_sum_static -- the same max/sum reduction for callers that must not allocate.
//
// Who this is for: code that wants the diagnostic (`max(T)`, `sum(T)`) but has no device memory to give
// and is not allowed to get any -- a monitoring hook that runs after the application has claimed the
// whole GPU for its fields, a host program whose memory policy reserves all allocation for its own
// pool, or a time loop that cannot afford the device-wide stall a cudaMalloc/cudaFree causes.  For
// them the scratch of igg_reduce_max_sum (block partials + ticket) lives in the array below, inside
// the library image: every GPU that loads the library carries these 16,640 bytes until the process
// ends, whether or not this entry point is ever called.
//
// Contract
//   * n may not exceed IGG_REDUCE_STATIC_MAX_N (2^29 values = 4 GiB of fp64, twice the 16384^2
//     flagship field).  The array has room for 1024 block partials and the launch uses at most that
//     many blocks of 256 threads, so the limit keeps the values folded by one thread -- and with them
//     the run time and the rounding-error growth of a partial sum -- to a few thousand.  Larger n is
//     refused with an error message; reduce it in pieces or use igg_reduce_max_sum with your own scratch.
//   * The first call on each device resolves the device address of the array (cudaGetSymbolAddress,
//     with that device current).  It is host-side set-up work, so make that first call during
//     start-up, before any stream capture begins.
//   * No cudaMalloc, cudaFree or other allocating routine is called, here or in the kernel.
//   * ALL calls on one device share the one array and it keeps whatever the previous call left in it.
//     The only part that has to start from zero, the ticket, is cleared on `stream` ahead of every
//     launch; the partials are rewritten before they are read.  Because of the sharing, two calls
//     whose kernels can overlap on the GPU (different streams, or different host threads without
//     ordering between them) race on the partials and the ticket: keep calls on one device ordered.
//   * Inputs and outputs are the caller's, exactly as for igg_reduce_max_sum.
// =====================================================================================================
namespace {
constexpr size_t STATIC_SCRATCH_BYTES = (sizeof(RedScratch) + 255) / 256 * 256;   // 16,640
static_assert(STATIC_SCRATCH_BYTES == 16640, "stated size of the built-in scratch");
}  // namespace

static __device__ __align__(256) unsigned char g_reduce_static_scratch[STATIC_SCRATCH_BYTES];

namespace {
// Device address of g_reduce_static_scratch per device.  In host code the symbol's own name is only a
// handle for the runtime, never something a kernel may dereference -- hence cudaGetSymbolAddress.
std::mutex g_static_mu;
void* g_static_ptr[64] = {nullptr};

int static_scratch_for_current_device(void** out) {
    int dev = -1;
    IGG_CUDA_OK(cudaGetDevice(&dev));
    IGG_REQUIRE(dev >= 0 && dev < 64, "igg_reduce_max_sum_static: device ordinal %d is out of range", dev);
    std::lock_guard<std::mutex> lk(g_static_mu);
    if (g_static_ptr[dev] == nullptr) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress((void**)&p, g_reduce_static_scratch);
        if (e != cudaSuccess || p == nullptr) {
            igg::set_error("igg_reduce_max_sum_static: cudaGetSymbolAddress failed on device %d: %s", dev,
                           cudaGetErrorString(e));
            return e != cudaSuccess ? (int)e : IGG_ERR_INVALID;
        }
        g_static_ptr[dev] = p;
    }
    *out = g_static_ptr[dev];
    return 0;
}
}  // namespace

extern "C" int igg_reduce_max_sum_static(const double* x, size_t n, double* out, void* stream) {
    IGG_REQUIRE(x && out && n > 0, "igg_reduce_max_sum_static: bad arguments");
    if (n > (size_t)IGG_REDUCE_STATIC_MAX_N) {
        igg::set_error("igg_reduce_max_sum_static: n = %zu exceeds the built-in scratch limit of %zu values "
                       "(IGG_REDUCE_STATIC_MAX_N); split the input or call igg_reduce_max_sum with your own scratch",
                       n, (size_t)IGG_REDUCE_STATIC_MAX_N);
        return IGG_ERR_UNSUPPORTED;
    }
    void* scratch = nullptr;
    int rc = static_scratch_for_current_device(&scratch);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    // the array still holds the previous call's state: the ticket is the one field that must be zero
    IGG_CUDA_OK(cudaMemsetAsync(scratch, 0, 16, st));
    int blocks = grid_for(n, RED_THREADS, 4);
    if (blocks > RED_MAX_BLOCKS) blocks = RED_MAX_BLOCKS;
    reduce_max_sum_kernel<<<blocks, RED_THREADS, 0, st>>>(x, n, (RedScratch*)scratch, out);
    IGG_LAUNCH_CHECK();
    return 0;
}
