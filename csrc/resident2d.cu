// Cluster-resident multi-step diffusion kernel for small local grids (sm_100a).
//
// No counterpart in the reference: its smallest configuration (128 x 128 cells per GPU,
// scripts/diffusion_2D_ap.jl:15, scripts/diffusion_2D_kp.jl:62) pays three launches with a host wait each plus an
// MPI halo exchange per time step.  Here the whole local field lives in the shared memory of ONE
// thread-block cluster (up to 16 CTAs) for all `nsteps` steps:
//
//  * CTA c owns a contiguous block of rows of T, T2 and Cp in its shared memory;
//  * the rows that straddle two CTAs are read from the neighbouring CTA's shared memory (distributed
//    shared memory, cluster.map_shared_rank), steps are separated by a cluster barrier (hardware,
//    sub-microsecond) instead of a grid barrier in global memory;
//  * cells on the edge of the local domain are also stored straight into the neighbouring GPUs' halo
//    cells (NVLink peer stores into their global arrays) and announced with per-CTA epoch flags; the
//    matching incoming halo cells are imported from global into shared memory once the neighbour's flag
//    has been acquired.  Row partitions are identical on all ranks, so CTA c only ever waits for CTA c of
//    its y-neighbours, CTA 0 / C-1 for the x-neighbours -- no atomic counters.
//
// Field layout T[ix][iy] (iy contiguous) in global memory; in shared memory each CTA keeps its rows with the
// same pitch ny.
#include <cooperative_groups.h>
#include <string.h>

#include "common.cuh"

namespace cg = cooperative_groups;

namespace {
using namespace igg;

constexpr int RES_THREADS = 512;
constexpr int RES_MAX_CTAS = 16;

struct ResParams {
    double* T;                  // global arrays (this rank)
    double* T2;
    const double* Cp;
    int nx, ny, R, nsteps;      // R = rows per CTA
    double lam, dt, _dx, _dy;
    double* peer_T[9];          // neighbours' global arrays (NULL = no neighbour); [ (ox+1)*3 + (oy+1) ]
    double* peer_T2[9];
    uint64_t* peer_flags[9];    // base of the neighbour's flag region
    uint64_t* my_flags;         // base of my flag region: slot [o * 16 + cta], error word at [9 * 16]
    uint64_t epoch0;            // epoch of step 0
    // low-latency mode: every pushed value travels WITH its epoch (two 8-byte words {32 data bits, 32-bit
    // epoch tag}) into a landing buffer of the neighbour, so no fence / flag round trip is needed
    uint64_t* peer_land[9];     // base of the neighbour's landing buffer (NULL = flag mode or no neighbour)
    uint64_t* my_land;          // mine: [parity][source offset][index along the edge][2 words]
    int land_len;               // entries per (parity, offset) region
    int form;                   // 0 = perf form, 1 = kp form (update_cell_kp)
};

// landing-buffer helpers ---------------------------------------------------------------------------------
__device__ __forceinline__ void ll_store(uint64_t* slot, double v, uint32_t tag) {
    const uint64_t bits = (uint64_t)__double_as_longlong(v);
    const uint64_t w0 = (bits & 0xffffffffull) | ((uint64_t)tag << 32);
    const uint64_t w1 = (bits >> 32) | ((uint64_t)tag << 32);
    asm volatile("st.volatile.global.v2.u64 [%0], {%1, %2};" ::"l"(slot), "l"(w0), "l"(w1) : "memory");
}
// spins until both words carry `tag`; returns false on timeout
__device__ __forceinline__ bool ll_load(const uint64_t* slot, uint32_t tag, double& v) {
    uint64_t w0, w1;
    unsigned spins = 0;
    while (true) {
        asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(w0), "=l"(w1) : "l"(slot) : "memory");
        if ((uint32_t)(w0 >> 32) == tag && (uint32_t)(w1 >> 32) == tag) break;
        if (++spins > (1u << 26)) return false;
    }
    v = __longlong_as_double((long long)((w0 & 0xffffffffull) | (w1 << 32)));
    return true;
}

__device__ __forceinline__ double upd(double c, double xm, double xp, double ym, double yp, double cp, double lam,
                                      double dt, double _dx, double _dy) {
    const double qxr = -lam * (xp - c) * _dx;
    const double qxl = -lam * (c - xm) * _dx;
    const double qyt = -lam * (yp - c) * _dy;
    const double qyb = -lam * (c - ym) * _dy;
    return c + dt * (cp * (-(qxr - qxl) * _dx - (qyt - qyb) * _dy));
}

__global__ void __launch_bounds__(RES_THREADS, 1) diff2d_resident_kernel(const __grid_constant__ ResParams p) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    cg::cluster_group cluster = cg::this_cluster();
    const int c = (int)cluster.block_rank();
    const int C = (int)cluster.num_blocks();
    const int tid = threadIdx.x;
    const int nx = p.nx, ny = p.ny, R = p.R;
    const int r0 = c * R;                                   // first row I own
    const int nr = max(0, min(R, nx - r0));                 // rows I own
    double* bufA = reinterpret_cast<double*>(smem_raw);     // R x ny each
    double* bufB = bufA + (size_t)R * ny;
    double* bufC = bufB + (size_t)R * ny;
    uint64_t* err = p.my_flags + 9 * 16;

    // load my rows of T, T2, Cp
    for (int e = tid; e < nr * ny; e += RES_THREADS) {
        const size_t g = (size_t)r0 * ny + e;
        bufA[e] = __ldcg(p.T + g);
        bufB[e] = __ldcg(p.T2 + g);
        bufC[e] = p.Cp[g];
    }
    cluster.sync();

    // neighbours inside the cluster (rows above / below my block)
    const bool has_up = c > 0, has_dn = (c + 1 < C) && ((c + 1) * R < nx);
    const bool own_row1 = (1 >= r0 && 1 < r0 + nr), own_rowN = (nx - 2 >= r0 && nx - 2 < r0 + nr);
    const bool own_row0 = (c == 0), own_rowL = (nx - 1 >= r0 && nx - 1 < r0 + nr);
    const bool ll = p.my_land != nullptr;                   // low-latency (value + epoch tag) exchange
    bool has_peers = false;
    bool nbr[9];                                            // is there a neighbour at offset o
#pragma unroll
    for (int o = 0; o < 9; ++o) {
        nbr[o] = ll ? p.peer_land[o] != nullptr : p.peer_flags[o] != nullptr;
        has_peers |= nbr[o];
    }

    for (int s = 0; s < p.nsteps; ++s) {
        const bool even = (s & 1) == 0;
        double* src = even ? bufA : bufB;
        double* dst = even ? bufB : bufA;
        double* const* peers = even ? p.peer_T2 : p.peer_T;
        double* gdst = even ? p.T2 : p.T;                   // my global array the neighbours push into
        const uint64_t epoch = p.epoch0 + (uint64_t)s;
        const double* up_src = has_up ? cluster.map_shared_rank(src, c - 1) + (size_t)(R - 1) * ny : nullptr;
        const double* dn_src = has_dn ? cluster.map_shared_rank(src, c + 1) : nullptr;

        // ---- compute my interior cells: 128 columns x 4 row lanes, no index divisions ------------------
        const int tx = tid & 127, ty = tid >> 7;
        for (int y = 1 + tx; y <= ny - 2; y += 128) {
#pragma unroll 2
            for (int lr = ty; lr < nr; lr += RES_THREADS / 128) {
                const int x = r0 + lr;
                if (x < 1 || x > nx - 2) continue;
                const double* row = src + (size_t)lr * ny;
                const double cc = row[y];
                const double xm = lr > 0 ? row[y - ny] : up_src[y];
                const double xp = lr + 1 < nr ? row[y + ny] : dn_src[y];
                const double v = p.form ? update_cell_kp(cc, xm, xp, row[y - 1], row[y + 1], bufC[(size_t)lr * ny + y],
                                                         p.lam, p.dt, p._dx, p._dy)
                                        : upd(cc, xm, xp, row[y - 1], row[y + 1], bufC[(size_t)lr * ny + y], p.lam, p.dt,
                                              p._dx, p._dy);
                dst[(size_t)lr * ny + y] = v;
                // edge cells go to the neighbouring GPUs' halo cells
                const bool xl = x == 1, xh = x == nx - 2, yl = y == 1, yh = y == ny - 2;
                if (has_peers && (xl | xh | yl | yh)) {
#pragma unroll
                    for (int sx = -1; sx <= 1; ++sx) {
                        if ((sx == -1 && !xl) || (sx == 1 && !xh)) continue;
#pragma unroll
                        for (int sy = -1; sy <= 1; ++sy) {
                            if ((sy == -1 && !yl) || (sy == 1 && !yh) || (sx == 0 && sy == 0)) continue;
                            const int k = (sx + 1) * 3 + (sy + 1);
                            if (ll) {
                                // value + tag into the neighbour's landing region for source offset 8-k;
                                // index along the edge: y for x-edges, x for y-edges, 0 for corners
                                uint64_t* land = p.peer_land[k];
                                if (!land) continue;
                                const int idx = sx != 0 ? (sy == 0 ? y : 0) : x;
                                ll_store(land + (((size_t)(epoch & 1) * 9 + (8 - k)) * p.land_len + idx) * 2, v,
                                         (uint32_t)epoch);
                            } else {
                                double* d = peers[k];
                                if (!d) continue;
                                const int dx = sx == -1 ? nx - 1 : (sx == 1 ? 0 : x);
                                const int dy = sy == -1 ? ny - 1 : (sy == 1 ? 0 : y);
                                d[(size_t)dx * ny + dy] = v;
                            }
                        }
                    }
                }
            }
        }
        if (ll && has_peers) {
            // ---- low-latency import: every halo cell polls its own landing slot until the epoch tag matches ----
            const uint64_t* land = p.my_land + (size_t)(epoch & 1) * 9 * p.land_len * 2;
            const uint32_t tag = (uint32_t)epoch;
            bool ok = true;
            double v;
            if (nbr[3] || nbr[5]) {                                        // columns 0 / ny-1 of my rows
                for (int lr = tid; lr < nr; lr += RES_THREADS) {
                    const int x = r0 + lr;
                    if (x < 1 || x > nx - 2) continue;
                    if (nbr[3]) { ok &= ll_load(land + ((size_t)3 * p.land_len + x) * 2, tag, v); dst[(size_t)lr * ny] = v; }
                    if (nbr[5]) { ok &= ll_load(land + ((size_t)5 * p.land_len + x) * 2, tag, v); dst[(size_t)lr * ny + ny - 1] = v; }
                }
            }
            for (int side = 0; side < 2; ++side) {                          // row 0 / row nx-1
                const bool own = side == 0 ? own_row0 : own_rowL;
                if (!own || nr == 0) continue;
                const int x = side == 0 ? 0 : nx - 1, lr = x - r0, o0 = side == 0 ? 0 : 6;
                if (nbr[o0 + 1])
                    for (int y = 1 + tid; y <= ny - 2; y += RES_THREADS) {
                        ok &= ll_load(land + ((size_t)(o0 + 1) * p.land_len + y) * 2, tag, v);
                        dst[(size_t)lr * ny + y] = v;
                    }
                if (tid == 0 && nbr[o0]) { ok &= ll_load(land + ((size_t)o0 * p.land_len) * 2, tag, v); dst[(size_t)lr * ny] = v; }
                if (tid == 1 && nbr[o0 + 2]) { ok &= ll_load(land + ((size_t)(o0 + 2) * p.land_len) * 2, tag, v); dst[(size_t)lr * ny + ny - 1] = v; }
            }
            if (!ok) atomicExch((unsigned long long*)err, (unsigned long long)epoch | (1ull << 63));
        }
        if (!ll && has_peers) __syncthreads();
        // ---- announce my pushes: one flag per (neighbour offset, CTA) -------------------------------------
        if (!ll && tid < 9 && tid != 4 && p.peer_flags[tid]) {
            const int sx = tid / 3 - 1;
            const bool mine = (sx == -1 ? own_row1 : (sx == 1 ? own_rowN : true)) && nr > 0;
            // y-direction pushes come from every CTA; x-direction / corner pushes only from the row owner
            if (mine) {
                __threadfence_system();
                st_release_sys(p.peer_flags[tid] + (8 - tid) * 16 + c, epoch);
            }
        }
        // ---- import the halo cells the neighbours pushed into my global array ------------------------------
        // which neighbours write into cells that live in MY rows: (0,-1)/(0,+1) for every CTA; (-1,*) for
        // the owner of row 0; (+1,*) for the owner of row nx-1.
        if (!ll && tid < 9 && tid != 4 && p.peer_flags[tid]) {
            const int sx = tid / 3 - 1;
            const bool mine = (sx == -1 ? own_row0 : (sx == 1 ? own_rowL : true)) && nr > 0;
            // the sender of offset (sx, sy) cells is CTA c for sx == 0, and the owner of ITS row nx-2 / row 1
            // for sx == -1 / +1 -- which is the same CTA index as my row nx-1 / row 0 owner only if the
            // partitions agree; they do (same nx, same R), but the sender's row owner may differ from me:
            // row nx-2 can belong to CTA C'-1 while row nx-1 ... both are in the last non-empty block or the
            // one before; resolve by asking for the sender's index explicitly.
            if (mine) {
                int sender = c;
                if (sx == -1) sender = (nx - 2) / R;        // neighbour's row nx-2 lands in my row 0
                if (sx == 1) sender = 1 / R;                // neighbour's row 1 lands in my row nx-1
                wait_flag_ge(p.my_flags + tid * 16 + sender, epoch, err);
            }
        }
        if (!ll && has_peers) {
            __syncthreads();
            const bool nb_yl = p.peer_flags[1 * 3 + 0] != nullptr, nb_yh = p.peer_flags[1 * 3 + 2] != nullptr;
            // columns 0 and ny-1 of my rows (rows 1..nx-2 only; corners belong to the row imports below)
            if (nb_yl || nb_yh) {
                for (int lr = tid; lr < nr; lr += RES_THREADS) {
                    const int x = r0 + lr;
                    if (x < 1 || x > nx - 2) continue;
                    if (nb_yl) dst[(size_t)lr * ny] = __ldcg(gdst + (size_t)x * ny);
                    if (nb_yh) dst[(size_t)lr * ny + ny - 1] = __ldcg(gdst + (size_t)x * ny + ny - 1);
                }
            }
            // row 0 / row nx-1: the cells that some neighbour actually owns (interior part from (-1,0)/(+1,0),
            // corners from the diagonal neighbours)
            for (int side = 0; side < 2; ++side) {
                const bool own = side == 0 ? own_row0 : own_rowL;
                if (!own || nr == 0) continue;
                const int x = side == 0 ? 0 : nx - 1, lr = x - r0, sxo = side == 0 ? 0 : 2;
                const bool mid = p.peer_flags[sxo * 3 + 1] != nullptr;
                const bool cl = p.peer_flags[sxo * 3 + 0] != nullptr, ch = p.peer_flags[sxo * 3 + 2] != nullptr;
                if (mid)
                    for (int y = 1 + tid; y <= ny - 2; y += RES_THREADS)
                        dst[(size_t)lr * ny + y] = __ldcg(gdst + (size_t)x * ny + y);
                if (tid == 0 && cl) dst[(size_t)lr * ny] = __ldcg(gdst + (size_t)x * ny);
                if (tid == 1 && ch) dst[(size_t)lr * ny + ny - 1] = __ldcg(gdst + (size_t)x * ny + ny - 1);
            }
        }
        cluster.sync();   // dst complete everywhere (incl. halos); nobody reads src any more
    }

    // write the final field back (the buffer the next step would read)
    const bool res_in_B = (p.nsteps & 1) == 1;
    const double* fin = res_in_B ? bufB : bufA;
    double* gout = res_in_B ? p.T2 : p.T;
    for (int e = tid; e < nr * ny; e += RES_THREADS) gout[(size_t)r0 * ny + e] = fin[e];
}

}  // namespace

extern "C" int igg_diffusion2d_resident(double* T, double* T2, const double* Cp, int nx, int ny, double lam, double dt,
                                        double _dx, double _dy, int nsteps, double* const* peer_T,
                                        double* const* peer_T2, uint64_t* const* peer_flags, uint64_t* my_flags,
                                        uint64_t epoch0, uint64_t* const* peer_land, uint64_t* my_land, int land_len,
                                        int* cluster_size_out, void* stream) {
    return igg_diffusion2d_resident_ex(T, T2, Cp, nx, ny, lam, dt, _dx, _dy, nsteps, 0, peer_T, peer_T2, peer_flags,
                                       my_flags, epoch0, peer_land, my_land, land_len, cluster_size_out, stream);
}

extern "C" int igg_diffusion2d_resident_ex(double* T, double* T2, const double* Cp, int nx, int ny, double lam,
                                           double dt, double _dx, double _dy, int nsteps, int form,
                                           double* const* peer_T, double* const* peer_T2,
                                           uint64_t* const* peer_flags, uint64_t* my_flags, uint64_t epoch0,
                                           uint64_t* const* peer_land, uint64_t* my_land, int land_len,
                                           int* cluster_size_out, void* stream) {
    IGG_REQUIRE(T && T2 && Cp && nx >= 3 && ny >= 3 && nsteps >= 1, "igg_diffusion2d_resident: bad arguments");
    IGG_REQUIRE(form == 0 || form == 1, "igg_diffusion2d_resident: form must be 0 (perf) or 1 (kp)");
    const bool comm = peer_flags != nullptr || peer_land != nullptr;
    IGG_REQUIRE(!comm || (my_flags && epoch0 >= 1), "igg_diffusion2d_resident: communication needs my_flags and epoch0 >= 1");
    IGG_REQUIRE(!peer_flags || (peer_T && peer_T2), "igg_diffusion2d_resident: flag mode needs the peer array tables");
    IGG_REQUIRE(!peer_land || (my_land && land_len >= (nx > ny ? nx : ny)),
                "igg_diffusion2d_resident: landing buffers need land_len >= max(nx, ny)");
    int dev = 0, smem_max = 0;
    IGG_CUDA_OK(cudaGetDevice(&dev));
    IGG_CUDA_OK(cudaDeviceGetAttribute(&smem_max, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
    // smallest power-of-two cluster whose per-CTA share (3 arrays of R rows) fits, aiming at <= 8 rows per CTA
    int C = 1, R = nx;
    for (C = 1; C <= RES_MAX_CTAS; C *= 2) {
        R = (nx + C - 1) / C;
        const size_t need = (size_t)3 * R * ny * sizeof(double);
        if (need <= (size_t)smem_max - 1024 && (R <= 8 || C == RES_MAX_CTAS)) break;
    }
    if (C > RES_MAX_CTAS) C = RES_MAX_CTAS;
    R = (nx + C - 1) / C;
    const size_t smem = (size_t)3 * R * ny * sizeof(double);
    if (smem > (size_t)smem_max - 1024 || R < 2) {
        igg::set_error("igg_diffusion2d_resident: a %d x %d field does not fit in the shared memory of a %d-CTA cluster "
                       "(%zu bytes per CTA needed, %d available); use igg_diffusion2d_multistep", nx, ny, C, smem,
                       smem_max);
        return IGG_ERR_UNSUPPORTED;
    }
    ResParams p;
    memset(&p, 0, sizeof(p));
    p.T = T; p.T2 = T2; p.Cp = Cp; p.nx = nx; p.ny = ny; p.R = R; p.nsteps = nsteps;
    p.lam = lam; p.dt = dt; p._dx = _dx; p._dy = _dy;
    p.form = form;
    if (comm) {
        for (int o = 0; o < 9; ++o) {
            if (peer_flags && !peer_land) {
                p.peer_T[o] = peer_T[o];
                p.peer_T2[o] = peer_T2[o];
                p.peer_flags[o] = peer_flags[o];
            }
            if (peer_land) p.peer_land[o] = peer_land[o];
        }
        p.my_flags = my_flags;
        p.epoch0 = epoch0;
        if (peer_land) {
            p.my_land = my_land;
            p.land_len = land_len;
        }
    }
    static bool attr_set[64] = {false};
    if (dev < 0 || dev >= 64 || !attr_set[dev]) {
        IGG_CUDA_OK(cudaFuncSetAttribute(diff2d_resident_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         smem_max - 1024));
        IGG_CUDA_OK(cudaFuncSetAttribute(diff2d_resident_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
        if (dev >= 0 && dev < 64) attr_set[dev] = true;
    }
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3(C);
    cfg.blockDim = dim3(RES_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = C;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    IGG_CUDA_OK(cudaLaunchKernelEx(&cfg, diff2d_resident_kernel, p));
    IGG_LAUNCH_CHECK();
    if (cluster_size_out) *cluster_size_out = C;
    return 0;
}
