// 3-D heat-diffusion step (7-point), fused form, for sm_100a.
//
// The reference is 2-D only (`init_global_grid(nx, ny, 1)`, scripts/diffusion_2D_ap.jl:17); 3-D is the
// ImplicitGlobalGrid capability the north star asks for.  Same structure as stencil2d.cu: tiles of
// the boundary frame are scheduled first, edge cells are pushed to the up-to-26 neighbours through
// peer pointers, modes 0/1/2 = all / frame / inside-the-frame.
//
//  * variant 1: persistent, warp-specialised 2.5-D kernel.  A work item is a column of 32 x 128 (y x z)
//    cells marched along x for up to 32 planes; a producer warp streams the T planes (34 x 132 doubles,
//    one-cell halo, TMA zero fill at array edges) through a 4-slot ring and the Cp planes through a
//    2-slot ring; 16 consumer warps compute plane x from the three resident T planes x-1, x, x+1 and
//    hand the oldest plane back.  Every T plane is read once per column (halo overhead only in y, z).
//  * variant 0: direct global loads, 8 x 8 x 64 tiles, marching along x with a register window.
// Layout: T[ix][iy][iz], iz contiguous.
#include <string.h>

#include "push.cuh"

namespace {
using namespace igg;

constexpr int TX = 8, TY = 8, TZ = 64;       // tile cells of the direct kernel
// 2.5-D TMA kernel: plane tile PTY x PTZ, marched over CX planes.  The TMA box must start on a 16-byte
// boundary (even z for fp64): it starts at z0-2 and is PTZ+4 wide, so cell (y,z) of a plane tile sits at
// smem row y-y0+1, column z-z0+2.
constexpr int CX = 32, PTY = 32, PTZ = 128;
constexpr int PP = PTZ + 4;                                        // smem pitch of a T plane (doubles)
constexpr int TPL_BYTES = (PTY + 2) * PP * 8;                      // 35904
constexpr int TPL_PAD = (TPL_BYTES + 127) / 128 * 128;             // 35968
constexpr int CPL_BYTES = PTY * PTZ * 8;                           // 32768
constexpr int NT = 4, NC = 2;                                      // ring depths (T planes, Cp planes)
constexpr int SMEM3_BYTES = NT * TPL_PAD + NC * CPL_BYTES + 128;   // 209536

struct Seg3 {
    int i0, j0, k0, ni, nj, nk, start;
};
struct TileSet3 {
    Seg3 s[7];
    int nseg, ntiles;
};

struct Params3 {
    const double* T;
    double* T2;
    const double* Cp;
    int nx, ny, nz;
    double lam, dt, _dx, _dy, _dz;
    int mode, bw[3];
    int t_hi[3];              // tile index containing n-2 per dim
    TileSet3 ts;
    int has_push;
    int edge_tiles[27];
    int side_nb[3][2];        // does ANY neighbour exist on the low / high side of each dimension
    igg_push3d push;
};

__device__ __forceinline__ void tile_of3(const TileSet3& ts, int t, int& i, int& j, int& k) {
    i = j = k = 0;
#pragma unroll
    for (int q = 0; q < 7; ++q) {
        if (q < ts.nseg) {
            const Seg3& s = ts.s[q];
            const int r = t - s.start;
            if (r >= 0 && r < s.ni * s.nj * s.nk) {
                k = s.k0 + r % s.nk;
                j = s.j0 + (r / s.nk) % s.nj;
                i = s.i0 + r / (s.nk * s.nj);
            }
        }
    }
}

__device__ __forceinline__ double update_cell3(double c, double xm, double xp, double ym, double yp, double zm,
                                               double zp, double cp, const Params3& p) {
    const double qxr = -p.lam * (xp - c) * p._dx, qxl = -p.lam * (c - xm) * p._dx;
    const double qyr = -p.lam * (yp - c) * p._dy, qyl = -p.lam * (c - ym) * p._dy;
    const double qzr = -p.lam * (zp - c) * p._dz, qzl = -p.lam * (c - zm) * p._dz;
    return c + p.dt * (cp * (-(qxr - qxl) * p._dx - (qyr - qyl) * p._dy - (qzr - qzl) * p._dz));
}

__device__ __forceinline__ bool in_frame3(const Params3& p, int x, int y, int z) {
    return x <= p.bw[0] || x >= p.nx - 1 - p.bw[0] || y <= p.bw[1] || y >= p.ny - 1 - p.bw[1] ||
           z <= p.bw[2] || z >= p.nz - 1 - p.bw[2];
}

__device__ __forceinline__ bool writes(const Params3& p, int x, int y, int z) {
    bool w = x >= 1 && x <= p.nx - 2 && y >= 1 && y <= p.ny - 2 && z >= 1 && z <= p.nz - 2;
    if (w && p.mode != 0) {
        const bool f = in_frame3(p, x, y, z);
        w = (p.mode == 1) ? f : !f;
    }
    return w;
}

// ---- variant 1: 2.5-D TMA kernel ----------------------------------------------------------------------

constexpr int RPT3 = 4;                                   // y rows per consumer thread
constexpr int CONS3 = (PTZ / 2) * (PTY / RPT3);           // 512 consumer threads
constexpr int NCW3 = CONS3 / 32;

__global__ void __launch_bounds__(CONS3 + 32, 1)
diff3d_tma_kernel(const __grid_constant__ CUtensorMap mapT, const __grid_constant__ CUtensorMap mapCp,
                  const __grid_constant__ Params3 p) {
    extern __shared__ __align__(128) unsigned char smem[];
    unsigned char* ringT = smem;
    unsigned char* ringC = smem + NT * TPL_PAD;
    uint64_t* fullT = reinterpret_cast<uint64_t*>(smem + NT * TPL_PAD + NC * CPL_BYTES);
    uint64_t* emptyT = fullT + NT;
    uint64_t* fullC = emptyT + NT;
    uint64_t* emptyC = fullC + NC;
    const int tid = threadIdx.x;
    const int first = blockIdx.x;
    const int n_my = first < p.ts.ntiles ? (p.ts.ntiles - first + gridDim.x - 1) / gridDim.x : 0;
    if (tid == 0) {
        for (int s = 0; s < NT; ++s) { mbar_init(&fullT[s], 1); mbar_init(&emptyT[s], NCW3); }
        for (int s = 0; s < NC; ++s) { mbar_init(&fullC[s], 1); mbar_init(&emptyC[s], NCW3); }
        fence_mbar_init();
    }
    __syncthreads();

    auto flags_of = [&](int i, int j, int k, bool lo[3], bool hi[3]) {
        lo[0] = i == 0 && p.side_nb[0][0]; lo[1] = j == 0 && p.side_nb[1][0]; lo[2] = k == 0 && p.side_nb[2][0];
        hi[0] = i == p.t_hi[0] && p.side_nb[0][1]; hi[1] = j == p.t_hi[1] && p.side_nb[1][1];
        hi[2] = k == p.t_hi[2] && p.side_nb[2][1];
    };

    if (tid >= CONS3) {
        // ===== producer: T planes xa-1 .. xb and Cp planes xa .. xb-1 of every work item ==================
        if (tid == CONS3) {
            tma_prefetch_desc(&mapT);
            tma_prefetch_desc(&mapCp);
            unsigned tq = 0, cq = 0;                        // planes issued so far (ring positions)
            auto issue_T = [&](int x, int j, int k) {
                const unsigned s = tq % NT;
                if (tq >= NT) mbar_wait(&emptyT[s], ((tq / NT) - 1) & 1);
                mbar_arrive_expect_tx(&fullT[s], TPL_BYTES);
                tma_load_3d(ringT + s * TPL_PAD, &mapT, &fullT[s], k * PTZ - 2, j * PTY - 1, x);
                ++tq;
            };
            auto issue_C = [&](int x, int j, int k) {
                const unsigned s = cq % NC;
                if (cq >= NC) mbar_wait(&emptyC[s], ((cq / NC) - 1) & 1);
                mbar_arrive_expect_tx(&fullC[s], CPL_BYTES);
                tma_load_3d(ringC + s * CPL_BYTES, &mapCp, &fullC[s], k * PTZ, j * PTY, x);
                ++cq;
            };
            for (int it = 0; it < n_my; ++it) {
                int i, j, k;
                tile_of3(p.ts, first + it * gridDim.x, i, j, k);
                if (p.has_push) {
                    bool lo[3], hi[3];
                    flags_of(i, j, k, lo, hi);
                    if (wait_faces_3d(p.push, lo, hi)) fence_proxy_async();
                }
                const int xa = i * CX, n = min(CX, p.nx - xa);
                issue_T(xa - 1, j, k);
                issue_T(xa, j, k);
                for (int m = 0; m < n; ++m) {
                    issue_T(xa + m + 1, j, k);
                    issue_C(xa + m, j, k);
                }
            }
        }
        return;
    }

    // ===== consumers ====================================================================================
    const int zp = tid % (PTZ / 2);       // z pair
    const int rg = tid / (PTZ / 2);       // row group (RPT3 rows)
    unsigned tq = 0, cq = 0;              // ring position of the item's first T plane / Cp plane
    for (int it = 0; it < n_my; ++it) {
        int i, j, k;
        tile_of3(p.ts, first + it * gridDim.x, i, j, k);
        bool lo[3], hi[3];
        flags_of(i, j, k, lo, hi);
        const bool edge_tile = p.has_push && (lo[0] || lo[1] || lo[2] || hi[0] || hi[1] || hi[2]);
        const int xa = i * CX, n = min(CX, p.nx - xa);
        const int gz = k * PTZ + 2 * zp, gyb = j * PTY + rg * RPT3;
        // planes q = 0 (x-1) and q = 1 (x) of the first compute plane
        mbar_wait(&fullT[tq % NT], (tq / NT) & 1);
        mbar_wait(&fullT[(tq + 1) % NT], ((tq + 1) / NT) & 1);
        for (int m = 0; m < n; ++m) {
            const unsigned q0 = tq + m, q1 = q0 + 1, q2 = q0 + 2, qc = cq + m;
            mbar_wait(&fullT[q2 % NT], (q2 / NT) & 1);
            mbar_wait(&fullC[qc % NC], (qc / NC) & 1);
            const int gx = xa + m;
            if (gx >= 1 && gx <= p.nx - 2) {
                const size_t off = (size_t)(rg * RPT3) * PP + 2 * zp + 2;   // row above the thread's first row
                const double* P0 = reinterpret_cast<const double*>(ringT + (q0 % NT) * TPL_PAD) + off;
                const double* P1 = reinterpret_cast<const double*>(ringT + (q1 % NT) * TPL_PAD) + off;
                const double* P2 = reinterpret_cast<const double*>(ringT + (q2 % NT) * TPL_PAD) + off;
                const double* sC = reinterpret_cast<const double*>(ringC + (qc % NC) * CPL_BYTES) +
                                   (size_t)(rg * RPT3) * PTZ + 2 * zp;
                // 3-row window over y inside the centre plane
                double2 a = *reinterpret_cast<const double2*>(P1);
                double2 b = *reinterpret_cast<const double2*>(P1 + PP);
#pragma unroll
                for (int r = 0; r < RPT3; ++r) {
                    const double2 c = *reinterpret_cast<const double2*>(P1 + (r + 2) * PP);
                    const double zl = P1[(r + 1) * PP - 1], zr = P1[(r + 1) * PP + 2];
                    const double2 xm = *reinterpret_cast<const double2*>(P0 + (r + 1) * PP);
                    const double2 xp = *reinterpret_cast<const double2*>(P2 + (r + 1) * PP);
                    const double2 cc = *reinterpret_cast<const double2*>(sC + r * PTZ);
                    const int gy = gyb + r;
                    const double vA = update_cell3(b.x, xm.x, xp.x, a.x, c.x, zl, b.y, cc.x, p);
                    const double vB = update_cell3(b.y, xm.y, xp.y, a.y, c.y, b.x, zr, cc.y, p);
                    const bool wA = writes(p, gx, gy, gz), wB = writes(p, gx, gy, gz + 1);
                    double* out = p.T2 + ((size_t)gx * p.ny + gy) * p.nz + gz;
                    if (wA && wB) {
                        st_cs_f64x2(out, vA, vB);
                    } else {
                        if (wA) st_cs_f64(out, vA);
                        if (wB) st_cs_f64(out + 1, vB);
                    }
                    if (edge_tile) {
                        if (wA) push_cell_3d(p.push, p.nx, p.ny, p.nz, gx, gy, gz, vA);
                        if (wB) push_cell_3d(p.push, p.nx, p.ny, p.nz, gx, gy, gz + 1, vB);
                    }
                    a = b;
                    b = c;
                }
            }
            // plane x-1 and this Cp plane are no longer needed
            __syncwarp();
            if ((tid & 31) == 0) {
                mbar_arrive(&emptyT[q0 % NT]);
                mbar_arrive(&emptyC[qc % NC]);
            }
        }
        // the last two T planes of the column
        __syncwarp();
        if ((tid & 31) == 0) {
            mbar_arrive(&emptyT[(tq + n) % NT]);
            mbar_arrive(&emptyT[(tq + n + 1) % NT]);
        }
        tq += n + 2;
        cq += n;
        if (edge_tile) {
            consumer_sync(CONS3);
            if (tid == 0) signal_edges_3d(p.push, p.edge_tiles, lo, hi);
        }
    }
}

// ---- variant 0: direct -----------------------------------------------------------------------------
__global__ void __launch_bounds__(256) diff3d_direct_kernel(const __grid_constant__ Params3 p) {
    int i, j, k;
    tile_of3(p.ts, blockIdx.x, i, j, k);
    bool lo[3] = {i == 0 && p.side_nb[0][0], j == 0 && p.side_nb[1][0], k == 0 && p.side_nb[2][0]};
    bool hi[3] = {i == p.t_hi[0] && p.side_nb[0][1], j == p.t_hi[1] && p.side_nb[1][1],
                  k == p.t_hi[2] && p.side_nb[2][1]};
    const bool edge_tile = p.has_push && (lo[0] || lo[1] || lo[2] || hi[0] || hi[1] || hi[2]);
    const int tid = threadIdx.x;
    if (edge_tile) {
        if (tid == 0) wait_faces_3d(p.push, lo, hi);
        __syncthreads();
    }
    // 256 threads = 64 z x 4 y-groups; each thread does 2 y rows x TX planes
    const int z = k * TZ + (tid % TZ);
    const int nx = p.nx, ny = p.ny, nz = p.nz;
    if (z >= 1 && z <= nz - 2) {
        for (int yr = 0; yr < 2; ++yr) {
            const int y = j * TY + (tid / TZ) * 2 + yr;
            if (y < 1 || y > ny - 2) continue;
            int x = i * TX < 1 ? 1 : i * TX;
            const int xe = min(i * TX + TX, nx - 1);
            if (x >= xe) continue;
            const size_t sx = (size_t)ny * nz;
            const double* Tc = p.T + ((size_t)x * ny + y) * nz + z;
            double um = Tc[-(ptrdiff_t)sx], c = Tc[0];
            for (; x < xe; ++x) {
                const double d = Tc[sx];
                const double v = update_cell3(c, um, d, Tc[-nz], Tc[nz], Tc[-1], Tc[1],
                                              p.Cp[((size_t)x * ny + y) * nz + z], p);
                if (writes(p, x, y, z)) {
                    p.T2[((size_t)x * ny + y) * nz + z] = v;
                    if (edge_tile) push_cell_3d(p.push, nx, ny, nz, x, y, z, v);
                }
                um = c;
                c = d;
                Tc += sx;
            }
        }
    }
    if (edge_tile) {
        __syncthreads();
        if (tid == 0) signal_edges_3d(p.push, p.edge_tiles, lo, hi);
    }
}

// ---- host ------------------------------------------------------------------------------------------
static void add_seg3(TileSet3& ts, int i0, int i1, int j0, int j1, int k0, int k1) {
    if (i1 < i0 || j1 < j0 || k1 < k0) return;
    Seg3& s = ts.s[ts.nseg++];
    s.i0 = i0; s.j0 = j0; s.k0 = k0;
    s.ni = i1 - i0 + 1; s.nj = j1 - j0 + 1; s.nk = k1 - k0 + 1;
    s.start = ts.ntiles;
    ts.ntiles += s.ni * s.nj * s.nk;
}

static void build_tiles3(TileSet3& ts, const int n[3], int mode, const int bw[3], const int tsz[3]) {
    ts.nseg = 0;
    ts.ntiles = 0;
    int nt[3], a0[3], a1[3];
    for (int d = 0; d < 3; ++d) nt[d] = (n[d] - 2) / tsz[d] + 1;
    if (mode == 2) {
        int lo[3], hi[3];
        for (int d = 0; d < 3; ++d) {
            lo[d] = 1 + bw[d];
            hi[d] = n[d] - 2 - bw[d];
            if (hi[d] < lo[d]) return;
        }
        add_seg3(ts, lo[0] / tsz[0], hi[0] / tsz[0], lo[1] / tsz[1], hi[1] / tsz[1], lo[2] / tsz[2], hi[2] / tsz[2]);
        return;
    }
    for (int d = 0; d < 3; ++d) {
        a0[d] = (bw[d] < n[d] - 2 ? bw[d] : n[d] - 2) / tsz[d];
        a1[d] = (n[d] - 1 - bw[d] > 1 ? n[d] - 1 - bw[d] : 1) / tsz[d];
    }
    // peel slabs dimension by dimension; [l,h] is the still-unassigned tile range per dim
    int l[3] = {0, 0, 0}, h[3] = {nt[0] - 1, nt[1] - 1, nt[2] - 1};
    for (int d = 0; d < 3; ++d) {
        int r0[3] = {l[0], l[1], l[2]}, r1[3] = {h[0], h[1], h[2]};
        if (a1[d] <= a0[d]) {   // bands meet: the whole remaining range is frame
            add_seg3(ts, r0[0], r1[0], r0[1], r1[1], r0[2], r1[2]);
            return;
        }
        r1[d] = a0[d];
        add_seg3(ts, r0[0], r1[0], r0[1], r1[1], r0[2], r1[2]);
        r0[d] = a1[d];
        r1[d] = h[d];
        add_seg3(ts, r0[0], r1[0], r0[1], r1[1], r0[2], r1[2]);
        l[d] = a0[d] + 1;
        h[d] = a1[d] - 1;
        if (h[d] < l[d]) return;
    }
    if (mode == 0) add_seg3(ts, l[0], h[0], l[1], h[1], l[2], h[2]);
}

static bool g_attr3_set[64] = {false};

}  // namespace

extern "C" int igg_diffusion3d_step(const double* T, double* T2, const double* Cp, int nx, int ny, int nz, double lam,
                                    double dt, double _dx, double _dy, double _dz, int mode, int bwx, int bwy, int bwz,
                                    int variant, const igg_push3d* push, void* stream) {
    IGG_REQUIRE(T && T2 && Cp, "igg_diffusion3d_step: null field pointer");
    IGG_REQUIRE(nx >= 3 && ny >= 3 && nz >= 3, "igg_diffusion3d_step: need nx, ny, nz >= 3");
    IGG_REQUIRE(mode >= 0 && mode <= 2 && (variant == 0 || variant == 1), "igg_diffusion3d_step: bad mode/variant");
    IGG_REQUIRE(mode == 0 || (bwx >= 1 && bwy >= 1 && bwz >= 1), "igg_diffusion3d_step: b_width must be >= 1");
    IGG_REQUIRE(!(push && mode == 2), "igg_diffusion3d_step: the interior pass (mode 2) never pushes");
    Params3 p;
    p.T = T; p.T2 = T2; p.Cp = Cp; p.nx = nx; p.ny = ny; p.nz = nz;
    p.lam = lam; p.dt = dt; p._dx = _dx; p._dy = _dy; p._dz = _dz;
    p.mode = mode;
    p.bw[0] = bwx < 1 ? 1 : bwx; p.bw[1] = bwy < 1 ? 1 : bwy; p.bw[2] = bwz < 1 ? 1 : bwz;
    const int n[3] = {nx, ny, nz};
    const int tsz_direct[3] = {TX, TY, TZ}, tsz_tma[3] = {CX, PTY, PTZ};
    const int* tsz = variant == 1 ? tsz_tma : tsz_direct;
    int nt[3];
    for (int d = 0; d < 3; ++d) {
        p.t_hi[d] = (n[d] - 2) / tsz[d];
        nt[d] = p.t_hi[d] + 1;
    }
    build_tiles3(p.ts, n, mode, p.bw, tsz);
    if (p.ts.ntiles == 0) return 0;
    p.has_push = push ? 1 : 0;
    for (int sx = -1; sx <= 1; ++sx)
        for (int sy = -1; sy <= 1; ++sy)
            for (int sz = -1; sz <= 1; ++sz)
                p.edge_tiles[((sx + 1) * 3 + (sy + 1)) * 3 + (sz + 1)] =
                    (sx ? 1 : nt[0]) * (sy ? 1 : nt[1]) * (sz ? 1 : nt[2]);
    memset(p.side_nb, 0, sizeof(p.side_nb));
    if (push) {
        p.push = *push;
        IGG_REQUIRE(push->epoch >= 1, "igg_diffusion3d_step: push epoch must be >= 1");
        for (int o = 0; o < 27; ++o) {
            if (!push->peer_dst[o]) continue;
            const int s3[3] = {o / 9 - 1, (o / 3) % 3 - 1, o % 3 - 1};
            for (int d = 0; d < 3; ++d) {
                if (s3[d] == -1) p.side_nb[d][0] = 1;
                if (s3[d] == 1) p.side_nb[d][1] = 1;
            }
        }
    } else {
        memset(&p.push, 0, sizeof(p.push));
    }
    cudaStream_t st = (cudaStream_t)stream;
    if (variant == 1) {
        if ((nz & 1) || (((uintptr_t)T | (uintptr_t)T2 | (uintptr_t)Cp) & 15)) {
            igg::set_error("igg_diffusion3d_step: the TMA variant needs even nz and 16-byte aligned fields");
            return IGG_ERR_UNSUPPORTED;
        }
        CUtensorMap mT, mC;
        const uint64_t dims[3] = {(uint64_t)nz, (uint64_t)ny, (uint64_t)nx};
        const uint32_t boxT[3] = {PP, PTY + 2, 1};
        const uint32_t boxC[3] = {PTZ, PTY, 1};
        int rc = igg::get_tensor_map_f64(T, 3, dims, boxT, &mT);
        if (rc) return rc;
        rc = igg::get_tensor_map_f64(Cp, 3, dims, boxC, &mC);
        if (rc) return rc;
        int dev = 0;
        IGG_CUDA_OK(cudaGetDevice(&dev));
        if (dev < 0 || dev >= 64 || !g_attr3_set[dev]) {
            IGG_CUDA_OK(cudaFuncSetAttribute(diff3d_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM3_BYTES));
            if (dev >= 0 && dev < 64) g_attr3_set[dev] = true;
        }
        const int grid = p.ts.ntiles < igg::sm_count() ? p.ts.ntiles : igg::sm_count();
        diff3d_tma_kernel<<<grid, CONS3 + 32, SMEM3_BYTES, st>>>(mT, mC, p);
        IGG_LAUNCH_CHECK();
    } else {
        diff3d_direct_kernel<<<p.ts.ntiles, 256, 0, st>>>(p);
        IGG_LAUNCH_CHECK();
    }
    return 0;
}
