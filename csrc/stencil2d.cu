// 2-D heat-diffusion step, fused form, for sm_100a.
//
// Provides the behaviour of the reference's fused kernel `diffusion_step!`
// (scripts/diffusion_2D_perf.jl:3-13) and of its boundary/interior split
// (scripts/diffusion_2D_perf_hide.jl:15-29, predicates :20-21), plus -- when a push descriptor is
// given -- the halo exchange that the reference performs afterwards with update_halo!
// (scripts/diffusion_2D_perf.jl:51).  The design is new:
//
//  * variant 1 (product): a persistent, warp-specialised kernel, one CTA per SM.  A producer warp
//    walks the CTA's static list of 32x128-cell tiles and keeps a 3-stage ring of TMA loads in
//    flight (the T tile with its one-cell halo -- out-of-bounds zero fill handles array edges -- and
//    the matching Cp tile); "full" mbarriers carry the TMA completion to the consumers and "empty"
//    mbarriers hand finished stages back.  Sixteen consumer warps compute from shared memory with a
//    3-row register window and 16-byte loads and write T2 with 16-byte streaming stores.  There is no
//    CTA-wide barrier in the steady state.
//  * boundary-frame tiles come first in the list, so edge cells are pushed into the neighbours'
//    halo cells (peer stores over NVLink) and their epoch flag is released while the bulk of the
//    interior is still being computed: communication is hidden inside ONE launch.  Waiting for the
//    neighbours' previous epoch is done by the producer, ahead of the consumers.
//  * optional in-kernel diagnostic {max, sum} of the written cells (no extra pass over the field).
//  * variant 0: straightforward direct-load kernel with the same tiling, masks and push logic;
//    it is the "recompiled baseline" flavour and the fallback for odd ny.
//
// Layout: T[ix][iy], iy contiguous.  "x" rows are strided, "y" columns contiguous.
#include <stdlib.h>
#include <string.h>

#include "push.cuh"

namespace {

using namespace igg;

constexpr int BX = 32;          // tile rows
constexpr int BY = 128;         // tile columns (contiguous)
// TMA needs the box to start on a 16-byte boundary in global memory, i.e. at an EVEN column for fp64.
// Tiles start at even columns, so the halo'd box starts two columns to the left (y0-2) and is BY+4
// wide; the cell (x, y) of a tile sits at smem row x-x0+1, column y-y0+2.
constexpr int TP = BY + 4;      // smem pitch of the T tile (doubles)
constexpr int T_TILE_BYTES = (BX + 2) * TP * 8;                      // 35904
constexpr int T_TILE_PAD = (T_TILE_BYTES + 127) / 128 * 128;         // 35968
constexpr int CP_TILE_BYTES = BX * BY * 8;                           // 32768
constexpr int STAGE_BYTES = T_TILE_PAD + CP_TILE_BYTES;              // 68736
constexpr int STAGES = 3;
constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 128;               // + mbarriers + tile ring

// same layout as the scratch of igg_reduce_max_sum (csrc/misc.cu)
struct DiagScratch {
    unsigned int ticket;
    unsigned int pad[3];
    double pmax[1024];
    double psum[1024];
};

struct Seg {
    int i0, j0, w, h, start;    // tiles [i0,i0+h) x [j0,j0+w); `start` = linear id of its first tile
    int band;                   // > 0: walk the segment in vertical bands of `band` tile columns
};
struct TileSet {
    Seg s[5];
    int nseg, ntiles;
};

struct Params {
    const double* T;
    double* T2;
    const double* Cp;
    int nx, ny;
    double lam, dt, _dx, _dy;
    const double* dev_params;   // optional {lam, dt, _dx, _dy} in device memory
    double* diag_out;           // optional {max, sum} of the written cells
    DiagScratch* diag;          // block partials + ticket for diag_out
    int mode, bwx, bwy;
    int i_hi, j_hi;             // tile row/column that contains x==nx-2 / y==ny-2
    TileSet ts;
    int has_push;
    int edge_tiles[9];
    int side_nb[4];             // does ANY neighbour exist on the x-low / x-high / y-low / y-high side
    int scx;                    // strip kernel: rows per work item (multiple of its block height)
    igg_push2d push;
};

__device__ __forceinline__ void tile_of(const TileSet& ts, int t, int& i, int& j) {
    i = 0;
    j = 0;
#pragma unroll
    for (int k = 0; k < 5; ++k) {
        if (k < ts.nseg) {
            const Seg& s = ts.s[k];
            const int r = t - s.start;
            if (r >= 0 && r < s.w * s.h) {
                if (s.band > 0) {
                    // banded order: consecutive tile ids sweep `band` columns, then drop one tile row
                    const int per_band = s.band * s.h;
                    const int b = r / per_band, rr = r - b * per_band;
                    const int jw = min(s.band, s.w - b * s.band);
                    i = s.i0 + rr / jw;
                    j = s.j0 + b * s.band + rr % jw;
                } else {
                    i = s.i0 + r / s.w;
                    j = s.j0 + r % s.w;
                }
            }
        }
    }
}

// The arithmetic of reference scripts/diffusion_2D_perf.jl:8-10, operation for operation.
__device__ __forceinline__ double update_cell(double c, double xm, double xp, double ym, double yp, double cp,
                                              double lam, double dt, double _dx, double _dy) {
    const double qxr = -lam * (xp - c) * _dx;
    const double qxl = -lam * (c - xm) * _dx;
    const double qyt = -lam * (yp - c) * _dy;
    const double qyb = -lam * (c - ym) * _dy;
    return c + dt * (cp * (-(qxr - qxl) * _dx - (qyt - qyb) * _dy));
}

// is (x,y) in the boundary frame?  (interior test done by the caller)
__device__ __forceinline__ bool in_frame(const Params& p, int x, int y) {
    return x <= p.bwx || x >= p.nx - 1 - p.bwx || y <= p.bwy || y >= p.ny - 1 - p.bwy;
}


// Fused diagnostic {max, sum}, shared by the tile and the strip kernel: every consumer thread brings its partial;
// CTA partial -> scratch, the last CTA (ticket) folds all partials in a fixed order (deterministic sum) and
// re-arms the ticket.  The hot loops fold with fmax (one instruction, drops NaN); a NaN anywhere makes the SUM
// NaN, and then the max is reported as NaN too -- what the reference's maximum(T_v) would show
// (scripts/diffusion_2D_perf_hide.jl:115).
template <int CONSUMERS>
__device__ __forceinline__ void diag_finish(DiagScratch* scratch, double* diag_out, double dmax, double dsum) {
    constexpr int NCW = CONSUMERS / 32;
    const int tid = threadIdx.x;
    __shared__ double red[64];
    __shared__ bool is_last;
    auto block_fold = [&](double& mx, double& sm) {
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            mx = igg::max_fast(mx, __shfl_xor_sync(0xffffffffu, mx, off));
            sm += __shfl_xor_sync(0xffffffffu, sm, off);
        }
        if ((tid & 31) == 0) { red[tid >> 5] = mx; red[32 + (tid >> 5)] = sm; }
        consumer_sync(CONSUMERS);
        if (tid < 32) {
            mx = tid < NCW ? red[tid] : -INFINITY;
            sm = tid < NCW ? red[32 + tid] : 0.0;
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) {
                mx = igg::max_fast(mx, __shfl_xor_sync(0xffffffffu, mx, off));
                sm += __shfl_xor_sync(0xffffffffu, sm, off);
            }
        }
        consumer_sync(CONSUMERS);
    };
    block_fold(dmax, dsum);
    if (tid == 0) {
        scratch->pmax[blockIdx.x] = dmax;
        scratch->psum[blockIdx.x] = dsum;
        __threadfence();
        is_last = atomicAdd(&scratch->ticket, 1u) == gridDim.x - 1;
    }
    consumer_sync(CONSUMERS);
    if (is_last) {
        __threadfence();
        dmax = -INFINITY;
        dsum = 0.0;
        for (unsigned int b2 = tid; b2 < gridDim.x; b2 += CONSUMERS) {
            dmax = igg::max_fast(dmax, scratch->pmax[b2]);
            dsum += scratch->psum[b2];
        }
        block_fold(dmax, dsum);
        if (tid == 0) {
            diag_out[0] = dsum != dsum ? dsum : dmax;
            diag_out[1] = dsum;
            scratch->ticket = 0;
        }
    }
}


// ------------------------------------------------------------------------------------------------
// variant 1: persistent, warp-specialised, TMA + mbarrier pipeline
// ------------------------------------------------------------------------------------------------
template <int RPT>   // rows per consumer thread
__global__ void __launch_bounds__((BY / 2) * (BX / RPT) + 32, 1)
diff2d_tma_kernel(const __grid_constant__ CUtensorMap mapT, const __grid_constant__ CUtensorMap mapCp,
                  const __grid_constant__ Params p) {
    constexpr int CONSUMERS = (BY / 2) * (BX / RPT);
    constexpr int NCW = CONSUMERS / 32;
    extern __shared__ __align__(128) unsigned char smem[];
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE_BYTES);   // TMA landed
    uint64_t* empty = full + STAGES;                                              // consumers done
    int2* tinfo = reinterpret_cast<int2*>(empty + STAGES);                        // (i, j) of the staged tile
    const int tid = threadIdx.x;
    const int first = blockIdx.x;
    const int n_my = first < p.ts.ntiles ? (p.ts.ntiles - first + gridDim.x - 1) / gridDim.x : 0;

    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], NCW);
        }
        fence_mbar_init();
    }
    __syncthreads();

    if (tid >= CONSUMERS) {
        // ===== producer warp: one lane walks the tile list, waits for halos, issues TMA ===========
        if (tid == CONSUMERS) {
            tma_prefetch_desc(&mapT);
            tma_prefetch_desc(&mapCp);
            for (int k = 0; k < n_my; ++k) {
                const int s = k % STAGES;
                if (k >= STAGES) mbar_wait(&empty[s], ((k / STAGES) - 1) & 1);   // stage handed back
                int i, j;
                tile_of(p.ts, first + k * gridDim.x, i, j);
                tinfo[s] = make_int2(i, j);
                // only tiles that read a neighbour's halo plane wait; the proxy fence orders the acquired
                // flag (generic proxy) before the TMA read (async proxy)
                if (p.has_push && wait_faces_2d(p.push, i == 0, i == p.i_hi, j == 0, j == p.j_hi))
                    fence_proxy_async();
                unsigned char* st = smem + s * STAGE_BYTES;
                mbar_arrive_expect_tx(&full[s], T_TILE_BYTES + CP_TILE_BYTES);   // also publishes tinfo[s]
                tma_load_2d(st, &mapT, &full[s], j * BY - 2, i * BX - 1);
                tma_load_2d(st + T_TILE_PAD, &mapCp, &full[s], j * BY, i * BX);
            }
        }
        return;
    }

    // ===== consumer warps ============================================================================
    const int cp = tid % (BY / 2);   // column pair
    const int rg = tid / (BY / 2);   // row group
    const double lam = p.dev_params ? p.dev_params[0] : p.lam, dt = p.dev_params ? p.dev_params[1] : p.dt;
    const double _dx = p.dev_params ? p.dev_params[2] : p._dx, _dy = p.dev_params ? p.dev_params[3] : p._dy;
    const bool diag = p.diag_out != nullptr;
    double dmax = -INFINITY, dsum = 0.0;

    for (int k = 0; k < n_my; ++k) {
        const int s = k % STAGES;
        mbar_wait(&full[s], (k / STAGES) & 1);
        const int2 ij = tinfo[s];
        const int i = ij.x, j = ij.y;

        // this thread's two cells sit at smem columns 2cp+2, 2cp+3 (16-byte aligned pair)
        const double* sT = reinterpret_cast<const double*>(smem + s * STAGE_BYTES) + (rg * RPT) * TP + 2 * cp + 2;
        const double* sC = reinterpret_cast<const double*>(smem + s * STAGE_BYTES + T_TILE_PAD) +
                           (rg * RPT) * BY + 2 * cp;
        const int y = j * BY + 2 * cp;
        const int xb = i * BX + rg * RPT;
        // tile touches a side of the local domain behind which a neighbour exists
        const bool e_xl = i == 0 && p.side_nb[0], e_xh = i == p.i_hi && p.side_nb[1];
        const bool e_yl = j == 0 && p.side_nb[2], e_yh = j == p.j_hi && p.side_nb[3];
        const bool edge_tile = p.has_push && (e_xl || e_xh || e_yl || e_yh);
        // column validity is fixed per thread
        const bool vy0 = (y >= 1) && (y <= p.ny - 2);
        const bool vy1 = (y + 1 <= p.ny - 2);

        // 3-row register window over x: a = row x-1, b = row x, c = row x+1 (centre pairs)
        double2 a = *reinterpret_cast<const double2*>(sT);
        double2 b = *reinterpret_cast<const double2*>(sT + TP);
#pragma unroll
        for (int r = 0; r < RPT; ++r) {
            const double2 c = *reinterpret_cast<const double2*>(sT + (r + 2) * TP);
            const double bl = sT[(r + 1) * TP - 1];
            const double br = sT[(r + 1) * TP + 2];
            const double2 cc = *reinterpret_cast<const double2*>(sC + r * BY);
            const int x = xb + r;
            const double vA = update_cell(b.x, a.x, c.x, bl, b.y, cc.x, lam, dt, _dx, _dy);
            const double vB = update_cell(b.y, a.y, c.y, b.x, br, cc.y, lam, dt, _dx, _dy);
            const bool vx = (x >= 1) && (x <= p.nx - 2);
            bool wA = vx && vy0, wB = vx && vy1;
            if (p.mode != 0) {
                const bool fA = in_frame(p, x, y), fB = in_frame(p, x, y + 1);
                wA = wA && (p.mode == 1 ? fA : !fA);
                wB = wB && (p.mode == 1 ? fB : !fB);
            }
            double* out = p.T2 + (size_t)x * p.ny + y;
            if (wA && wB) {
                st_cs_f64x2(out, vA, vB);
            } else {
                if (wA) st_cs_f64(out, vA);
                if (wB) st_cs_f64(out + 1, vB);
            }
            if (edge_tile) {
                if (wA) push_cell_2d(p.push, p.nx, p.ny, x, y, vA);
                if (wB) push_cell_2d(p.push, p.nx, p.ny, x, y + 1, vB);
            }
            if (diag) {
                if (wA) { dmax = igg::max_fast(dmax, vA); dsum += vA; }
                if (wB) { dmax = igg::max_fast(dmax, vB); dsum += vB; }
            }
            a = b;
            b = c;
        }
        // hand the stage back to the producer: one arrival per consumer warp
        __syncwarp();
        if ((tid & 31) == 0) mbar_arrive(&empty[s]);
        if (edge_tile) {
            consumer_sync(CONSUMERS);   // every remote store of this tile has been issued
            if (tid == 0) signal_edges_2d(p.push, p.edge_tiles, e_xl, e_xh, e_yl, e_yh);
        }
    }

    // fused diagnostic: CTA partial -> scratch, last CTA (ticket) folds all partials in a fixed order
    if (diag) diag_finish<CONSUMERS>(p.diag, p.diag_out, dmax, dsum);
}

// ------------------------------------------------------------------------------------------------
// variant 1, whole-interior steps (mode 0): strip-marching kernel
// ------------------------------------------------------------------------------------------------
// A work item is a strip of BY columns marched down x for p.scx rows in blocks of SRB rows.  The producer
// streams the T blocks (SRB x (BY+4), no row halo) through a 4-slot ring, the Cp blocks through a 2-slot
// ring and the two halo rows of the chunk (above the first / below the last block) into a small
// double-buffered slot; consumers compute block b from blocks b-1, b, b+1 and hand block b-1 back.  Every
// row of T is fetched once per strip: the two halo rows per 32-row tile that the tile kernel re-fetches
// (6 % of its T traffic) shrink to two per 256 rows.
constexpr int SRB = 32;                     // rows per block
constexpr int SCX_DEFAULT = 8 * SRB;        // rows per work item (256); IGG_STRIP_ROWS overrides
constexpr int SNT = 4, SNC = 2;             // ring depths
constexpr int ST_BYTES = SRB * TP * 8;      // 33792 (multiple of 128)
constexpr int SH_BYTES = TP * 8;            // 1056: one halo row
constexpr int SH_PAD = 1152;
constexpr int SSMEM_BYTES = SNT * ST_BYTES + SNC * CP_TILE_BYTES + 4 * SH_PAD + 256;

template <int RPT>
__global__ void __launch_bounds__((BY / 2) * (SRB / RPT) + 32, 1)
diff2d_strip_kernel(const __grid_constant__ CUtensorMap mapT, const __grid_constant__ CUtensorMap mapRow,
                    const __grid_constant__ CUtensorMap mapCp, const __grid_constant__ Params p) {
    constexpr int CONSUMERS = (BY / 2) * (SRB / RPT);
    constexpr int NCW = CONSUMERS / 32;
    extern __shared__ __align__(128) unsigned char smem[];
    unsigned char* ringT = smem;
    unsigned char* ringC = smem + SNT * ST_BYTES;
    unsigned char* halo = ringC + SNC * CP_TILE_BYTES;          // [parity][top, bottom]
    uint64_t* fullT = reinterpret_cast<uint64_t*>(halo + 4 * SH_PAD);
    uint64_t* emptyT = fullT + SNT;
    uint64_t* fullC = emptyT + SNT;
    uint64_t* emptyC = fullC + SNC;
    uint64_t* fullH = emptyC + SNC;
    uint64_t* emptyH = fullH + 2;
    const int tid = threadIdx.x;
    const int first = blockIdx.x;
    const int n_my = first < p.ts.ntiles ? (p.ts.ntiles - first + gridDim.x - 1) / gridDim.x : 0;
    if (tid == 0) {
        for (int s = 0; s < SNT; ++s) { mbar_init(&fullT[s], 1); mbar_init(&emptyT[s], NCW); }
        for (int s = 0; s < SNC; ++s) { mbar_init(&fullC[s], 1); mbar_init(&emptyC[s], NCW); }
        for (int s = 0; s < 2; ++s) { mbar_init(&fullH[s], 1); mbar_init(&emptyH[s], NCW); }
        fence_mbar_init();
    }
    __syncthreads();

    if (tid >= CONSUMERS) {
        // ===== producer ===================================================================================
        if (tid == CONSUMERS) {
            tma_prefetch_desc(&mapT);
            tma_prefetch_desc(&mapRow);
            tma_prefetch_desc(&mapCp);
            unsigned tq = 0, cq = 0;
            auto issue_T = [&](int x, int j) {
                const unsigned s = tq % SNT;
                if (tq >= SNT) mbar_wait(&emptyT[s], ((tq / SNT) - 1) & 1);
                mbar_arrive_expect_tx(&fullT[s], ST_BYTES);
                tma_load_2d(ringT + s * ST_BYTES, &mapT, &fullT[s], j * BY - 2, x);
                ++tq;
            };
            auto issue_C = [&](int x, int j) {
                const unsigned s = cq % SNC;
                if (cq >= SNC) mbar_wait(&emptyC[s], ((cq / SNC) - 1) & 1);
                mbar_arrive_expect_tx(&fullC[s], CP_TILE_BYTES);
                tma_load_2d(ringC + s * CP_TILE_BYTES, &mapCp, &fullC[s], j * BY, x);
                ++cq;
            };
            for (int it = 0; it < n_my; ++it) {
                int ci, j;
                tile_of(p.ts, first + it * gridDim.x, ci, j);
                if (p.has_push && wait_faces_2d(p.push, ci == 0, ci == p.i_hi, j == 0, j == p.j_hi))
                    fence_proxy_async();
                const int x0 = ci * p.scx;
                const int nb = (min(p.scx, p.nx - x0) + SRB - 1) / SRB;
                const int hp = it & 1;
                if (it >= 2) mbar_wait(&emptyH[hp], ((it / 2) - 1) & 1);
                mbar_arrive_expect_tx(&fullH[hp], 2 * SH_BYTES);
                tma_load_2d(halo + (2 * hp) * SH_PAD, &mapRow, &fullH[hp], j * BY - 2, x0 - 1);
                tma_load_2d(halo + (2 * hp + 1) * SH_PAD, &mapRow, &fullH[hp], j * BY - 2, x0 + nb * SRB);
                issue_T(x0, j);
                for (int b = 0; b < nb; ++b) {
                    if (b + 1 < nb) issue_T(x0 + (b + 1) * SRB, j);
                    issue_C(x0 + b * SRB, j);
                }
            }
        }
        return;
    }

    // ===== consumers ====================================================================================
    const int cp = tid % (BY / 2);   // column pair
    const int rg = tid / (BY / 2);   // row group inside a block
    const double lam = p.dev_params ? p.dev_params[0] : p.lam, dt = p.dev_params ? p.dev_params[1] : p.dt;
    const double _dx = p.dev_params ? p.dev_params[2] : p._dx, _dy = p.dev_params ? p.dev_params[3] : p._dy;
    const bool diag = p.diag_out != nullptr;
    double dmax = -INFINITY, dsum = 0.0;
    unsigned tq = 0, cq = 0;
    const int col = 2 * cp + 2;      // smem column of the thread's first cell

    for (int it = 0; it < n_my; ++it) {
        int ci, j;
        tile_of(p.ts, first + it * gridDim.x, ci, j);
        const int x0 = ci * p.scx;
        const int nb = (min(p.scx, p.nx - x0) + SRB - 1) / SRB;
        const int hp = it & 1;
        const bool e_xl = ci == 0 && p.side_nb[0], e_xh = ci == p.i_hi && p.side_nb[1];
        const bool e_yl = j == 0 && p.side_nb[2], e_yh = j == p.j_hi && p.side_nb[3];
        const bool edge_tile = p.has_push && (e_xl || e_xh || e_yl || e_yh);
        const int y = j * BY + 2 * cp;
        const bool vy0 = (y >= 1) && (y <= p.ny - 2);
        const bool vy1 = (y + 1 <= p.ny - 2);
        mbar_wait(&fullH[hp], (it / 2) & 1);
        mbar_wait(&fullT[tq % SNT], (tq / SNT) & 1);
        const double* hTop = reinterpret_cast<const double*>(halo + (2 * hp) * SH_PAD);
        const double* hBot = reinterpret_cast<const double*>(halo + (2 * hp + 1) * SH_PAD);
        for (int b = 0; b < nb; ++b) {
            const unsigned qb = tq + b, qc = cq + b;
            if (b + 1 < nb) mbar_wait(&fullT[(qb + 1) % SNT], ((qb + 1) / SNT) & 1);
            mbar_wait(&fullC[qc % SNC], (qc / SNC) & 1);
            const double* cur = reinterpret_cast<const double*>(ringT + (qb % SNT) * ST_BYTES);
            // row above the block's first row / below its last row live in the neighbouring ring slots
            // (or in the chunk's halo rows)
            const double* above = b > 0 ? reinterpret_cast<const double*>(ringT + ((qb - 1) % SNT) * ST_BYTES) +
                                              (SRB - 1) * TP
                                        : hTop;
            const double* below = b + 1 < nb ? reinterpret_cast<const double*>(ringT + ((qb + 1) % SNT) * ST_BYTES)
                                             : hBot;
            const double* sC = reinterpret_cast<const double*>(ringC + (qc % SNC) * CP_TILE_BYTES) +
                               (rg * RPT) * BY + 2 * cp;
            const int lr0 = rg * RPT;
            const double* r_m1 = lr0 > 0 ? cur + (lr0 - 1) * TP : above;
            double2 a = *reinterpret_cast<const double2*>(r_m1 + col);
            double2 bb = *reinterpret_cast<const double2*>(cur + lr0 * TP + col);
            const int xb = x0 + b * SRB + lr0;
#pragma unroll
            for (int r = 0; r < RPT; ++r) {
                const double* rc = cur + (lr0 + r) * TP;                                   // this row
                const double* rn = (lr0 + r + 1 < SRB) ? rc + TP : below;                  // next row
                const double2 c = *reinterpret_cast<const double2*>(rn + col);
                const double bl = rc[col - 1];
                const double br = rc[col + 2];
                const double2 cc = *reinterpret_cast<const double2*>(sC + r * BY);
                const int x = xb + r;
                const double vA = update_cell(bb.x, a.x, c.x, bl, bb.y, cc.x, lam, dt, _dx, _dy);
                const double vB = update_cell(bb.y, a.y, c.y, bb.x, br, cc.y, lam, dt, _dx, _dy);
                const bool vx = (x >= 1) && (x <= p.nx - 2);
                bool wA = vx && vy0, wB = vx && vy1;
                if (p.mode != 0) {
                    const bool fA = in_frame(p, x, y), fB = in_frame(p, x, y + 1);
                    wA = wA && (p.mode == 1 ? fA : !fA);
                    wB = wB && (p.mode == 1 ? fB : !fB);
                }
                double* out = p.T2 + (size_t)x * p.ny + y;
                if (wA && wB) {
                    st_cs_f64x2(out, vA, vB);
                } else {
                    if (wA) st_cs_f64(out, vA);
                    if (wB) st_cs_f64(out + 1, vB);
                }
                if (edge_tile) {
                    if (wA) push_cell_2d(p.push, p.nx, p.ny, x, y, vA);
                    if (wB) push_cell_2d(p.push, p.nx, p.ny, x, y + 1, vB);
                }
                if (diag) {
                    if (wA) { dmax = igg::max_fast(dmax, vA); dsum += vA; }
                    if (wB) { dmax = igg::max_fast(dmax, vB); dsum += vB; }
                }
                a = bb;
                bb = c;
            }
            // block b-1 and this Cp block are no longer needed
            __syncwarp();
            if ((tid & 31) == 0) {
                if (b > 0) mbar_arrive(&emptyT[(qb - 1) % SNT]);
                mbar_arrive(&emptyC[qc % SNC]);
            }
        }
        __syncwarp();
        if ((tid & 31) == 0) {
            mbar_arrive(&emptyT[(tq + nb - 1) % SNT]);
            mbar_arrive(&emptyH[hp]);
        }
        tq += nb;
        cq += nb;
        if (edge_tile) {
            consumer_sync(CONSUMERS);
            if (tid == 0) signal_edges_2d(p.push, p.edge_tiles, e_xl, e_xh, e_yl, e_yh);
        }
    }

    if (diag) diag_finish<CONSUMERS>(p.diag, p.diag_out, dmax, dsum);
}

// ------------------------------------------------------------------------------------------------
// variant 0: direct global loads, one CTA per tile, one cell per thread per row
// ------------------------------------------------------------------------------------------------
constexpr int DBX = 16;
__global__ void __launch_bounds__(256) diff2d_direct_kernel(const __grid_constant__ Params p) {
    int i, j;
    tile_of(p.ts, blockIdx.x, i, j);
    // tiles are BX x BY as well; 256 threads = 2 row groups x 128 columns, 16 rows each
    const int tid = threadIdx.x;
    const int y = j * BY + (tid % BY);
    const int xb = i * BX + (tid / BY) * DBX;
    const bool e_xl = i == 0 && p.side_nb[0], e_xh = i == p.i_hi && p.side_nb[1];
    const bool e_yl = j == 0 && p.side_nb[2], e_yh = j == p.j_hi && p.side_nb[3];
    const bool edge_tile = p.has_push && (e_xl || e_xh || e_yl || e_yh);
    if (edge_tile) {
        if (tid == 0) wait_faces_2d(p.push, e_xl, e_xh, e_yl, e_yh);
        __syncthreads();
    }
    const bool vy = (y >= 1) && (y <= p.ny - 2);
    const double lam = p.dev_params ? p.dev_params[0] : p.lam, dt = p.dev_params ? p.dev_params[1] : p.dt;
    const double _dx = p.dev_params ? p.dev_params[2] : p._dx, _dy = p.dev_params ? p.dev_params[3] : p._dy;
    if (vy) {
        const int nx = p.nx, ny = p.ny;
        int x = xb < 1 ? 1 : xb;
        const int xe = min(xb + DBX, nx - 1);
        if (x < xe) {
            const double* Tc = p.T + (size_t)x * ny + y;
            double up = Tc[-ny], c = Tc[0];
            for (; x < xe; ++x) {
                const double dn = Tc[ny];
                const double v = update_cell(c, up, dn, Tc[-1], Tc[1], p.Cp[(size_t)x * ny + y], lam, dt, _dx,
                                             _dy);
                bool w = true;
                if (p.mode != 0) {
                    const bool f = in_frame(p, x, y);
                    w = (p.mode == 1) ? f : !f;
                }
                if (w) {
                    p.T2[(size_t)x * ny + y] = v;
                    if (edge_tile) push_cell_2d(p.push, nx, ny, x, y, v);
                }
                up = c;
                c = dn;
                Tc += ny;
            }
        }
    }
    if (edge_tile) {
        __syncthreads();
        if (tid == 0) signal_edges_2d(p.push, p.edge_tiles, e_xl, e_xh, e_yl, e_yh);
    }
}

// ------------------------------------------------------------------------------------------------
// host: tile lists
// ------------------------------------------------------------------------------------------------
static void add_seg(TileSet& ts, int i0, int i1, int j0, int j1, int band = 0) {   // inclusive tile ranges
    if (i1 < i0 || j1 < j0) return;
    Seg& s = ts.s[ts.nseg++];
    s.band = band;
    s.i0 = i0;
    s.j0 = j0;
    s.h = i1 - i0 + 1;
    s.w = j1 - j0 + 1;
    s.start = ts.ntiles;
    ts.ntiles += s.w * s.h;
}

static int env_int(const char* name, int dflt) {
    const char* e = getenv(name);
    return e ? atoi(e) : dflt;
}

// tuning knobs (see scripts/setenv.sh); read once
static int tma_rows_per_thread() {
    static int v = -1;
    if (v < 0) v = env_int("IGG_TMA_RPT", 4);
    return v;
}
static int tile_band() {
    static int v = -1;
    if (v < 0) {
        v = env_int("IGG_TILE_BAND", 0);
        if (v < 0) v = 0;
    }
    return v;
}

static void build_tiles(TileSet& ts, int nx, int ny, int mode, int bwx, int bwy, const int BX = ::BX,
                        const int BY = ::BY) {
    ts.nseg = 0;
    ts.ntiles = 0;
    const int band = tile_band();
    const int nti = (nx - 2) / BX + 1, ntj = (ny - 2) / BY + 1;
    if (mode == 2) {
        const int x0 = 1 + bwx, x1 = nx - 2 - bwx, y0 = 1 + bwy, y1 = ny - 2 - bwy;
        if (x1 < x0 || y1 < y0) return;
        add_seg(ts, x0 / BX, x1 / BX, y0 / BY, y1 / BY, band);
        return;
    }
    // tile rows / columns that intersect the boundary frame
    int a0 = (bwx < nx - 2 ? bwx : nx - 2) / BX;                 // last row of the low band
    int a1 = (nx - 1 - bwx > 1 ? nx - 1 - bwx : 1) / BX;         // first row of the high band
    int b0 = (bwy < ny - 2 ? bwy : ny - 2) / BY;
    int b1 = (ny - 1 - bwy > 1 ? ny - 1 - bwy : 1) / BY;
    if (a1 <= a0) {  // bands meet: everything is frame
        add_seg(ts, 0, nti - 1, 0, ntj - 1, band);
        return;
    }
    add_seg(ts, 0, a0, 0, ntj - 1);
    add_seg(ts, a1, nti - 1, 0, ntj - 1);
    if (b1 <= b0) {
        add_seg(ts, a0 + 1, a1 - 1, 0, ntj - 1);
        return;
    }
    add_seg(ts, a0 + 1, a1 - 1, 0, b0);
    add_seg(ts, a0 + 1, a1 - 1, b1, ntj - 1);
    if (mode == 0) add_seg(ts, a0 + 1, a1 - 1, b0 + 1, b1 - 1, band);
}

template <int RPT>
static int launch_tma(const CUtensorMap& mT, const CUtensorMap& mC, const Params& p, cudaStream_t st) {
    auto kern = diff2d_tma_kernel<RPT>;
    static bool attr_set[64] = {false};
    int dev = 0;
    IGG_CUDA_OK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !attr_set[dev]) {
        IGG_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
        if (dev >= 0 && dev < 64) attr_set[dev] = true;
    }
    const int sms = igg::sm_count();
    const int grid = p.ts.ntiles < sms ? p.ts.ntiles : sms;
    kern<<<grid, (BY / 2) * (BX / RPT) + 32, SMEM_BYTES, st>>>(mT, mC, p);
    return 0;
}

static int launch_strip(const CUtensorMap& mT, const CUtensorMap& mRow, const CUtensorMap& mC, const Params& p,
                        cudaStream_t st) {
    auto kern = diff2d_strip_kernel<4>;
    static bool attr_set[64] = {false};
    int dev = 0;
    IGG_CUDA_OK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !attr_set[dev]) {
        IGG_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SSMEM_BYTES));
        if (dev >= 0 && dev < 64) attr_set[dev] = true;
    }
    const int sms = igg::sm_count();
    const int grid = p.ts.ntiles < sms ? p.ts.ntiles : sms;
    kern<<<grid, (BY / 2) * (SRB / 4) + 32, SSMEM_BYTES, st>>>(mT, mRow, mC, p);
    return 0;
}

// which TMA kernel serves whole-interior steps: 1 = strip marching (default), 0 = tiles
static int use_strip_kernel() {
    static int v = -1;
    if (v < 0) v = env_int("IGG_TMA_STRIP", 1);
    return v;
}

}  // namespace

extern "C" int igg_diffusion2d_step(const double* T, double* T2, const double* Cp, int nx, int ny, double lam,
                                    double dt, double _dx, double _dy, const igg_step_opts* opts, int mode, int bwx,
                                    int bwy, int variant, const igg_push2d* push, void* stream) {
    IGG_REQUIRE(T && T2 && Cp, "igg_diffusion2d_step: null field pointer");
    IGG_REQUIRE(nx >= 3 && ny >= 3, "igg_diffusion2d_step: need nx, ny >= 3 (got %d x %d)", nx, ny);
    IGG_REQUIRE(mode >= 0 && mode <= 2, "igg_diffusion2d_step: mode must be 0, 1 or 2");
    IGG_REQUIRE(variant == 0 || variant == 1, "igg_diffusion2d_step: variant must be 0 or 1");
    IGG_REQUIRE(mode == 0 || (bwx >= 1 && bwy >= 1), "igg_diffusion2d_step: b_width must be >= 1 in split modes");
    IGG_REQUIRE(!(push && mode == 2), "igg_diffusion2d_step: the interior pass (mode 2) never pushes");
    if (bwx < 1) bwx = 1;
    if (bwy < 1) bwy = 1;
    Params p;
    p.T = T; p.T2 = T2; p.Cp = Cp; p.nx = nx; p.ny = ny;
    p.lam = lam; p.dt = dt; p._dx = _dx; p._dy = _dy;
    p.dev_params = opts ? opts->dev_params : nullptr;
    p.diag_out = opts ? opts->diag_out : nullptr;
    p.diag = opts ? (DiagScratch*)opts->diag_scratch : nullptr;
    IGG_REQUIRE(!p.diag_out || (p.diag && variant == 1),
                "igg_diffusion2d_step: diag_out needs diag_scratch and the TMA variant (variant 1)");
    p.mode = mode; p.bwx = bwx; p.bwy = bwy;
    // whole-interior (mode 0) and inside-the-frame (mode 2) steps of large grids go to the strip-marching
    // kernel (work items of scx x BY cells); the thin boundary frame (mode 1) stays with the tile kernel
    static int scx_env = -1;
    if (scx_env < 0) {
        scx_env = env_int("IGG_STRIP_ROWS", SCX_DEFAULT);
        if (scx_env < SRB || scx_env % SRB) scx_env = SCX_DEFAULT;
    }
    p.scx = scx_env;
    const bool strip = variant == 1 && (mode == 0 || mode == 2) && use_strip_kernel() && nx >= 2 * SCX_DEFAULT;
    const int tile_x = strip ? p.scx : BX;
    p.i_hi = (nx - 2) / tile_x;
    p.j_hi = (ny - 2) / BY;
    build_tiles(p.ts, nx, ny, mode, bwx, bwy, tile_x, BY);
    if (p.ts.ntiles == 0) return 0;
    p.has_push = push ? 1 : 0;
    const int nti = p.i_hi + 1, ntj = p.j_hi + 1;
    for (int sx = -1; sx <= 1; ++sx)
        for (int sy = -1; sy <= 1; ++sy)
            p.edge_tiles[(sx + 1) * 3 + (sy + 1)] = (sx != 0 && sy != 0) ? 1 : (sx != 0 ? ntj : nti);
    for (int k = 0; k < 4; ++k) p.side_nb[k] = 0;
    if (push) {
        p.push = *push;
        IGG_REQUIRE(push->epoch >= 1, "igg_diffusion2d_step: push epoch must be >= 1");
        for (int sx = -1; sx <= 1; ++sx)
            for (int sy = -1; sy <= 1; ++sy) {
                if (!push->peer_dst[(sx + 1) * 3 + (sy + 1)]) continue;
                if (sx == -1) p.side_nb[0] = 1;
                if (sx == 1) p.side_nb[1] = 1;
                if (sy == -1) p.side_nb[2] = 1;
                if (sy == 1) p.side_nb[3] = 1;
            }
    } else {
        memset(&p.push, 0, sizeof(p.push));
    }
    cudaStream_t st = (cudaStream_t)stream;
    if (variant == 1) {
        if ((ny & 1) || (((uintptr_t)T | (uintptr_t)T2 | (uintptr_t)Cp) & 15)) {
            igg::set_error("igg_diffusion2d_step: the TMA variant needs even ny and 16-byte aligned fields "
                           "(ny=%d); use variant 0", ny);
            return IGG_ERR_UNSUPPORTED;
        }
        CUtensorMap mT, mC;
        const uint64_t dims[2] = {(uint64_t)ny, (uint64_t)nx};
        const uint32_t boxT[2] = {TP, BX + 2};
        const uint32_t boxC[2] = {BY, BX};
        int rc = igg::get_tensor_map_f64(T, 2, dims, boxT, &mT);
        if (rc) return rc;
        rc = igg::get_tensor_map_f64(Cp, 2, dims, boxC, &mC);
        if (rc) return rc;
        if (strip) {
            CUtensorMap mB, mRow;
            const uint32_t boxB[2] = {TP, SRB};
            const uint32_t boxR[2] = {TP, 1};
            rc = igg::get_tensor_map_f64(T, 2, dims, boxB, &mB);
            if (rc) return rc;
            rc = igg::get_tensor_map_f64(T, 2, dims, boxR, &mRow);
            if (rc) return rc;
            rc = launch_strip(mB, mRow, mC, p, st);
        } else {
            switch (tma_rows_per_thread()) {
                case 4: rc = launch_tma<4>(mT, mC, p, st); break;    // 16 consumer warps
                case 16: rc = launch_tma<16>(mT, mC, p, st); break;  // 4 consumer warps
                default: rc = launch_tma<8>(mT, mC, p, st); break;   // 8 consumer warps
            }
        }
        if (rc) return rc;
        IGG_LAUNCH_CHECK();
    } else {
        diff2d_direct_kernel<<<p.ts.ntiles, 256, 0, st>>>(p);
        IGG_LAUNCH_CHECK();
    }
    return 0;
}

// ================================================================================================
// Persistent multi-step kernel for latency-bound (small) local grids -- beyond the reference.
//
// The reference pays, per time step, three kernel launches with a host wait after each
// (scripts/diffusion_2D_kp.jl:88-90) or one launch + a blocking update_halo! (scripts/diffusion_2D_perf.jl:49-51);
// at 128 x 128 cells per GPU that is pure launch / synchronisation latency.  Here ONE cooperative launch
// runs `nsteps` steps: every CTA owns a fixed set of tiles, edge cells are pushed to the neighbours'
// halo cells and the epoch released exactly as in the per-step kernels, and the steps are separated by
// a grid barrier (a monotonic counter in global memory).  Fields stay L2-resident; reads of the field
// use ld.global.cg because L1 is not coherent with what other SMs (or peers) wrote in the previous step.
// ================================================================================================
namespace {

struct MSParams {
    Params p;                   // step 0 reads p.T and writes p.T2; push.peer_dst = neighbours' T2
    double* peer_dst_odd[9];    // odd steps write T: neighbours' T
    int nsteps;
    int form;                   // 0 = fused perf form (perf.jl:8-10), 1 = kp form (kp.jl:20-51, divides by Cp)
    unsigned long long* gbar;   // grid-barrier counter (monotonic)
    unsigned long long gbar_base;
};

constexpr int MBX = 8;          // tile rows of the multi-step kernel: small tiles = many CTAs on a small grid
constexpr int MROWS = MBX / 2;  // rows per thread (256 threads = 2 row groups x 128 columns)

__global__ void __launch_bounds__(256) diff2d_multistep_kernel(const __grid_constant__ MSParams q) {
    const Params& p = q.p;
    const int tid = threadIdx.x;
    const int nx = p.nx, ny = p.ny;
    for (int step = 0; step < q.nsteps; ++step) {
        const bool even = (step & 1) == 0;
        const double* __restrict__ src = even ? p.T : p.T2;
        double* __restrict__ dst = even ? p.T2 : const_cast<double*>(p.T);
        double* const* peers = even ? p.push.peer_dst : q.peer_dst_odd;
        const uint64_t epoch = p.push.epoch + (uint64_t)step;
        for (int t = blockIdx.x; t < p.ts.ntiles; t += gridDim.x) {
            int i, j;
            tile_of(p.ts, t, i, j);
            const int y = j * BY + (tid % BY);
            const int xb = i * MBX + (tid / BY) * MROWS;
            const bool e_xl = i == 0 && p.side_nb[0], e_xh = i == p.i_hi && p.side_nb[1];
            const bool e_yl = j == 0 && p.side_nb[2], e_yh = j == p.j_hi && p.side_nb[3];
            const bool edge_tile = p.has_push && (e_xl || e_xh || e_yl || e_yh);
            if (edge_tile) {
                if (tid == 0) wait_faces_2d_epoch(p.push, e_xl, e_xh, e_yl, e_yh, epoch - 1);
                __syncthreads();
            }
            if (y >= 1 && y <= ny - 2) {
                // all loads of this thread are independent: one L2 round trip per step, not one per row
                double col[MROWS + 2], lf[MROWS], rt[MROWS], cpv[MROWS];
#pragma unroll
                for (int r = 0; r < MROWS + 2; ++r) {
                    int x = xb - 1 + r;
                    x = x < 0 ? 0 : (x > nx - 1 ? nx - 1 : x);
                    col[r] = __ldcg(src + (size_t)x * ny + y);
                }
#pragma unroll
                for (int r = 0; r < MROWS; ++r) {
                    int x = xb + r;
                    x = x > nx - 1 ? nx - 1 : x;
                    lf[r] = __ldcg(src + (size_t)x * ny + y - 1);
                    rt[r] = __ldcg(src + (size_t)x * ny + y + 1);
                    cpv[r] = p.Cp[(size_t)x * ny + y];
                }
#pragma unroll
                for (int r = 0; r < MROWS; ++r) {
                    const int x = xb + r;
                    if (x >= 1 && x <= nx - 2) {
                        const double v = q.form ? update_cell_kp(col[r + 1], col[r], col[r + 2], lf[r], rt[r], cpv[r],
                                                                 p.lam, p.dt, p._dx, p._dy)
                                                : update_cell(col[r + 1], col[r], col[r + 2], lf[r], rt[r], cpv[r], p.lam,
                                                              p.dt, p._dx, p._dy);
                        dst[(size_t)x * ny + y] = v;
                        if (edge_tile) push_cell_2d_to(peers, nx, ny, x, y, v);
                    }
                }
            }
            if (edge_tile) {
                __syncthreads();
                if (tid == 0) signal_edges_2d_epoch(p.push, p.edge_tiles, e_xl, e_xh, e_yl, e_yh, epoch);
            }
        }
        // grid barrier: everybody has finished reading `src` and writing `dst`
        __syncthreads();
        if (tid == 0) {
            __threadfence();
            atomicAdd(q.gbar, 1ull);
            const unsigned long long want = q.gbar_base + (unsigned long long)(step + 1) * gridDim.x;
            while (*((volatile unsigned long long*)q.gbar) < want) {
            }
            __threadfence();
        }
        __syncthreads();
    }
}

}  // namespace

extern "C" int igg_diffusion2d_multistep(double* T, double* T2, const double* Cp, int nx, int ny, double lam, double dt,
                                         double _dx, double _dy, int nsteps, const igg_push2d* push_even,
                                         double* const* peer_dst_odd, unsigned long long* grid_barrier,
                                         unsigned long long grid_barrier_base, int* grid_size_out, void* stream) {
    return igg_diffusion2d_multistep_ex(T, T2, Cp, nx, ny, lam, dt, _dx, _dy, nsteps, 0, push_even, peer_dst_odd,
                                        grid_barrier, grid_barrier_base, grid_size_out, stream);
}

extern "C" int igg_diffusion2d_multistep_ex(double* T, double* T2, const double* Cp, int nx, int ny, double lam,
                                            double dt, double _dx, double _dy, int nsteps, int form,
                                            const igg_push2d* push_even, double* const* peer_dst_odd,
                                            unsigned long long* grid_barrier, unsigned long long grid_barrier_base,
                                            int* grid_size_out, void* stream) {
    IGG_REQUIRE(T && T2 && Cp && grid_barrier, "igg_diffusion2d_multistep: null pointer");
    IGG_REQUIRE(form == 0 || form == 1, "igg_diffusion2d_multistep: form must be 0 (perf) or 1 (kp)");
    IGG_REQUIRE(nx >= 3 && ny >= 3 && nsteps >= 1, "igg_diffusion2d_multistep: bad sizes");
    IGG_REQUIRE(!push_even || peer_dst_odd, "igg_diffusion2d_multistep: a push needs both destination tables");
    MSParams q;
    Params& p = q.p;
    memset(&q, 0, sizeof(q));
    p.T = T; p.T2 = T2; p.Cp = Cp; p.nx = nx; p.ny = ny;
    p.lam = lam; p.dt = dt; p._dx = _dx; p._dy = _dy;
    p.mode = 0; p.bwx = 1; p.bwy = 1;
    p.i_hi = (nx - 2) / MBX;
    p.j_hi = (ny - 2) / BY;
    build_tiles(p.ts, nx, ny, 0, 1, 1, MBX, BY);
    p.has_push = push_even ? 1 : 0;
    const int nti = p.i_hi + 1, ntj = p.j_hi + 1;
    for (int sx = -1; sx <= 1; ++sx)
        for (int sy = -1; sy <= 1; ++sy)
            p.edge_tiles[(sx + 1) * 3 + (sy + 1)] = (sx != 0 && sy != 0) ? 1 : (sx != 0 ? ntj : nti);
    if (push_even) {
        p.push = *push_even;
        IGG_REQUIRE(push_even->epoch >= 1, "igg_diffusion2d_multistep: push epoch must be >= 1");
        for (int o = 0; o < 9; ++o) {
            q.peer_dst_odd[o] = peer_dst_odd[o];
            if (!push_even->peer_dst[o]) continue;
            IGG_REQUIRE(peer_dst_odd[o] != nullptr, "igg_diffusion2d_multistep: destination tables disagree");
            const int sx = o / 3 - 1, sy = o % 3 - 1;
            if (sx == -1) p.side_nb[0] = 1;
            if (sx == 1) p.side_nb[1] = 1;
            if (sy == -1) p.side_nb[2] = 1;
            if (sy == 1) p.side_nb[3] = 1;
        }
    }
    q.nsteps = nsteps;
    q.form = form;
    q.gbar = grid_barrier;
    q.gbar_base = grid_barrier_base;
    // all CTAs must be co-resident (they wait for one another): cooperative launch, grid <= what fits
    int dev = 0, coop = 0, per_sm = 0;
    IGG_CUDA_OK(cudaGetDevice(&dev));
    IGG_CUDA_OK(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev));
    IGG_REQUIRE(coop, "igg_diffusion2d_multistep: device does not support cooperative launches");
    IGG_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, diff2d_multistep_kernel, 256, 0));
    int grid = igg::sm_count() * (per_sm > 0 ? 1 : 0);
    IGG_REQUIRE(grid > 0, "igg_diffusion2d_multistep: kernel does not fit on an SM");
    if (grid > p.ts.ntiles) grid = p.ts.ntiles;
    if (grid_size_out) *grid_size_out = grid;
    void* args[] = {(void*)&q};
    IGG_CUDA_OK(cudaLaunchCooperativeKernel((const void*)diff2d_multistep_kernel, dim3(grid), dim3(256), args, 0,
                                            (cudaStream_t)stream));
    IGG_LAUNCH_CHECK();
    return 0;
}

// Native time loop: `nsteps` launches of igg_diffusion2d_step with the buffers (and push descriptors)
// alternating, issued back to back from C++ so that small grids are not bound by per-step interpreter
// overhead (the reference's loop: scripts/diffusion_2D_perf.jl:47-52 without the host wait).
extern "C" int igg_diffusion2d_run(double* T, double* T2, const double* Cp, int nx, int ny, double lam, double dt,
                                   double _dx, double _dy, int bwx, int bwy, int variant, int nsteps,
                                   const igg_push2d* push_even, const igg_push2d* push_odd, void* stream) {
    IGG_REQUIRE(nsteps >= 0, "igg_diffusion2d_run: nsteps must be >= 0");
    IGG_REQUIRE((push_even == nullptr) == (push_odd == nullptr), "igg_diffusion2d_run: give both descriptors or none");
    igg_push2d pe, po;
    if (push_even) {
        pe = *push_even;
        po = *push_odd;
    }
    for (int s = 0; s < nsteps; ++s) {
        const bool even = (s & 1) == 0;
        igg_push2d* ps = nullptr;
        if (push_even) {
            ps = even ? &pe : &po;
            ps->epoch = push_even->epoch + (uint64_t)s;
        }
        const int rc = igg_diffusion2d_step(even ? T : T2, even ? T2 : T, Cp, nx, ny, lam, dt, _dx, _dy, nullptr, 0,
                                            bwx, bwy, variant, ps, stream);
        if (rc) return rc;
    }
    return 0;
}

// One end-to-end controlled step, entirely from native code: parameters travel host(pinned) -> device, the
// fused step runs (reading them from device memory and producing the in-kernel {max, sum} diagnostic), the
// diagnostic travels device -> host(pinned).  This is what a host control loop (adaptive time stepping,
// monitoring) executes per step.  Two forms:
//   igg_diffusion2d_step_controlled_submit   queues the three operations and records `done_event`; the caller
//                                            waits for it (igg_event_wait) whenever it needs the diagnostic.  A
//                                            loop that submits step k+1 BEFORE it waits for step k keeps the GPU
//                                            busy (the reference's loop never blocks the device either:
//                                            scripts/diffusion_2D_perf.jl:47-52); it needs two pinned slots.
//   igg_diffusion2d_step_controlled          submit + wait (the host sees step k before step k+1 is queued).
extern "C" int igg_event_create(void** event_out) {
    IGG_REQUIRE(event_out, "igg_event_create: null pointer");
    cudaEvent_t ev = nullptr;
    IGG_CUDA_OK(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    *event_out = (void*)ev;
    return 0;
}

extern "C" int igg_event_destroy(void* event) {
    if (event) IGG_CUDA_OK(cudaEventDestroy((cudaEvent_t)event));
    return 0;
}

// Polling keeps the host thread on the CPU (no blocking-sync wake-up latency), which matters when the wait is
// made once per ~1 ms step; spin = 0 blocks in cudaEventSynchronize instead.
extern "C" int igg_event_wait(void* event, int spin) {
    IGG_REQUIRE(event, "igg_event_wait: null event");
    cudaEvent_t ev = (cudaEvent_t)event;
    if (spin) {
        cudaError_t q;
        while ((q = cudaEventQuery(ev)) == cudaErrorNotReady) {
        }
        IGG_CUDA_OK(q);
    } else {
        IGG_CUDA_OK(cudaEventSynchronize(ev));
    }
    return 0;
}

extern "C" int igg_diffusion2d_step_controlled_submit(const double* T, double* T2, const double* Cp, int nx, int ny,
                                                      const double* params_host_pinned, double* params_dev,
                                                      double* diag_dev, void* diag_scratch, double* diag_host_pinned,
                                                      int bwx, int bwy, const igg_push2d* push, void* done_event,
                                                      void* stream) {
    IGG_REQUIRE(params_host_pinned && params_dev && diag_dev && diag_scratch && diag_host_pinned && done_event,
                "igg_diffusion2d_step_controlled_submit: null pointer");
    cudaStream_t st = (cudaStream_t)stream;
    IGG_CUDA_OK(cudaMemcpyAsync(params_dev, params_host_pinned, 4 * sizeof(double), cudaMemcpyHostToDevice, st));
    igg_step_opts o;
    o.dev_params = params_dev;
    o.diag_out = diag_dev;
    o.diag_scratch = diag_scratch;
    const int rc = igg_diffusion2d_step(T, T2, Cp, nx, ny, 0.0, 0.0, 0.0, 0.0, &o, 0, bwx, bwy, 1, push, stream);
    if (rc) return rc;
    IGG_CUDA_OK(cudaMemcpyAsync(diag_host_pinned, diag_dev, 2 * sizeof(double), cudaMemcpyDeviceToHost, st));
    IGG_CUDA_OK(cudaEventRecord((cudaEvent_t)done_event, st));
    return 0;
}

extern "C" int igg_diffusion2d_step_controlled(const double* T, double* T2, const double* Cp, int nx, int ny,
                                               const double* params_host_pinned, double* params_dev,
                                               double* diag_dev, void* diag_scratch, double* diag_host_pinned,
                                               int bwx, int bwy, const igg_push2d* push, void* stream) {
    static thread_local cudaEvent_t ev = nullptr;
    if (!ev) IGG_CUDA_OK(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    const int rc = igg_diffusion2d_step_controlled_submit(T, T2, Cp, nx, ny, params_host_pinned, params_dev, diag_dev,
                                                          diag_scratch, diag_host_pinned, bwx, bwy, push, (void*)ev,
                                                          stream);
    if (rc) return rc;
    static int spin = -1;
    if (spin < 0) spin = env_int("IGG_E2E_SPIN", 1);
    return igg_event_wait((void*)ev, spin);
}
