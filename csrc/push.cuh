// Edge push + flag protocol shared by the fused stencil kernels (2-D and 3-D).
//
// Replaces the pack -> MPI Isend/Irecv -> unpack pipeline that ImplicitGlobalGrid runs inside
// update_halo! (reference call site scripts/diffusion_2D_perf.jl:51): the thread that
// computed an edge cell stores it straight into the neighbour's halo cell through a peer-mapped
// pointer (NVLink), and the last tile of an edge publishes the step's epoch to the neighbour with a
// system-scope release store.  Readers of a halo plane acquire that epoch before touching it.
//
// Offsets: neighbour (ox,oy,oz) in {-1,0,1}^3, slot = ((ox+1)*3 + (oy+1))*3 + (oz+1) in 3-D and
// (ox+1)*3 + (oy+1) in 2-D.  A cell at local index 1 along a dimension is an edge cell towards the
// low neighbour of that dimension and lands in the neighbour's index n-1; a cell at n-2 goes to the
// high neighbour's index 0 (overlap 2, halo width 1).
#pragma once
#include "common.cuh"

namespace igg {

// ---- 2-D -------------------------------------------------------------------------------------------
// `peer_dst` is the table of destination arrays to use (ps.peer_dst, or an alternate table when one
// launch alternates between two buffers).
__device__ __forceinline__ void push_cell_2d_to(double* const* peer_dst, int nx, int ny, int x, int y, double v) {
    const bool xl = (x == 1), xh = (x == nx - 2), yl = (y == 1), yh = (y == ny - 2);
    if (!(xl | xh | yl | yh)) return;
#pragma unroll
    for (int sx = -1; sx <= 1; ++sx) {
        if ((sx == -1 && !xl) || (sx == 1 && !xh)) continue;
#pragma unroll
        for (int sy = -1; sy <= 1; ++sy) {
            if ((sy == -1 && !yl) || (sy == 1 && !yh)) continue;
            if (sx == 0 && sy == 0) continue;
            double* dst = peer_dst[(sx + 1) * 3 + (sy + 1)];
            if (dst == nullptr) continue;
            const int dx = (sx == -1) ? nx - 1 : (sx == 1 ? 0 : x);
            const int dy = (sy == -1) ? ny - 1 : (sy == 1 ? 0 : y);
            dst[(size_t)dx * ny + dy] = v;
        }
    }
}

__device__ __forceinline__ void push_cell_2d(const igg_push2d& ps, int nx, int ny, int x, int y, double v) {
    push_cell_2d_to(ps.peer_dst, nx, ny, x, y, v);
}

// One thread per CTA, after a CTA-wide barrier that follows the tile's last remote store.
// (lo_x, hi_x, lo_y, hi_y): does this tile contain cells with x==1 / x==nx-2 / y==1 / y==ny-2.
__device__ __forceinline__ void signal_edges_2d_epoch(const igg_push2d& ps, const int* edge_tiles, bool lo_x,
                                                      bool hi_x, bool lo_y, bool hi_y, uint64_t epoch) {
    __threadfence_system();
#pragma unroll
    for (int sx = -1; sx <= 1; ++sx) {
        if ((sx == -1 && !lo_x) || (sx == 1 && !hi_x)) continue;
#pragma unroll
        for (int sy = -1; sy <= 1; ++sy) {
            if ((sy == -1 && !lo_y) || (sy == 1 && !hi_y)) continue;
            if (sx == 0 && sy == 0) continue;
            const int o = (sx + 1) * 3 + (sy + 1);
            if (ps.peer_flag[o] == nullptr) continue;
            const unsigned long long old = atomicAdd((unsigned long long*)ps.my_count[o], 1ull);
            if ((old + 1ull) % (unsigned long long)edge_tiles[o] == 0ull) {
                __threadfence_system();
                st_release_sys(ps.peer_flag[o], epoch);
            }
        }
    }
}

__device__ __forceinline__ bool wait_faces_2d_epoch(const igg_push2d& ps, bool lo_x, bool hi_x, bool lo_y,
                                                    bool hi_y, uint64_t want) {
    bool any = false;
    if (lo_x && ps.my_flag[0 * 3 + 1]) { wait_flag_ge(ps.my_flag[0 * 3 + 1], want, ps.err_word); any = true; }
    if (hi_x && ps.my_flag[2 * 3 + 1]) { wait_flag_ge(ps.my_flag[2 * 3 + 1], want, ps.err_word); any = true; }
    if (lo_y && ps.my_flag[1 * 3 + 0]) { wait_flag_ge(ps.my_flag[1 * 3 + 0], want, ps.err_word); any = true; }
    if (hi_y && ps.my_flag[1 * 3 + 2]) { wait_flag_ge(ps.my_flag[1 * 3 + 2], want, ps.err_word); any = true; }
    return any;
}

// the step's own epoch: what the per-step kernels use (the multi-step kernel passes the epoch of each step)
__device__ __forceinline__ void signal_edges_2d(const igg_push2d& ps, const int* edge_tiles, bool lo_x,
                                                bool hi_x, bool lo_y, bool hi_y) {
    signal_edges_2d_epoch(ps, edge_tiles, lo_x, hi_x, lo_y, hi_y, ps.epoch);
}

// Before a tile reads halo planes: acquire the previous epoch from the face neighbours it touches.
// Returns true if any flag was actually acquired.
__device__ __forceinline__ bool wait_faces_2d(const igg_push2d& ps, bool lo_x, bool hi_x, bool lo_y,
                                              bool hi_y) {
    return wait_faces_2d_epoch(ps, lo_x, hi_x, lo_y, hi_y, ps.epoch - 1);
}

// ---- 3-D -------------------------------------------------------------------------------------------
__device__ __forceinline__ void push_cell_3d(const igg_push3d& ps, int nx, int ny, int nz, int x, int y,
                                             int z, double v) {
    const bool xl = (x == 1), xh = (x == nx - 2), yl = (y == 1), yh = (y == ny - 2), zl = (z == 1),
               zh = (z == nz - 2);
    if (!(xl | xh | yl | yh | zl | zh)) return;
    for (int sx = -1; sx <= 1; ++sx) {
        if ((sx == -1 && !xl) || (sx == 1 && !xh)) continue;
        for (int sy = -1; sy <= 1; ++sy) {
            if ((sy == -1 && !yl) || (sy == 1 && !yh)) continue;
            for (int sz = -1; sz <= 1; ++sz) {
                if ((sz == -1 && !zl) || (sz == 1 && !zh)) continue;
                if (sx == 0 && sy == 0 && sz == 0) continue;
                double* dst = ps.peer_dst[((sx + 1) * 3 + (sy + 1)) * 3 + (sz + 1)];
                if (dst == nullptr) continue;
                const int dx = (sx == -1) ? nx - 1 : (sx == 1 ? 0 : x);
                const int dy = (sy == -1) ? ny - 1 : (sy == 1 ? 0 : y);
                const int dz = (sz == -1) ? nz - 1 : (sz == 1 ? 0 : z);
                dst[((size_t)dx * ny + dy) * nz + dz] = v;
            }
        }
    }
}

__device__ __forceinline__ void signal_edges_3d(const igg_push3d& ps, const int* edge_tiles, const bool lo[3],
                                                const bool hi[3]) {
    __threadfence_system();
    for (int sx = -1; sx <= 1; ++sx) {
        if ((sx == -1 && !lo[0]) || (sx == 1 && !hi[0])) continue;
        for (int sy = -1; sy <= 1; ++sy) {
            if ((sy == -1 && !lo[1]) || (sy == 1 && !hi[1])) continue;
            for (int sz = -1; sz <= 1; ++sz) {
                if ((sz == -1 && !lo[2]) || (sz == 1 && !hi[2])) continue;
                if (sx == 0 && sy == 0 && sz == 0) continue;
                const int o = ((sx + 1) * 3 + (sy + 1)) * 3 + (sz + 1);
                if (ps.peer_flag[o] == nullptr) continue;
                const unsigned long long old = atomicAdd((unsigned long long*)ps.my_count[o], 1ull);
                if ((old + 1ull) % (unsigned long long)edge_tiles[o] == 0ull) {
                    __threadfence_system();
                    st_release_sys(ps.peer_flag[o], ps.epoch);
                }
            }
        }
    }
}

__device__ __forceinline__ bool wait_faces_3d(const igg_push3d& ps, const bool lo[3], const bool hi[3]) {
    const uint64_t want = ps.epoch - 1;
    const int face_lo[3] = {(0 * 3 + 1) * 3 + 1, (1 * 3 + 0) * 3 + 1, (1 * 3 + 1) * 3 + 0};
    const int face_hi[3] = {(2 * 3 + 1) * 3 + 1, (1 * 3 + 2) * 3 + 1, (1 * 3 + 1) * 3 + 2};
    bool any = false;
    for (int d = 0; d < 3; ++d) {
        if (lo[d] && ps.my_flag[face_lo[d]]) { wait_flag_ge(ps.my_flag[face_lo[d]], want, ps.err_word); any = true; }
        if (hi[d] && ps.my_flag[face_hi[d]]) { wait_flag_ge(ps.my_flag[face_hi[d]], want, ps.err_word); any = true; }
    }
    return any;
}

}  // namespace igg
