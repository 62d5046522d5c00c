// 2-D heat diffusion, three-kernel ("kernel programming") form for sm_100a.
//
// Behaviour of reference scripts/diffusion_2D_kp.jl: Flux! (:16-26), Residual! (:33-40) and
// Update! (:47-54).  The reference launches each with 32x8 work-groups over nx x ny work-items
// and blocks the host after every launch (:88-90); here the three kernels are queued on one
// stream without host waits, 32 lanes run along the contiguous index, and Update! can carry the
// edge push so that no separate halo exchange follows it.
#include <string.h>

#include "push.cuh"

namespace {
using namespace igg;

constexpr int KBX = 8, KBY = 32;   // block = 32 (contiguous y) x 8 (x)

__global__ void __launch_bounds__(KBX* KBY) flux2d_kernel(double* __restrict__ qx, double* __restrict__ qy,
                                                          const double* __restrict__ T, int nx, int ny,
                                                          double lam, double _dx, double _dy) {
    const int iy = blockIdx.x * KBY + threadIdx.x;
    const int ix = blockIdx.y * KBX + threadIdx.y;
    // qx: (nx-1) x (ny-2);  qx[ix][iy] = -lam*(T[ix+1][iy+1] - T[ix][iy+1])*_dx
    if (ix < nx - 1 && iy < ny - 2)
        qx[(size_t)ix * (ny - 2) + iy] = -lam * (T[(size_t)(ix + 1) * ny + iy + 1] - T[(size_t)ix * ny + iy + 1]) * _dx;
    // qy: (nx-2) x (ny-1);  qy[ix][iy] = -lam*(T[ix+1][iy+1] - T[ix+1][iy])*_dy
    if (ix < nx - 2 && iy < ny - 1)
        qy[(size_t)ix * (ny - 1) + iy] = -lam * (T[(size_t)(ix + 1) * ny + iy + 1] - T[(size_t)(ix + 1) * ny + iy]) * _dy;
}

__global__ void __launch_bounds__(KBX* KBY) residual2d_kernel(double* __restrict__ dTdt, const double* __restrict__ qx,
                                                              const double* __restrict__ qy,
                                                              const double* __restrict__ Cp, int nx, int ny,
                                                              double _dx, double _dy) {
    const int iy = blockIdx.x * KBY + threadIdx.x;
    const int ix = blockIdx.y * KBX + threadIdx.y;
    if (ix < nx - 2 && iy < ny - 2) {
        const double dqx = qx[(size_t)(ix + 1) * (ny - 2) + iy] - qx[(size_t)ix * (ny - 2) + iy];
        const double dqy = qy[(size_t)ix * (ny - 1) + iy + 1] - qy[(size_t)ix * (ny - 1) + iy];
        dTdt[(size_t)ix * (ny - 2) + iy] = 1.0 / Cp[(size_t)(ix + 1) * ny + iy + 1] * -(dqx * _dx + dqy * _dy);
    }
}

struct UpdParams {
    double* T;
    const double* dTdt;
    int nx, ny;
    double dt;
    int has_push;
    int nbx, nby;          // blocks along x / y
    int edge_tiles[9];
    igg_push2d push;
};

__global__ void __launch_bounds__(KBX* KBY) update2d_kernel(const __grid_constant__ UpdParams p) {
    const int iy = blockIdx.x * KBY + threadIdx.x;   // index into dTdt
    const int ix = blockIdx.y * KBX + threadIdx.y;
    const bool lo_x = blockIdx.y == 0, hi_x = (int)blockIdx.y == p.nbx - 1;
    const bool lo_y = blockIdx.x == 0, hi_y = (int)blockIdx.x == p.nby - 1;
    const bool edge = p.has_push && (lo_x || hi_x || lo_y || hi_y);
    if (ix < p.nx - 2 && iy < p.ny - 2) {
        const size_t k = (size_t)(ix + 1) * p.ny + iy + 1;
        const double v = p.T[k] + p.dt * p.dTdt[(size_t)ix * (p.ny - 2) + iy];
        p.T[k] = v;
        if (edge) push_cell_2d(p.push, p.nx, p.ny, ix + 1, iy + 1, v);
    }
    if (edge) {
        __syncthreads();
        if (threadIdx.x == 0 && threadIdx.y == 0) signal_edges_2d(p.push, p.edge_tiles, lo_x, hi_x, lo_y, hi_y);
    }
}

}  // namespace

extern "C" int igg_flux2d(double* qx, double* qy, const double* T, int nx, int ny, double lam, double _dx, double _dy,
                          void* stream) {
    IGG_REQUIRE(qx && qy && T && nx >= 3 && ny >= 3, "igg_flux2d: bad arguments");
    dim3 blk(KBY, KBX), grd(igg::ceil_div(ny - 1, KBY), igg::ceil_div(nx - 1, KBX));
    flux2d_kernel<<<grd, blk, 0, (cudaStream_t)stream>>>(qx, qy, T, nx, ny, lam, _dx, _dy);
    IGG_LAUNCH_CHECK();
    return 0;
}

extern "C" int igg_residual2d(double* dTdt, const double* qx, const double* qy, const double* Cp, int nx, int ny,
                              double _dx, double _dy, void* stream) {
    IGG_REQUIRE(dTdt && qx && qy && Cp && nx >= 3 && ny >= 3, "igg_residual2d: bad arguments");
    dim3 blk(KBY, KBX), grd(igg::ceil_div(ny - 2, KBY), igg::ceil_div(nx - 2, KBX));
    residual2d_kernel<<<grd, blk, 0, (cudaStream_t)stream>>>(dTdt, qx, qy, Cp, nx, ny, _dx, _dy);
    IGG_LAUNCH_CHECK();
    return 0;
}

// NOTE on the in-place update with a push: the kp form updates T in place, so the neighbour's halo
// cells of the SAME array are overwritten while the neighbour may still be reading them in its own
// Flux! of this step.  The Python layer therefore brackets kp steps with the neighbour barrier
// (igg_neighbor_barrier) before Flux!, which is still one tiny kernel instead of pack/send/recv/unpack.
extern "C" int igg_update2d(double* T, const double* dTdt, int nx, int ny, double dt, const igg_push2d* push,
                            void* stream) {
    IGG_REQUIRE(T && dTdt && nx >= 3 && ny >= 3, "igg_update2d: bad arguments");
    UpdParams p;
    p.T = T; p.dTdt = dTdt; p.nx = nx; p.ny = ny; p.dt = dt;
    p.nby = igg::ceil_div(ny - 2, KBY);
    p.nbx = igg::ceil_div(nx - 2, KBX);
    p.has_push = push ? 1 : 0;
    for (int sx = -1; sx <= 1; ++sx)
        for (int sy = -1; sy <= 1; ++sy)
            p.edge_tiles[(sx + 1) * 3 + (sy + 1)] = (sx != 0 && sy != 0) ? 1 : (sx != 0 ? p.nby : p.nbx);
    if (push) p.push = *push; else memset(&p.push, 0, sizeof(p.push));
    dim3 blk(KBY, KBX), grd(p.nby, p.nbx);
    update2d_kernel<<<grd, blk, 0, (cudaStream_t)stream>>>(p);
    IGG_LAUNCH_CHECK();
    return 0;
}
