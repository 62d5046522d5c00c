// 2-D heat diffusion, TWO time steps per pass over memory, for sm_100a.
//
// The reference does one step per launch and per trip through HBM (scripts/diffusion_2D_perf.jl:47-52); its
// metric T_eff = 3*nx*ny*8 B / t_it (scripts/diffusion_2D_perf.jl:55-58) is therefore capped by the memory
// bandwidth.  This kernel keeps step k+1 in shared memory and writes only step k+2, so t_it per step drops
// towards half a memory pass.  No counterpart in the reference.
//
// Structure (same skeleton as the strip kernel in stencil2d.cu): one persistent CTA per SM; a producer warp
// streams 16-row x 128-column blocks of T and of Cp through mbarrier rings with TMA; the consumer warps
//   step 1  block b of T  ->  block b of the mid ring          (shared -> shared)
//   bar.sync among consumers
//   step 2  mid blocks b-1, b  ->  T2 rows one behind the block (shared -> global, streaming stores)
// The index arithmetic of both steps lives in temporal2d_core.h and is executed unchanged by the host
// emulator that the CPU tests run.
//
// Multi-GPU (halo width 1 is kept, so step k+1 of the neighbours' edge cells is needed half way):
//   igg_diffusion2d_midlines   small kernel: waits for the neighbours' pushes of the previous pass, computes
//                              step k+1 on this rank's four send lines and stores them into the neighbours'
//                              mid-line buffers, then releases the mid flag of this pass;
//   igg_diffusion2d_step2      the pass: items that read a mid line first acquire that neighbour's mid flag;
//                              step-k+2 edge cells are pushed into the neighbours' halo cells and the DATA
//                              flag of the epoch is released exactly as in the one-step kernel (push.cuh).
// The two kernels of a rank run in stream order; waits only ever target OTHER ranks' kernels.
#include <stdlib.h>
#include <string.h>

#include "push.cuh"
#include "temporal2d_core.h"

namespace {

using namespace igg;
using namespace igg_t2;

constexpr int t2_smem_bytes(int srb) { return RING_PAD * 8 + ring_doubles(srb) * 8 + RING_PAD * 8 + 128; }  // pad | rings | pad | mbarriers

// block partials + ticket of the optional diagnostic {max, sum} of step k+1 and of step k+2
struct DiagScratch2 {
    unsigned int ticket;
    unsigned int pad[3];
    double part[4][1024];          // [max1, sum1, max2, sum2][CTA]
};

struct T2Params {
    double* T2;
    Geom g;
    Phys ph;
    const double* dev_params;      // optional {lam, dt, _dx, _dy} x 2 (step k+1, step k+2) in device memory
    double* diag_out;              // optional {max1, sum1, max2, sum2}
    DiagScratch2* diag;
    MidLines ml;
    const uint64_t* mid_flag[4];   // my mid flags, one per line (xlo, xhi, ylo, yhi); NULL = no neighbour
    int has_push;
    int lean;                      // 1: inner items take the lean path (IGG_T2_LEAN=0 switches it off for A/B runs)
    int yprefetch;                 // 1: mid-line values of a y-halo column are requested one block ahead (IGG_T2_YPREFETCH)
    int edge_tiles[9];
    int side_nb[4];
    igg_push2d push;
};


// ---- lean path for items whose whole box lies inside the updated region (all but the items on the array edge) ----
// Same arithmetic and the same ring / barrier protocol as the generic path below (step1_rows / step2_rows of
// temporal2d_core.h), but written against 32-bit shared-memory addresses that are advanced incrementally: a
// thread's row and column offsets never change, so per block only the three slot bases move.  The generic path
// spends ~430 instructions per thread and block of which 100 are FP64 (ncu, profiles/ncu_temporal2_r2a.md); this one
// has no boundary rules, no pushes, no per-row pointer selection and no divisions.
__device__ __forceinline__ d2 lds128(uint32_t a) {
    d2 v;
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "r"(a));
    return v;
}
__device__ __forceinline__ double lds64(uint32_t a) {
    double v;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(a));
    return v;
}
__device__ __forceinline__ void sts128(uint32_t a, double x, double y) {
    asm volatile("st.shared.v2.f64 [%0], {%1, %2};" ::"r"(a), "d"(x), "d"(y) : "memory");
}
__device__ __forceinline__ void mbar_wait_u32(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@!p bra WAIT_%=;\n\t}"
        ::"r"(bar), "r"(parity)
        : "memory");
}
__device__ __forceinline__ void mbar_arrive_u32(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}

// Diagnostics {max1, sum1, max2, sum2}: a thread accumulates over ONE work item in registers; at the end of the item
// the warp folds its lanes and lane 0 merges the result into the warp's own slot in shared memory (no other warp
// touches it).  Nothing diagnostic is carried in registers from item to item.
struct T2Acc {
    double* slot;         // [4] doubles of this warp, stride T2_ACC_STRIDE
};
constexpr int T2_ACC_STRIDE = 32;
__device__ __forceinline__ void t2_flush_diag(const T2Acc& acc, double m1, double s1, double m2, double s2) {
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        m1 = igg::max_fast(m1, __shfl_xor_sync(0xffffffffu, m1, off));
        s1 += __shfl_xor_sync(0xffffffffu, s1, off);
        m2 = igg::max_fast(m2, __shfl_xor_sync(0xffffffffu, m2, off));
        s2 += __shfl_xor_sync(0xffffffffu, s2, off);
    }
    if ((threadIdx.x & 31) == 0) {
        acc.slot[0 * T2_ACC_STRIDE] = igg::max_fast(acc.slot[0 * T2_ACC_STRIDE], m1);
        acc.slot[1 * T2_ACC_STRIDE] += s1;
        acc.slot[2 * T2_ACC_STRIDE] = igg::max_fast(acc.slot[2 * T2_ACC_STRIDE], m2);
        acc.slot[3 * T2_ACC_STRIDE] += s2;
    }
}

// End of a kernel with diagnostics: every warp's slot -> CTA partial -> scratch; the last CTA (ticket) folds all
// partials in a fixed order, restores a NaN that the hot loops' compare-select max dropped (it is still in the sum) and
// re-arms the ticket.  THREADS = the threads that take part (the consumers: the whole CTA in the self-fed kernel).
template <int THREADS>
__device__ __forceinline__ void t2_diag_finish(DiagScratch2* scratch, double* diag_out, const T2Acc& acc) {
    constexpr int NCW = THREADS / 32;
    const int tid = threadIdx.x;
    __shared__ double red[4][32];
    __shared__ bool is_last;
    __syncwarp();
    const bool l0 = (tid & 31) == 0;
    double v[4] = {l0 ? acc.slot[0 * T2_ACC_STRIDE] : -INFINITY, l0 ? acc.slot[1 * T2_ACC_STRIDE] : 0.0,
                   l0 ? acc.slot[2 * T2_ACC_STRIDE] : -INFINITY, l0 ? acc.slot[3 * T2_ACC_STRIDE] : 0.0};
    auto block_fold = [&]() {
#pragma unroll
        for (int q = 0; q < 4; ++q)
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) {
                const double o = __shfl_xor_sync(0xffffffffu, v[q], off);
                v[q] = (q & 1) ? v[q] + o : igg::max_nan(v[q], o);
            }
        if ((tid & 31) == 0)
            for (int q = 0; q < 4; ++q) red[q][tid >> 5] = v[q];
        consumer_sync(THREADS);
        if (tid < 32) {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                v[q] = tid < NCW ? red[q][tid] : ((q & 1) ? 0.0 : -INFINITY);
#pragma unroll
                for (int off = 16; off > 0; off >>= 1) {
                    const double o = __shfl_xor_sync(0xffffffffu, v[q], off);
                    v[q] = (q & 1) ? v[q] + o : igg::max_nan(v[q], o);
                }
            }
        }
        consumer_sync(THREADS);
    };
    block_fold();
    if (tid == 0) {
        for (int q = 0; q < 4; ++q) scratch->part[q][blockIdx.x] = v[q];
        __threadfence();
        is_last = atomicAdd(&scratch->ticket, 1u) == gridDim.x - 1;
    }
    consumer_sync(THREADS);
    if (is_last) {
        __threadfence();
        for (int q = 0; q < 4; ++q) v[q] = (q & 1) ? 0.0 : -INFINITY;
        for (unsigned int b2 = tid; b2 < gridDim.x; b2 += THREADS)
            for (int q = 0; q < 4; ++q) {
                const double o = scratch->part[q][b2];
                v[q] = (q & 1) ? v[q] + o : igg::max_nan(v[q], o);
            }
        block_fold();
        if (tid == 0) {
            if (v[1] != v[1]) v[0] = v[1];
            if (v[3] != v[3]) v[2] = v[3];
            for (int q = 0; q < 4; ++q) diag_out[q] = v[q];
            scratch->ticket = 0;
        }
    }
}

// Step 1 of one block for one thread: mid rows rg*RPT ... rg*RPT+RPT-1.  EDGE = the block may contain rows outside the
// item's own output range [x_lo, x_hi] (first block and the last two): only then are rows tested one by one.
// ---- self-fed pipeline ---------------------------------------------------------------------------------------
// Variant without a producer warp: thread 0 of the CTA issues the TMA loads itself, right after the per-block
// barrier.  That barrier already orders every slot re-use (T block q-1 is last read in step 1 of block q; the Cp row
// that step 2 of block q needs from block q-1 is carried in registers), so the "empty" mbarriers and the producer's
// polling disappear, and a CTA is four warps: five CTAs per SM are 20 warps = 5 per scheduler, i.e. up to 102
// registers per thread, where 4 + 1 warps per CTA capped a thread at 72.  At barrier(q): T block q+3 -> slot of
// q-1, Cp block q+2 -> slot of q-1.
__device__ __forceinline__ void tma_load_2d_u32(uint32_t dst, const CUtensorMap* m, uint32_t bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4}], [%2];"
        ::"r"(dst), "l"(m), "r"(bar), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void mbar_expect_tx_u32(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}

struct T2Feed {                    // position of the next block to request, per ring (used by thread 0 only)
    const CUtensorMap* mapT;
    const CUtensorMap* mapC;
    const T2Params* p;
    int first, stride, n_my;
    int kT, bT, kC, bC;
    Item itT, itC;
    unsigned qT, qC;
    uint32_t sT, sC, fullT, fullC;
};
// request the next T block of this CTA's stream (no-op at the end of the stream)
template <int SRB>
__device__ __forceinline__ void t2_feed_T(T2Feed& f) {
    if (f.kT >= f.n_my) return;
    const T2Params& p = *f.p;
    if (f.bT == 0) {
        f.itT = make_item(p.g, f.first + f.kT * f.stride);
        if (p.has_push) {       // mid lines this item may read: acquire the neighbours' mid flag of this pass
            const Item& it = f.itT;
            const int x_last = it.xin0 + it.nbi * SRB - 1;
            if (it.xin0 <= 0 && p.mid_flag[0]) wait_flag_ge(p.mid_flag[0], p.push.epoch, p.push.err_word);
            if (x_last >= p.g.nx - 1 && p.mid_flag[1]) wait_flag_ge(p.mid_flag[1], p.push.epoch, p.push.err_word);
            if (it.yb <= 0 && p.mid_flag[2]) wait_flag_ge(p.mid_flag[2], p.push.epoch, p.push.err_word);
            if (it.yb + BW - 1 >= p.g.ny - 1 && p.mid_flag[3]) wait_flag_ge(p.mid_flag[3], p.push.epoch, p.push.err_word);
        }
    }
    constexpr uint32_t BLKB = SRB * BW * 8;
    const uint32_t slot = f.qT % SNT;
    mbar_expect_tx_u32(f.fullT + slot * 8, BLKB);
    tma_load_2d_u32(f.sT + slot * BLKB, f.mapT, f.fullT + slot * 8, f.itT.yb, f.itT.xin0 + f.bT * SRB);
    ++f.qT;
    if (++f.bT == f.itT.nbi) {
        f.bT = 0;
        ++f.kT;
    }
}
template <int SRB>
__device__ __forceinline__ void t2_feed_C(T2Feed& f) {
    if (f.kC >= f.n_my) return;
    if (f.bC == 0) f.itC = make_item(f.p->g, f.first + f.kC * f.stride);
    constexpr uint32_t BLKB = SRB * BW * 8;
    const uint32_t slot = f.qC % SNC;
    mbar_expect_tx_u32(f.fullC + slot * 8, BLKB);
    tma_load_2d_u32(f.sC + slot * BLKB, f.mapC, f.fullC + slot * 8, f.itC.yb, f.itC.xin0 + f.bC * SRB);
    ++f.qC;
    if (++f.bC == f.itC.nbi) {
        f.bC = 0;
        ++f.kC;
    }
}

// What an item on the array edge needs on top of an inner one (INNER = false): the boundary rules of step 1 (global
// boundary keeps its value, halo cells take the neighbours' mid lines), per-cell column validity, and the push of the
// new edge cells to the neighbours.
struct T2Border {
    int nx, ny;
    int y;                 // global column of the thread's first cell
    bool vy0, vy1;         // are the thread's two columns inside [1, ny-2]
    bool pair_ok;          // not one of the box's two garbage column pairs
    bool edge_tile;        // does this item push
    bool e_xl, e_xh, e_yl, e_yh;
    const MidLines* ml;
    const igg_push2d* push;
    const int* edge_tiles;
    // One of the thread's two columns may be a y-halo column with a neighbour (y == 0: cell A, y == ny-1: cell B): its
    // step-(k+1) value comes from the neighbour's mid line, one global load per row.  Loaded when needed it would
    // stall the thread -- and, through the per-block barrier, its whole CTA -- for a memory latency in every block
    // (measured with a y-periodic self-neighbour: +18 us per pass), so the values of block b+1 are requested
    // during block b and carried in registers.
    const double* yline;   // ml->ylo or ml->yhi for this thread, or NULL
    bool yline_B;          // the halo column is the thread's second cell
};

template <int SRB, int RPT, bool DIAG, bool EDGE, bool INNER>
__device__ __forceinline__ void t2_lean_step1(uint32_t pa, uint32_t pb, uint32_t plast, uint32_t pc, uint32_t pm,
                                              const Phys& ph1, int xr0, int x_lo, int x_hi, double& imax, double& isum,
                                              const T2Border& bd, const double* ymid = nullptr) {
    constexpr uint32_t ROWB = BW * 8;
    d2 a = lds128(pa);
    d2 bb = lds128(pb);
#pragma unroll
    for (int r = 0; r < RPT; ++r) {
        const d2 c = lds128(r + 1 < RPT ? pb + (r + 1) * ROWB : plast);
        const double bl = lds64(pb + r * ROWB - 8);
        const double br = lds64(pb + r * ROWB + 16);
        const d2 cc = lds128(pc + r * ROWB);
        double mA = update_cell(bb.x, a.x, c.x, bl, bb.y, cc.x, ph1);
        double mB = update_cell(bb.y, a.y, c.y, bb.x, br, cc.y, ph1);
        if (!INNER) {
            const int xg = xr0 + r;
            const bool vx = xg >= 1 && xg <= bd.nx - 2;
            if (vx && bd.yline) {                 // y-halo column with a neighbour: the prefetched mid-line value
                if (bd.yline_B) mB = ymid[r]; else mA = ymid[r];
                if (bd.yline_B) { if (!bd.vy0) mA = boundary_mid(*bd.ml, xg, bd.y, bd.nx, bd.ny, bb.x); }
                else            { if (!bd.vy1) mB = boundary_mid(*bd.ml, xg, bd.y + 1, bd.nx, bd.ny, bb.y); }
            } else {
                if (!(vx && bd.vy0)) mA = boundary_mid(*bd.ml, xg, bd.y, bd.nx, bd.ny, bb.x);
                if (!(vx && bd.vy1)) mB = boundary_mid(*bd.ml, xg, bd.y + 1, bd.nx, bd.ny, bb.y);
            }
        }
        sts128(pm + r * ROWB, mA, mB);
        if (DIAG) {
            const bool row_ok = !EDGE || (xr0 + r >= x_lo && xr0 + r <= x_hi);
            if (INNER) {
                if (row_ok) {
                    imax = igg::max_fast(imax, igg::max_fast(mA, mB));
                    isum += mA + mB;
                }
            } else if (row_ok && bd.pair_ok) {
                if (bd.vy0) { imax = igg::max_fast(imax, mA); isum += mA; }
                if (bd.vy1) { imax = igg::max_fast(imax, mB); isum += mB; }
            }
        }
        a = bb;
        bb = c;
    }
}

// Step 2 of one block for one thread: output rows (xb - 1 + rg*RPT + r) from the mid rows around them.
template <int SRB, int RPT, bool DIAG, bool EDGE, bool INNER>
__device__ __forceinline__ void t2_lean_step2(uint32_t mm2, uint32_t mm1, uint32_t m0, uint32_t cm1, uint32_t c0,
                                              const Phys& ph2, double* __restrict__ out, int ny, bool pair_ok, int x,
                                              int x_lo, int x_hi, double& imax, double& isum, const T2Border& bd,
                                              bool use_cc0, const d2& cc0) {
    constexpr uint32_t ROWB = BW * 8;
    d2 a = lds128(mm2);
    d2 bb = lds128(mm1);
#pragma unroll
    for (int r = 0; r < RPT; ++r) {
        const uint32_t rc = r == 0 ? mm1 : m0 + (r - 1) * ROWB;        // row of the output cell
        const d2 c = lds128(m0 + r * ROWB);
        const double bl = lds64(rc - 8);
        const double br = lds64(rc + 16);
        d2 cc;
        if (r == 0 && use_cc0) cc = cc0;        // the row of the previous Cp block, carried in registers (self-fed)
        else cc = lds128(r == 0 ? cm1 : c0 + (r - 1) * ROWB);
        const double oA = update_cell(bb.x, a.x, c.x, bl, bb.y, cc.x, ph2);
        const double oB = update_cell(bb.y, a.y, c.y, bb.x, br, cc.y, ph2);
        const bool row_ok = !EDGE || (x + r >= x_lo && x + r <= x_hi);
        if (INNER) {
            if (pair_ok && row_ok) st_cs_f64x2(out + (size_t)r * ny, oA, oB);
            if (DIAG) {
                if (row_ok) {
                    imax = igg::max_fast(imax, igg::max_fast(oA, oB));
                    isum += oA + oB;
                }
            }
        } else {
            const bool wa = row_ok && pair_ok && bd.vy0, wb = row_ok && pair_ok && bd.vy1;
            double* o = out + (size_t)r * ny;
            if (wa && wb) {
                st_cs_f64x2(o, oA, oB);
            } else {
                if (wa) st_cs_f64(o, oA);
                if (wb) st_cs_f64(o + 1, oB);
            }
            if (bd.edge_tile) {
                if (wa) push_cell_2d(*bd.push, bd.nx, bd.ny, x + r, bd.y, oA);
                if (wb) push_cell_2d(*bd.push, bd.nx, bd.ny, x + r, bd.y + 1, oB);
            }
            if (DIAG) {
                if (wa) { imax = igg::max_fast(imax, oA); isum += oA; }
                if (wb) { imax = igg::max_fast(imax, oB); isum += oB; }
            }
        }
        a = bb;
        bb = c;
    }
}

template <int SRB, int RPT, bool DIAG, bool INNER, bool SELF = false>
__device__ __forceinline__ void t2_lean_item(const Item& it, double* __restrict__ T2, int ny, const Phys& ph1,
                                             const Phys& ph2, uint32_t sT, uint32_t sC, uint32_t sM, uint32_t fullT,
                                             uint32_t emptyT, uint32_t fullC, uint32_t emptyC, unsigned tq, unsigned cq,
                                             unsigned mq, int cp, int rg, int consumers, const T2Acc& acc,
                                             const T2Border& bd, T2Feed* feed = nullptr, d2* ccarry = nullptr) {
    constexpr int NRG = SRB / RPT;
    constexpr uint32_t ROWB = BW * 8, BLKB = SRB * ROWB;
    const uint32_t colB = (uint32_t)cp * 16u;
    const uint32_t myB = (uint32_t)(rg * RPT) * ROWB + colB;      // my first row, my column pair, inside a block
    const bool lane0 = (threadIdx.x & 31) == 0;
    const bool pair_ok = cp >= 1 && cp <= NPAIR - 2;
    const int x_lo = it.x0, x_hi = it.x0 + it.rows_out - 1;
    // diagnostics of this item: folded without any per-cell column test (the two garbage column pairs of the box
    // accumulate garbage) and merged into the thread's running values at the end only where the pair is valid
    double i1max = -INFINITY, i1sum = 0.0, i2max = -INFINITY, i2sum = 0.0;
    // mid-line values of my y-halo column for the rows of the current block (ymid) and of the next one (ynext)
    double ymid[RPT], ynext[RPT];
#pragma unroll
    for (int r = 0; r < RPT; ++r) ymid[r] = ynext[r] = 0.0;
    auto fetch_yline = [&](int xrow0, double (&dst)[RPT]) {
#pragma unroll
        for (int r = 0; r < RPT; ++r) {
            const int xg = xrow0 + r;
            if (xg >= 1 && xg <= bd.nx - 2) dst[r] = ld_mid(bd.yline + xg);
        }
    };
    // slots and parities of block 0
    uint32_t ts = tq % SNT, tph = (tq / SNT) & 1u;
    uint32_t cs = cq % SNC, cph = (cq / SNC) & 1u;
    uint32_t ms = mq % SNM;
    uint32_t tprev = ts, cprev = cs, mprev = (ms + SNM - 1) % SNM;
    mbar_wait_u32(fullT + ts * 8, tph);
    // Only now: the first T block of the item was requested AFTER its mid flags had been acquired, so its arrival
    // orders these loads behind the neighbours' mid-line stores (read earlier they can see the previous pass' lines:
    // found on 4 GPUs with a CTA whose first item lies on a y face).
    if (!INNER && bd.yline) fetch_yline(it.xin0 + rg * RPT, ymid);
    // output pointer of my first step-2 row of block 0 (rows xb - 1 + rg*RPT + r)
    int x = it.xin0 - 1 + rg * RPT;
    double* out = T2 + (size_t)x * ny + (it.yb + 2 * cp);
    for (int b = 0; b < it.nbi; ++b) {
        const uint32_t tn = (ts + 1) & (SNT - 1);
        const uint32_t tnph = tn == 0 ? tph ^ 1u : tph;
        const bool has_next = b + 1 < it.nbi;
        // rows outside [x_lo, x_hi] exist only in the first block and the last two (items on the array edge test
        // every row: their blocks are few)
        const bool edge = !INNER || b == 0 || b + 2 >= it.nbi;
        if (has_next) mbar_wait_u32(fullT + tn * 8, tnph);
        mbar_wait_u32(fullC + cs * 8, cph);
        const uint32_t tc = sT + ts * BLKB, tp = sT + tprev * BLKB, tnx = sT + (has_next ? tn : ts) * BLKB;
        const uint32_t cc_ = sC + cs * BLKB, cpv = sC + cprev * BLKB;
        const uint32_t mc = sM + ms * BLKB, mp = sM + mprev * BLKB;
        // ---- step 1: mid rows rg*RPT ... rg*RPT+RPT-1 of block b
        {
            const uint32_t pb = tc + myB;
            const uint32_t pa = rg == 0 ? tp + (SRB - 1) * ROWB + colB : pb - ROWB;
            const uint32_t plast = rg == NRG - 1 ? tnx + colB : pb + RPT * ROWB;
            if (!INNER && bd.yline && has_next) fetch_yline(x + 1 + SRB, ynext);     // rows of block b+1
            if (edge)
                t2_lean_step1<SRB, RPT, DIAG, true, INNER>(pa, pb, plast, cc_ + myB, mc + myB, ph1, x + 1, x_lo, x_hi, i1max, i1sum, bd, ymid);
            else
                t2_lean_step1<SRB, RPT, DIAG, false, true>(pa, pb, plast, cc_ + myB, mc + myB, ph1, x + 1, x_lo, x_hi, i1max, i1sum, bd);
        }
        if (!SELF) {
            __syncwarp();
            if (lane0 && b > 0) mbar_arrive_u32(emptyT + tprev * 8);  // T block b-1 is done
        }
        consumer_sync(consumers);                                   // mid block b complete
        if constexpr (SELF) {
            if (threadIdx.x == 0) {                                  // the barrier has freed one T and one Cp slot
                t2_feed_T<SRB>(*feed);
                t2_feed_C<SRB>(*feed);
            }
        }
        // ---- step 2: output rows xb - 1 + rg*RPT + r from mid rows (rg*RPT - 2 + r) ... (rg*RPT + r)
        {
            const uint32_t m0 = mc + myB;                                      // mid row rg*RPT of block b
            const uint32_t mm1 = rg == 0 ? mp + (SRB - 1) * ROWB + colB : m0 - ROWB;       // row rg*RPT - 1
            const uint32_t mm2 = rg == 0 ? mp + (SRB - 2) * ROWB + colB : m0 - 2 * ROWB;   // row rg*RPT - 2
            const uint32_t cm1 = rg == 0 ? cpv + (SRB - 1) * ROWB + colB : cc_ + myB - ROWB;
            const bool use_cc0 = SELF && rg == 0;
            d2 cc0;
            cc0.x = 0.0; cc0.y = 0.0;
            if (use_cc0) cc0 = *ccarry;
            if (edge)
                t2_lean_step2<SRB, RPT, DIAG, true, INNER>(mm2, mm1, m0, cm1, cc_ + myB, ph2, out, ny, pair_ok, x, x_lo, x_hi, i2max, i2sum, bd, use_cc0, cc0);
            else
                t2_lean_step2<SRB, RPT, DIAG, false, true>(mm2, mm1, m0, cm1, cc_ + myB, ph2, out, ny, pair_ok, x, x_lo, x_hi, i2max, i2sum, bd, use_cc0, cc0);
            if (use_cc0) *ccarry = lds128(cc_ + (SRB - 1) * ROWB + colB);   // last Cp row of this block, for block b+1
        }
        if (!SELF) {
            __syncwarp();
            if (lane0 && b > 0) mbar_arrive_u32(emptyC + cprev * 8);  // Cp block b-1 is done
        }
        if (!INNER) {
#pragma unroll
            for (int r = 0; r < RPT; ++r) ymid[r] = ynext[r];
        }
        // advance the rings
        tprev = ts; ts = tn; tph = tnph;
        cprev = cs; cs = cs + 1 == SNC ? 0 : cs + 1; cph = cs == 0 ? cph ^ 1u : cph;
        mprev = ms; ms = ms + 1 == SNM ? 0 : ms + 1;
        x += SRB;
        out += (size_t)SRB * ny;
    }
    if (!SELF) {
        __syncwarp();
        if (lane0) {
            mbar_arrive_u32(emptyT + tprev * 8);
            mbar_arrive_u32(emptyC + cprev * 8);
        }
    }
    if (DIAG) {
        const bool keep = !INNER || pair_ok;       // items on the edge have masked cell by cell already
        t2_flush_diag(acc, keep ? i1max : -INFINITY, keep ? i1sum : 0.0, keep ? i2max : -INFINITY, keep ? i2sum : 0.0);
    }
    if (!INNER && bd.edge_tile) {
        consumer_sync(consumers);   // every remote store of this item has been issued
        if (threadIdx.x == 0) signal_edges_2d(*bd.push, bd.edge_tiles, bd.e_xl, bd.e_xh, bd.e_yl, bd.e_yh);
    }
}

// Resident CTAs per SM.  With 4-row blocks a CTA has 5 warps (4 consumer warps + the producer): four CTAs put 5 warps
// on each of the SM's four schedulers (<= 102 registers per thread), five CTAs put 7 on one of them, which caps a
// thread at 72 registers (each scheduler owns a quarter of the register file).  The plain kernel fits (0.534 vs
// 0.547 ms per step, measured in one call); the kernel with the in-kernel diagnostics does not without spilling in
// its inner loop, so it stays at four.
#ifndef IGG_T2_CTAS_SRB4
#define IGG_T2_CTAS_SRB4 5
#endif
constexpr int t2_ctas_per_sm(int srb, bool diag) {
    return srb == 4 ? (diag ? 4 : IGG_T2_CTAS_SRB4) : (srb == 8 ? 2 : 1);
}
template <int SRB, int RPT, bool DIAG>
__global__ void __launch_bounds__(NPAIR*(SRB / RPT) + 32, t2_ctas_per_sm(SRB, DIAG))
diff2d_temporal2_kernel(const __grid_constant__ CUtensorMap mapT, const __grid_constant__ CUtensorMap mapC,
                        const __grid_constant__ T2Params p) {
    constexpr int CONSUMERS = NPAIR * (SRB / RPT);
    constexpr int NCW = CONSUMERS / 32;
    constexpr int BLK = blk_doubles(SRB), BLK_BYTES = BLK * 8;
    extern __shared__ __align__(128) unsigned char smem[];
    double* ringT = reinterpret_cast<double*>(smem) + RING_PAD;      // 128 B in: still 128-byte aligned for TMA
    double* ringC = ringT + SNT * BLK;
    double* ringM = ringC + SNC * BLK;
    uint64_t* fullT = reinterpret_cast<uint64_t*>(ringM + SNM * BLK + RING_PAD);
    uint64_t* emptyT = fullT + SNT;
    uint64_t* fullC = emptyT + SNT;
    uint64_t* emptyC = fullC + SNC;
    const int tid = threadIdx.x;
    const int first = blockIdx.x;
    const int nitems = num_items(p.g);
    const int n_my = first < nitems ? (nitems - first + gridDim.x - 1) / gridDim.x : 0;
    if (tid == 0) {
        for (int s = 0; s < SNT; ++s) { mbar_init(&fullT[s], 1); mbar_init(&emptyT[s], NCW); }
        for (int s = 0; s < SNC; ++s) { mbar_init(&fullC[s], 1); mbar_init(&emptyC[s], NCW); }
        fence_mbar_init();
    }
    __syncthreads();
    const int nx = p.g.nx, ny = p.g.ny;

    if (tid >= CONSUMERS) {
        // ===== producer ===================================================================================
        if (tid == CONSUMERS) {
            tma_prefetch_desc(&mapT);
            tma_prefetch_desc(&mapC);
            unsigned tq = 0, cq = 0;
            for (int k = 0; k < n_my; ++k) {
                const Item it = make_item(p.g, first + k * gridDim.x);
                if (p.has_push) {
                    // mid lines this item may read: acquire the neighbours' mid flag of this pass
                    const int x_last = it.xin0 + it.nbi * SRB - 1;
                    if (it.xin0 <= 0 && p.mid_flag[0]) wait_flag_ge(p.mid_flag[0], p.push.epoch, p.push.err_word);
                    if (x_last >= nx - 1 && p.mid_flag[1]) wait_flag_ge(p.mid_flag[1], p.push.epoch, p.push.err_word);
                    if (it.yb <= 0 && p.mid_flag[2]) wait_flag_ge(p.mid_flag[2], p.push.epoch, p.push.err_word);
                    if (it.yb + BW - 1 >= ny - 1 && p.mid_flag[3])
                        wait_flag_ge(p.mid_flag[3], p.push.epoch, p.push.err_word);
                }
                auto issue_T = [&](int b) {
                    const unsigned s = tq % SNT;
                    if (tq >= SNT) mbar_wait(&emptyT[s], ((tq / SNT) - 1) & 1);
                    mbar_arrive_expect_tx(&fullT[s], BLK_BYTES);
                    tma_load_2d(ringT + s * BLK, &mapT, &fullT[s], it.yb, it.xin0 + b * SRB);
                    ++tq;
                };
                auto issue_C = [&](int b) {
                    const unsigned s = cq % SNC;
                    if (cq >= SNC) mbar_wait(&emptyC[s], ((cq / SNC) - 1) & 1);
                    mbar_arrive_expect_tx(&fullC[s], BLK_BYTES);
                    tma_load_2d(ringC + s * BLK, &mapC, &fullC[s], it.yb, it.xin0 + b * SRB);
                    ++cq;
                };
                issue_T(0);
                for (int b = 0; b < it.nbi; ++b) {
                    if (b + 1 < it.nbi) issue_T(b + 1);
                    issue_C(b);
                }
            }
        }
        return;
    }

    // ===== consumers ====================================================================================
    const int cp = tid % NPAIR;   // column pair
    const int rg = tid / NPAIR;   // row group inside a block
    Phys ph1 = p.ph, ph2 = p.ph;   // step k+1 / step k+2 (a host control loop may change lam, dt per step)
    if (p.dev_params) {
        ph1.lam = p.dev_params[0]; ph1.dt = p.dev_params[1]; ph1._dx = p.dev_params[2]; ph1._dy = p.dev_params[3];
        ph2.lam = p.dev_params[4]; ph2.dt = p.dev_params[5]; ph2._dx = p.dev_params[6]; ph2._dy = p.dev_params[7];
    }
    constexpr bool diag = DIAG;
    __shared__ double wacc[DIAG ? 4 * T2_ACC_STRIDE : 1];      // [max1 | sum1 | max2 | sum2][warp]
    T2Acc acc;
    acc.slot = &wacc[DIAG ? (tid >> 5) : 0];
    if (DIAG && (tid & 31) == 0) {
        acc.slot[0 * T2_ACC_STRIDE] = -INFINITY;
        acc.slot[1 * T2_ACC_STRIDE] = 0.0;
        acc.slot[2 * T2_ACC_STRIDE] = -INFINITY;
        acc.slot[3 * T2_ACC_STRIDE] = 0.0;
    }
    __syncwarp();
    unsigned tq = 0, cq = 0, mq = 0;
    const uint32_t sT = smem_u32(ringT), sC = smem_u32(ringC), sM = smem_u32(ringM);
    const uint32_t aFullT = smem_u32(fullT), aEmptyT = smem_u32(emptyT), aFullC = smem_u32(fullC),
                   aEmptyC = smem_u32(emptyC);

    for (int k = 0; k < n_my; ++k) {
        const Item it = make_item(p.g, first + k * gridDim.x);
        const bool e_xl = it.ci == 0 && p.side_nb[0], e_xh = it.ci == p.g.nti - 1 && p.side_nb[1];
        const bool e_yl = it.j == 0 && p.side_nb[2], e_yh = it.j == p.g.ntj - 1 && p.side_nb[3];
        const bool edge_tile = p.has_push && (e_xl || e_xh || e_yl || e_yh);
        const int x_lo = it.x0, x_hi = it.x0 + it.rows_out - 1;      // inside [1, nx-2] by construction
        const bool inner = item_is_inner(p.g, it);
        if (p.lean) {
            T2Border bd;
            bd.nx = nx; bd.ny = ny;
            bd.y = it.yb + 2 * cp;
            bd.vy0 = bd.y >= 1 && bd.y <= ny - 2;
            bd.vy1 = bd.y + 1 >= 1 && bd.y + 1 <= ny - 2;
            bd.pair_ok = cp >= 1 && cp <= NPAIR - 2;
            bd.edge_tile = edge_tile;
            bd.e_xl = e_xl; bd.e_xh = e_xh; bd.e_yl = e_yl; bd.e_yh = e_yh;
            bd.ml = &p.ml;
            bd.push = &p.push;
            bd.edge_tiles = p.edge_tiles;
            bd.yline = nullptr;
            bd.yline_B = false;
            if (p.yprefetch && bd.y == 0 && p.ml.ylo) bd.yline = p.ml.ylo;
            if (p.yprefetch && bd.y + 1 == ny - 1 && p.ml.yhi) { bd.yline = p.ml.yhi; bd.yline_B = true; }
            if (inner)
                t2_lean_item<SRB, RPT, DIAG, true>(it, p.T2, ny, ph1, ph2, sT, sC, sM, aFullT, aEmptyT, aFullC, aEmptyC,
                                                   tq, cq, mq, cp, rg, CONSUMERS, acc, bd);
            else
                t2_lean_item<SRB, RPT, DIAG, false>(it, p.T2, ny, ph1, ph2, sT, sC, sM, aFullT, aEmptyT, aFullC, aEmptyC,
                                                    tq, cq, mq, cp, rg, CONSUMERS, acc, bd);
            tq += it.nbi;
            cq += it.nbi;
            mq += it.nbi;
            continue;
        }
        const int y = it.yb + 2 * cp;
        double d1max = -INFINITY, d1sum = 0.0, d2max = -INFINITY, d2sum = 0.0;      // of this item
        mbar_wait(&fullT[tq % SNT], (tq / SNT) & 1);
        for (int b = 0; b < it.nbi; ++b) {
            const unsigned qb = tq + b, qc = cq + b;
            if (b + 1 < it.nbi) mbar_wait(&fullT[(qb + 1) % SNT], ((qb + 1) / SNT) & 1);
            mbar_wait(&fullC[qc % SNC], (qc / SNC) & 1);
            const BlockPtrs bp = block_ptrs(ringT, ringC, ringM, BLK, tq, cq, mq, b, it.nbi);
            const int xb = it.xin0 + b * SRB;

            double oA[RPT], oB[RPT];
            unsigned wA, wB;
            step1_rows<SRB, RPT>(bp.prevT, bp.curT, bp.nextT, bp.curC, bp.curM, cp, rg, xb, it.yb, nx, ny, ph1, p.ml, inner, oA, oB);
            if (diag) {
                step1_own_masks<RPT>(cp, rg, xb, it.yb, x_lo, x_hi, ny, wA, wB);
#pragma unroll
                for (int r = 0; r < RPT; ++r) {
                    if ((wA >> r) & 1u) { d1max = igg::max_fast(d1max, oA[r]); d1sum += oA[r]; }
                    if ((wB >> r) & 1u) { d1max = igg::max_fast(d1max, oB[r]); d1sum += oB[r]; }
                }
            }
            __syncwarp();
            if ((tid & 31) == 0 && b > 0) mbar_arrive(&emptyT[(qb + SNT - 1) % SNT]);   // T block b-1 is done
            consumer_sync(CONSUMERS);                                                   // mid block b complete

            step2_rows<SRB, RPT>(bp.prevM, bp.curM, bp.prevC, bp.curC, cp, rg, xb, it.yb, x_lo, x_hi, ny, ph2, oA, oB, wA, wB);
#pragma unroll
            for (int r = 0; r < RPT; ++r) {
                const bool a = (wA >> r) & 1u, c = (wB >> r) & 1u;
                if (a | c) {
                    const int x = step2_row(xb, rg, RPT, r);
                    double* out = p.T2 + (size_t)x * ny + y;
                    if (a && c) {
                        st_cs_f64x2(out, oA[r], oB[r]);
                    } else {
                        if (a) st_cs_f64(out, oA[r]);
                        if (c) st_cs_f64(out + 1, oB[r]);
                    }
                    if (edge_tile) {
                        if (a) push_cell_2d(p.push, nx, ny, x, y, oA[r]);
                        if (c) push_cell_2d(p.push, nx, ny, x, y + 1, oB[r]);
                    }
                    if (diag) {
                        if (a) { d2max = igg::max_fast(d2max, oA[r]); d2sum += oA[r]; }
                        if (c) { d2max = igg::max_fast(d2max, oB[r]); d2sum += oB[r]; }
                    }
                }
            }
            __syncwarp();
            if ((tid & 31) == 0 && b > 0) mbar_arrive(&emptyC[(qc + SNC - 1) % SNC]);   // Cp block b-1 is done
        }
        __syncwarp();
        if ((tid & 31) == 0) {
            mbar_arrive(&emptyT[(tq + it.nbi - 1) % SNT]);
            mbar_arrive(&emptyC[(cq + it.nbi - 1) % SNC]);
        }
        tq += it.nbi;
        cq += it.nbi;
        mq += it.nbi;
        if (diag) t2_flush_diag(acc, d1max, d1sum, d2max, d2sum);
        if (edge_tile) {
            consumer_sync(CONSUMERS);   // every remote store of this item has been issued
            if (tid == 0) signal_edges_2d(p.push, p.edge_tiles, e_xl, e_xh, e_yl, e_yh);
        }
    }

    if (diag) t2_diag_finish<CONSUMERS>(p.diag, p.diag_out, acc);
}

// ---- the self-fed kernel ------------------------------------------------------------------------------------
// Same work items, rings, step functions and protocol as diff2d_temporal2_kernel (lean path only), without the producer
// warp: see "self-fed pipeline" above.
constexpr int t2_sf_ctas_per_sm(int srb) { return srb == 4 ? 5 : (srb == 8 ? 2 : 1); }

template <int SRB, int RPT, bool DIAG>
__global__ void __launch_bounds__(NPAIR*(SRB / RPT), t2_sf_ctas_per_sm(SRB))
diff2d_temporal2_sf_kernel(const __grid_constant__ CUtensorMap mapT, const __grid_constant__ CUtensorMap mapC,
                           const __grid_constant__ T2Params p) {
    constexpr int CONSUMERS = NPAIR * (SRB / RPT);
    constexpr int NCW = CONSUMERS / 32;
    constexpr int BLK = blk_doubles(SRB);
    extern __shared__ __align__(128) unsigned char smem[];
    double* ringT = reinterpret_cast<double*>(smem) + RING_PAD;
    double* ringC = ringT + SNT * BLK;
    double* ringM = ringC + SNC * BLK;
    uint64_t* fullT = reinterpret_cast<uint64_t*>(ringM + SNM * BLK + RING_PAD);
    uint64_t* fullC = fullT + 2 * SNT;               // same offsets as the producer kernel's layout
    const int tid = threadIdx.x;
    const int first = blockIdx.x;
    const int nitems = num_items(p.g);
    const int n_my = first < nitems ? (nitems - first + gridDim.x - 1) / gridDim.x : 0;
    if (tid == 0) {
        for (int s = 0; s < SNT; ++s) mbar_init(&fullT[s], 1);
        for (int s = 0; s < SNC; ++s) mbar_init(&fullC[s], 1);
        fence_mbar_init();
    }
    __syncthreads();
    const int nx = p.g.nx, ny = p.g.ny;
    const int cp = tid % NPAIR, rg = tid / NPAIR;
    Phys ph1 = p.ph, ph2 = p.ph;
    if (p.dev_params) {
        ph1.lam = p.dev_params[0]; ph1.dt = p.dev_params[1]; ph1._dx = p.dev_params[2]; ph1._dy = p.dev_params[3];
        ph2.lam = p.dev_params[4]; ph2.dt = p.dev_params[5]; ph2._dx = p.dev_params[6]; ph2._dy = p.dev_params[7];
    }
    __shared__ double wacc[DIAG ? 4 * T2_ACC_STRIDE : 1];
    T2Acc acc;
    acc.slot = &wacc[DIAG ? (tid >> 5) : 0];
    if (DIAG && (tid & 31) == 0) {
        acc.slot[0 * T2_ACC_STRIDE] = -INFINITY;
        acc.slot[1 * T2_ACC_STRIDE] = 0.0;
        acc.slot[2 * T2_ACC_STRIDE] = -INFINITY;
        acc.slot[3 * T2_ACC_STRIDE] = 0.0;
    }
    __syncwarp();
    const uint32_t sT = smem_u32(ringT), sC = smem_u32(ringC), sM = smem_u32(ringM);
    const uint32_t aFullT = smem_u32(fullT), aFullC = smem_u32(fullC);
    T2Feed feed;
    feed.mapT = &mapT; feed.mapC = &mapC; feed.p = &p;
    feed.first = first; feed.stride = gridDim.x; feed.n_my = n_my;
    feed.kT = feed.bT = feed.kC = feed.bC = 0;
    feed.qT = feed.qC = 0;
    feed.sT = sT; feed.sC = sC; feed.fullT = aFullT; feed.fullC = aFullC;
    if (tid == 0) {
        tma_prefetch_desc(&mapT);
        tma_prefetch_desc(&mapC);
        // prologue: T blocks 0..2 and Cp blocks 0..1 of the stream; from then on barrier(q) requests T q+3 and Cp q+2
        t2_feed_T<SRB>(feed); t2_feed_T<SRB>(feed); t2_feed_T<SRB>(feed);
        t2_feed_C<SRB>(feed); t2_feed_C<SRB>(feed);
    }
    d2 ccarry;
    ccarry.x = 0.0; ccarry.y = 0.0;
    unsigned tq = 0, cq = 0, mq = 0;
    for (int k = 0; k < n_my; ++k) {
        const Item it = make_item(p.g, first + k * gridDim.x);
        const bool e_xl = it.ci == 0 && p.side_nb[0], e_xh = it.ci == p.g.nti - 1 && p.side_nb[1];
        const bool e_yl = it.j == 0 && p.side_nb[2], e_yh = it.j == p.g.ntj - 1 && p.side_nb[3];
        T2Border bd;
        bd.nx = nx; bd.ny = ny;
        bd.y = it.yb + 2 * cp;
        bd.vy0 = bd.y >= 1 && bd.y <= ny - 2;
        bd.vy1 = bd.y + 1 >= 1 && bd.y + 1 <= ny - 2;
        bd.pair_ok = cp >= 1 && cp <= NPAIR - 2;
        bd.edge_tile = p.has_push && (e_xl || e_xh || e_yl || e_yh);
        bd.e_xl = e_xl; bd.e_xh = e_xh; bd.e_yl = e_yl; bd.e_yh = e_yh;
        bd.ml = &p.ml;
        bd.push = &p.push;
        bd.edge_tiles = p.edge_tiles;
        bd.yline = nullptr;
        bd.yline_B = false;
        if (p.yprefetch && bd.y == 0 && p.ml.ylo) bd.yline = p.ml.ylo;
        if (p.yprefetch && bd.y + 1 == ny - 1 && p.ml.yhi) { bd.yline = p.ml.yhi; bd.yline_B = true; }
        if (item_is_inner(p.g, it))
            t2_lean_item<SRB, RPT, DIAG, true, true>(it, p.T2, ny, ph1, ph2, sT, sC, sM, aFullT, 0u, aFullC, 0u, tq, cq, mq,
                                                     cp, rg, CONSUMERS, acc, bd, &feed, &ccarry);
        else
            t2_lean_item<SRB, RPT, DIAG, false, true>(it, p.T2, ny, ph1, ph2, sT, sC, sM, aFullT, 0u, aFullC, 0u, tq, cq, mq,
                                                      cp, rg, CONSUMERS, acc, bd, &feed, &ccarry);
        tq += it.nbi;
        cq += it.nbi;
        mq += it.nbi;
    }
    if (DIAG) t2_diag_finish<CONSUMERS>(p.diag, p.diag_out, acc);
}

// ---- mid lines -------------------------------------------------------------------------------------------
struct MidParams {
    const double* T;
    const double* Cp;
    int nx, ny;
    Phys ph;
    const double* dev_params;
    igg_mid2d mid;
    const uint64_t* data_flag[4];   // DATA flags of the four face neighbours (previous pass), NULL = none
    uint64_t* err_word;
    uint64_t epoch;
};

// CTAs per face: one per 256 cells of the longer edge, at most 64 (a line is pure latency -- six dependent-free loads,
// one remote store per cell -- so it is spread wide; the arrival counter is taken modulo this number, which depends
// only on the field shape and is therefore the same for every launch of a time loop)
constexpr int MID_MAX_CTAS_PER_FACE = 64;
static int mid_ctas_per_face(int nx, int ny) {
    const int n = nx > ny ? nx : ny;
    int c = (n + 255) / 256;
    return c < 1 ? 1 : (c > MID_MAX_CTAS_PER_FACE ? MID_MAX_CTAS_PER_FACE : c);
}

__global__ void __launch_bounds__(256) diff2d_midlines_kernel(const __grid_constant__ MidParams p) {
    const int face = blockIdx.y;
    if (p.mid.peer_mid[face] == nullptr) return;
    // the cells read here include halo cells of all four sides: wait for every face neighbour's pushes of
    // the previous pass
    if (threadIdx.x < 4 && p.data_flag[threadIdx.x])
        wait_flag_ge(p.data_flag[threadIdx.x], p.epoch - 1, p.err_word);
    __syncthreads();
    Phys ph = p.ph;
    if (p.dev_params) {
        ph.lam = p.dev_params[0];
        ph.dt = p.dev_params[1];
        ph._dx = p.dev_params[2];
        ph._dy = p.dev_params[3];
    }
    const int n = face_length(face, p.nx, p.ny);
    double* dst = p.mid.peer_mid[face] + mid_line_offset(mid_line_of_face(face), p.nx, p.ny);
    for (int i = 1 + blockIdx.x * blockDim.x + threadIdx.x; i <= n - 2; i += gridDim.x * blockDim.x)
        dst[i] = mid_line_value(p.T, p.Cp, p.nx, p.ny, face, i, ph);
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence_system();
        const unsigned long long old = atomicAdd((unsigned long long*)p.mid.my_count[face], 1ull);
        if ((old + 1ull) % (unsigned long long)gridDim.x == 0ull) {
            __threadfence_system();
            st_release_sys(p.mid.peer_flag[face], p.epoch);
        }
    }
}

static int env_int_(const char* name, int dflt) {
    const char* e = getenv(name);
    return e ? atoi(e) : dflt;
}

template <int SRB, int RPT, bool DIAG>
static int launch_t2_self_fed(const CUtensorMap& mT, const CUtensorMap& mC, const T2Params& p, cudaStream_t st) {
    auto kern = diff2d_temporal2_sf_kernel<SRB, RPT, DIAG>;
    constexpr int T2_SMEM_BYTES = t2_smem_bytes(SRB);
    static bool attr_set[64] = {false};
    int dev = 0;
    IGG_CUDA_OK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !attr_set[dev]) {
        IGG_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, T2_SMEM_BYTES));
        if (dev >= 0 && dev < 64) attr_set[dev] = true;
    }
    const int slots = igg::sm_count() * t2_sf_ctas_per_sm(SRB);
    const int nitems = num_items(p.g);
    const int grid = nitems < slots ? nitems : slots;
    if (grid > 1024) {
        igg::set_error("igg_diffusion2d_step2: more than 1024 CTAs (%d) do not fit the diagnostic scratch", grid);
        return IGG_ERR_UNSUPPORTED;
    }
    static int lead_on = -1;
    if (lead_on < 0) lead_on = env_int_("IGG_T2_LEAD", 1);
    T2Params q = p;
    set_lead(q.g, lead_on ? grid : 0);
    kern<<<grid, NPAIR*(SRB / RPT), T2_SMEM_BYTES, st>>>(mT, mC, q);
    return 0;
}

template <int SRB, int RPT, bool DIAG>
static int launch_t2d(const CUtensorMap& mT, const CUtensorMap& mC, const T2Params& p, cudaStream_t st) {
    // default: the variant without a producer warp (lean path, two rows per thread: 0.515 vs 0.526 ms per step in one call);
    // IGG_T2_SELF_FED=0 selects the producer-warp kernel
    static int self_fed = -1;
    if (self_fed < 0) self_fed = env_int_("IGG_T2_SELF_FED", 1);
    if (self_fed && p.lean && RPT == 2) return launch_t2_self_fed<SRB, 2, DIAG>(mT, mC, p, st);
    auto kern = diff2d_temporal2_kernel<SRB, RPT, DIAG>;
    constexpr int T2_SMEM_BYTES = t2_smem_bytes(SRB);
    static bool attr_set[64] = {false};
    int dev = 0;
    IGG_CUDA_OK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !attr_set[dev]) {
        IGG_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, T2_SMEM_BYTES));
        if (dev >= 0 && dev < 64) attr_set[dev] = true;
    }
    const int slots = igg::sm_count() * t2_ctas_per_sm(SRB, DIAG);   // small blocks: several CTAs per SM
    const int nitems = num_items(p.g);
    const int grid = nitems < slots ? nitems : slots;
    if (grid > 1024) {
        igg::set_error("igg_diffusion2d_step2: more than 1024 CTAs (%d) do not fit the diagnostic scratch", grid);
        return IGG_ERR_UNSUPPORTED;
    }
    // one interior item per CTA before the edge items (IGG_T2_LEAD=0: edge items first)
    static int lead_on = -1;
    if (lead_on < 0) lead_on = env_int_("IGG_T2_LEAD", 1);
    T2Params q = p;
    set_lead(q.g, lead_on ? grid : 0);
    kern<<<grid, NPAIR*(SRB / RPT) + 32, T2_SMEM_BYTES, st>>>(mT, mC, q);
    return 0;
}

template <int SRB, int RPT>
static int launch_t2(const CUtensorMap& mT, const CUtensorMap& mC, const T2Params& p, cudaStream_t st) {
    return p.diag_out ? launch_t2d<SRB, RPT, true>(mT, mC, p, st) : launch_t2d<SRB, RPT, false>(mT, mC, p, st);
}

}  // namespace

extern "C" int igg_temporal2d_default_rows(void) {
    static int v = -1;
    if (v < 0) {
        v = env_int_("IGG_T2_ROWS", 252);
        if (v < 12 || (v + 4) % 16) v = 252;        // (rows + 4) must be a multiple of the block height (4, 8 or 16)
    }
    return v;
}

extern "C" size_t igg_temporal2d_diag_scratch_bytes(void) { return sizeof(DiagScratch2); }

extern "C" int igg_diffusion2d_step2(const double* T, double* T2, const double* Cp, int nx, int ny, double lam,
                                     double dt, double _dx, double _dy, const igg_step_opts* opts,
                                     const igg_push2d* push, const igg_mid2d* mid, void* stream) {
    IGG_REQUIRE(T && T2 && Cp, "igg_diffusion2d_step2: null field pointer");
    IGG_REQUIRE(nx >= 3 && ny >= 4, "igg_diffusion2d_step2: need nx >= 3, ny >= 4 (got %d x %d)", nx, ny);
    IGG_REQUIRE((push == nullptr) == (mid == nullptr),
                "igg_diffusion2d_step2: push and mid descriptors come together (or both NULL)");
    if ((ny & 1) || (((uintptr_t)T | (uintptr_t)T2 | (uintptr_t)Cp) & 15)) {
        igg::set_error("igg_diffusion2d_step2: needs even ny and 16-byte aligned fields (ny=%d)", ny);
        return IGG_ERR_UNSUPPORTED;
    }
    T2Params p;
    memset(&p, 0, sizeof(p));
    p.T2 = T2;
    static int srb = -1, rpt = -1, lean = 1;
    if (srb < 0) {
        // rows per block: 4 -> four CTAs of 160 threads per SM (0.539 ms per step at 16384^2), 8 -> two of 288
        // (0.570), 16 -> one of 544 (0.612); measured in one call, profiles/r2/t2_lean_sweep.txt
        const int e = env_int_("IGG_T2_SRB", 4);
        rpt = env_int_("IGG_T2_RPT", 2) == 4 ? 4 : 2;
        lean = env_int_("IGG_T2_LEAN", 1);
        srb = e == 16 ? 16 : (e == 8 ? 8 : 4);
        if (srb == 4) rpt = 2;
    }
    p.lean = lean;
    static int yprefetch = -1;
    if (yprefetch < 0) yprefetch = env_int_("IGG_T2_YPREFETCH", 1);
    p.yprefetch = yprefetch;
    p.g = make_geom(nx, ny, srb, igg_temporal2d_default_rows(), push ? 1 : 0);
    p.ph.lam = lam; p.ph.dt = dt; p.ph._dx = _dx; p.ph._dy = _dy;
    p.dev_params = opts ? opts->dev_params : nullptr;
    p.diag_out = opts ? opts->diag_out : nullptr;
    p.diag = opts ? (DiagScratch2*)opts->diag_scratch : nullptr;
    IGG_REQUIRE(!p.diag_out || p.diag, "igg_diffusion2d_step2: diag_out needs diag_scratch");
    p.has_push = push ? 1 : 0;
    for (int sx = -1; sx <= 1; ++sx)
        for (int sy = -1; sy <= 1; ++sy)
            p.edge_tiles[(sx + 1) * 3 + (sy + 1)] = (sx != 0 && sy != 0) ? 1 : (sx != 0 ? p.g.ntj : p.g.nti);
    if (push) {
        IGG_REQUIRE(push->epoch >= 1, "igg_diffusion2d_step2: push epoch must be >= 1");
        p.push = *push;
        for (int sx = -1; sx <= 1; ++sx)
            for (int sy = -1; sy <= 1; ++sy) {
                if (!push->peer_dst[(sx + 1) * 3 + (sy + 1)]) continue;
                if (sx == -1) p.side_nb[0] = 1;
                if (sx == 1) p.side_nb[1] = 1;
                if (sy == -1) p.side_nb[2] = 1;
                if (sy == 1) p.side_nb[3] = 1;
            }
        // line l of my buffer is filled by the neighbour on that side (xlo by the x-low neighbour, ...)
        const int face_slot[4] = {0 * 3 + 1, 2 * 3 + 1, 1 * 3 + 0, 1 * 3 + 2};
        const double* base = mid->my_mid;
        IGG_REQUIRE(base != nullptr, "igg_diffusion2d_step2: mid->my_mid is NULL");
        for (int l = 0; l < 4; ++l) {
            const bool nb = push->peer_dst[face_slot[l]] != nullptr;
            p.mid_flag[l] = nb ? mid->my_flag[l] : nullptr;
            IGG_REQUIRE(!nb || mid->my_flag[l], "igg_diffusion2d_step2: mid flag missing for an existing neighbour");
        }
        p.ml.xlo = p.mid_flag[0] ? base + mid_line_offset(0, nx, ny) : nullptr;
        p.ml.xhi = p.mid_flag[1] ? base + mid_line_offset(1, nx, ny) : nullptr;
        p.ml.ylo = p.mid_flag[2] ? base + mid_line_offset(2, nx, ny) : nullptr;
        p.ml.yhi = p.mid_flag[3] ? base + mid_line_offset(3, nx, ny) : nullptr;
    }
    CUtensorMap mT, mC;
    const uint64_t dims[2] = {(uint64_t)ny, (uint64_t)nx};
    const uint32_t box[2] = {BW, (uint32_t)srb};
    int rc = igg::get_tensor_map_f64(T, 2, dims, box, &mT);
    if (rc) return rc;
    rc = igg::get_tensor_map_f64(Cp, 2, dims, box, &mC);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    if (srb == 16 && rpt == 4) rc = launch_t2<16, 4>(mT, mC, p, st);
    else if (srb == 4) rc = launch_t2<4, 2>(mT, mC, p, st);
    else if (srb == 16) rc = launch_t2<16, 2>(mT, mC, p, st);
    else if (rpt == 4) rc = launch_t2<8, 4>(mT, mC, p, st);
    else rc = launch_t2<8, 2>(mT, mC, p, st);
    if (rc) return rc;
    IGG_LAUNCH_CHECK();
    return 0;
}

extern "C" int igg_diffusion2d_midlines(const double* T, const double* Cp, int nx, int ny, double lam, double dt,
                                        double _dx, double _dy, const double* dev_params, const igg_push2d* push,
                                        const igg_mid2d* mid, void* stream) {
    IGG_REQUIRE(T && Cp && push && mid, "igg_diffusion2d_midlines: null argument");
    IGG_REQUIRE(nx >= 3 && ny >= 3, "igg_diffusion2d_midlines: need nx, ny >= 3");
    IGG_REQUIRE(push->epoch >= 1, "igg_diffusion2d_midlines: push epoch must be >= 1");
    MidParams p;
    memset(&p, 0, sizeof(p));
    p.T = T; p.Cp = Cp; p.nx = nx; p.ny = ny;
    p.ph.lam = lam; p.ph.dt = dt; p.ph._dx = _dx; p.ph._dy = _dy;
    p.dev_params = dev_params;
    p.mid = *mid;
    const int face_slot[4] = {0 * 3 + 1, 2 * 3 + 1, 1 * 3 + 0, 1 * 3 + 2};
    bool any = false;
    for (int f = 0; f < 4; ++f) {
        p.data_flag[f] = push->peer_dst[face_slot[f]] ? push->my_flag[face_slot[f]] : nullptr;
        if (mid->peer_mid[f]) {
            IGG_REQUIRE(mid->peer_flag[f] && mid->my_count[f], "igg_diffusion2d_midlines: incomplete mid descriptor");
            any = true;
        }
    }
    p.err_word = push->err_word;
    p.epoch = push->epoch;
    if (!any) return 0;
    dim3 grid(mid_ctas_per_face(nx, ny), 4);
    diff2d_midlines_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(p);
    IGG_LAUNCH_CHECK();
    return 0;
}

// One end-to-end PAIR of steps of a host control loop: H2D copy of both steps' {lam, dt, _dx, _dy} (64 B) from
// pinned host memory, [mid lines to the neighbours,] the two-step pass with both in-kernel diagnostics, D2H copy
// of {max1, sum1, max2, sum2} (32 B) into pinned host memory.  `_submit` records `done_event` and returns (wait
// with igg_event_wait; a loop that submits pass k+1 before waiting for pass k keeps the GPU busy); the plain form
// returns when the diagnostics have arrived.
extern "C" int igg_diffusion2d_step2_controlled_submit(const double* T, double* T2, const double* Cp, int nx, int ny,
                                                       const double* params_host_pinned, double* params_dev,
                                                       double* diag_dev, void* diag_scratch,
                                                       double* diag_host_pinned, const igg_push2d* push,
                                                       const igg_mid2d* mid, void* done_event, void* stream) {
    IGG_REQUIRE(params_host_pinned && params_dev && diag_dev && diag_scratch && diag_host_pinned && done_event,
                "igg_diffusion2d_step2_controlled_submit: null pointer");
    cudaStream_t st = (cudaStream_t)stream;
    IGG_CUDA_OK(cudaMemcpyAsync(params_dev, params_host_pinned, 8 * sizeof(double), cudaMemcpyHostToDevice, st));
    int rc = 0;
    if (push) {
        rc = igg_diffusion2d_midlines(T, Cp, nx, ny, 0.0, 0.0, 0.0, 0.0, params_dev, push, mid, stream);
        if (rc) return rc;
    }
    igg_step_opts o;
    o.dev_params = params_dev;
    o.diag_out = diag_dev;
    o.diag_scratch = diag_scratch;
    rc = igg_diffusion2d_step2(T, T2, Cp, nx, ny, 0.0, 0.0, 0.0, 0.0, &o, push, mid, stream);
    if (rc) return rc;
    IGG_CUDA_OK(cudaMemcpyAsync(diag_host_pinned, diag_dev, 4 * sizeof(double), cudaMemcpyDeviceToHost, st));
    IGG_CUDA_OK(cudaEventRecord((cudaEvent_t)done_event, st));
    return 0;
}

extern "C" int igg_diffusion2d_step2_controlled(const double* T, double* T2, const double* Cp, int nx, int ny,
                                                const double* params_host_pinned, double* params_dev,
                                                double* diag_dev, void* diag_scratch, double* diag_host_pinned,
                                                const igg_push2d* push, const igg_mid2d* mid, void* stream) {
    static thread_local cudaEvent_t ev = nullptr;
    if (!ev) IGG_CUDA_OK(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    const int rc = igg_diffusion2d_step2_controlled_submit(T, T2, Cp, nx, ny, params_host_pinned, params_dev, diag_dev,
                                                           diag_scratch, diag_host_pinned, push, mid, (void*)ev, stream);
    if (rc) return rc;
    return igg_event_wait((void*)ev, 1);
}
