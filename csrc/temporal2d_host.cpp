// Host emulation of the two-steps-per-pass kernel (temporal2d.cu).
//
// The kernel's index arithmetic -- work items, ring slots, the two per-thread row loops and their masks --
// lives in temporal2d_core.h and is compiled here for the CPU, so the tests can run exactly that code without
// a GPU: a "worker" plays one persistent CTA with its own ring counters, the TMA box load is a copy with zero
// fill, and the consumer barrier between step 1 and step 2 becomes two loops over the threads.
// No counterpart in the reference (it has neither native code nor a multi-step kernel).
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "../include/igg_b200.h"
#include "temporal2d_core.h"

using namespace igg_t2;

namespace {

// what cp.async.bulk.tensor does for a [srb x BW] box at (x, y): out-of-range elements read as zero
void load_box(const double* A, int nx, int ny, int x, int y, int srb, double* dst) {
    for (int r = 0; r < srb; ++r) {
        const int gx = x + r;
        for (int c = 0; c < BW; ++c) {
            const int gy = y + c;
            dst[r * BW + c] = (gx >= 0 && gx < nx && gy >= 0 && gy < ny) ? A[(size_t)gx * ny + gy] : 0.0;
        }
    }
}

template <int SRB, int RPT>
void emulate(const double* T, double* T2, const double* Cp, const Geom& g, const Phys& ph1, const Phys& ph2,
             int workers, const MidLines& ml, double* diag) {
    double d1max = -HUGE_VAL, d1sum = 0.0, d2max = -HUGE_VAL, d2sum = 0.0;
    constexpr int NRG = SRB / RPT;
    constexpr int BLK = blk_doubles(SRB), RING_DOUBLES = ring_doubles(SRB);
    const int nitems = num_items(g);
    // 16-byte aligned rings, poisoned so that a read of a slot that was never filled shows up
    // same layout as the kernel's shared memory: pad | T ring | Cp ring | mid ring | pad
    std::vector<d2> buf((RING_DOUBLES + 2 * RING_PAD) / 2);
    double* base = reinterpret_cast<double*>(buf.data());
    for (int i = 0; i < RING_DOUBLES + 2 * RING_PAD; ++i) base[i] = 1e300;
    double* ringT = base + RING_PAD;
    double* ringC = ringT + SNT * BLK;
    double* ringM = ringC + SNC * BLK;
    for (int w = 0; w < workers; ++w) {
        for (int i = 0; i < SNT * BLK; ++i) ringT[i] = 1e300;
        for (int i = 0; i < SNC * BLK; ++i) ringC[i] = 1e300;
        for (int i = 0; i < SNM * BLK; ++i) ringM[i] = 1e300;
        unsigned tq = 0, cq = 0, mq = 0;
        const int n_my = w < nitems ? (nitems - w + workers - 1) / workers : 0;
        for (int k = 0; k < n_my; ++k) {
            const Item it = make_item(g, w + k * workers);
            const int x_lo = it.x0, x_hi = it.x0 + it.rows_out - 1;
            const bool inner = item_is_inner(g, it);
            load_box(T, g.nx, g.ny, it.xin0, it.yb, SRB, ringT + (tq % SNT) * BLK);
            for (int b = 0; b < it.nbi; ++b) {
                const unsigned qb = tq + b, qc = cq + b;
                if (b + 1 < it.nbi)
                    load_box(T, g.nx, g.ny, it.xin0 + (b + 1) * SRB, it.yb, SRB, ringT + ((qb + 1) % SNT) * BLK);
                load_box(Cp, g.nx, g.ny, it.xin0 + b * SRB, it.yb, SRB, ringC + (qc % SNC) * BLK);
                const BlockPtrs bp = block_ptrs(ringT, ringC, ringM, BLK, tq, cq, mq, b, it.nbi);
                const int xb = it.xin0 + b * SRB;
                for (int rg = 0; rg < NRG; ++rg)
                    for (int cp = 0; cp < NPAIR; ++cp) {
                        double mA[RPT], mB[RPT];
                        unsigned wA, wB;
                        step1_rows<SRB, RPT>(bp.prevT, bp.curT, bp.nextT, bp.curC, bp.curM, cp, rg, xb, it.yb, g.nx, g.ny, ph1, ml, inner, mA, mB);
                        step1_own_masks<RPT>(cp, rg, xb, it.yb, x_lo, x_hi, g.ny, wA, wB);
                        for (int r = 0; r < RPT; ++r) {
                            if ((wA >> r) & 1u) { d1max = mA[r] > d1max ? mA[r] : d1max; d1sum += mA[r]; }
                            if ((wB >> r) & 1u) { d1max = mB[r] > d1max ? mB[r] : d1max; d1sum += mB[r]; }
                        }
                    }
                // ---- consumer barrier ----
                for (int rg = 0; rg < NRG; ++rg)
                    for (int cp = 0; cp < NPAIR; ++cp) {
                        double oA[RPT], oB[RPT];
                        unsigned wA, wB;
                        step2_rows<SRB, RPT>(bp.prevM, bp.curM, bp.prevC, bp.curC, cp, rg, xb, it.yb, x_lo, x_hi, g.ny, ph2, oA, oB,
                                        wA, wB);
                        const int y = it.yb + 2 * cp;
                        for (int r = 0; r < RPT; ++r) {
                            const int x = step2_row(xb, rg, RPT, r);
                            if ((wA >> r) & 1u) {
                                T2[(size_t)x * g.ny + y] = oA[r];
                                d2max = oA[r] > d2max ? oA[r] : d2max;
                                d2sum += oA[r];
                            }
                            if ((wB >> r) & 1u) {
                                T2[(size_t)x * g.ny + y + 1] = oB[r];
                                d2max = oB[r] > d2max ? oB[r] : d2max;
                                d2sum += oB[r];
                            }
                        }
                    }
            }
            tq += it.nbi;
            cq += it.nbi;
            mq += it.nbi;
        }
    }
    if (diag) {
        diag[0] = d1max;
        diag[1] = d1sum;
        diag[2] = d2max;
        diag[3] = d2sum;
    }
}

}  // namespace

extern "C" int igg_cpu_temporal2d_emulate(const double* T, double* T2, const double* Cp, int nx, int ny, double lam,
                                          double dt, double _dx, double _dy, int srb, int scx, int rpt,
                                          int workers, const double* mid_lines, int side_mask,
                                          const double* phys2, double* diag_out) {
    if (!T || !T2 || !Cp || nx < 3 || ny < 4 || (ny & 1)) return IGG_ERR_INVALID;
    if (srb != 4 && srb != 8 && srb != 16) return IGG_ERR_INVALID;
    if (scx < srb - 4 || scx < 4 || (scx + 4) % srb || workers < 1 || (rpt != 2 && rpt != 4)) return IGG_ERR_INVALID;
    if (srb == 4 && rpt != 2) return IGG_ERR_INVALID;
    if (side_mask && !mid_lines) return IGG_ERR_INVALID;
    Geom g = make_geom(nx, ny, srb, scx, side_mask ? 1 : 0);
    set_lead(g, workers);                  // as the kernel launch does: one interior item per worker first
    Phys ph;
    ph.lam = lam;
    ph.dt = dt;
    ph._dx = _dx;
    ph._dy = _dy;
    MidLines ml;
    ml.xlo = (side_mask & 1) ? mid_lines + mid_line_offset(0, nx, ny) : nullptr;
    ml.xhi = (side_mask & 2) ? mid_lines + mid_line_offset(1, nx, ny) : nullptr;
    ml.ylo = (side_mask & 4) ? mid_lines + mid_line_offset(2, nx, ny) : nullptr;
    ml.yhi = (side_mask & 8) ? mid_lines + mid_line_offset(3, nx, ny) : nullptr;
    Phys ph2 = ph;
    if (phys2) {
        ph2.lam = phys2[0];
        ph2.dt = phys2[1];
        ph2._dx = phys2[2];
        ph2._dy = phys2[3];
    }
    if (srb == 4) emulate<4, 2>(T, T2, Cp, g, ph, ph2, workers, ml, diag_out);
    else if (srb == 16 && rpt == 4) emulate<16, 4>(T, T2, Cp, g, ph, ph2, workers, ml, diag_out);
    else if (srb == 16) emulate<16, 2>(T, T2, Cp, g, ph, ph2, workers, ml, diag_out);
    else if (rpt == 4) emulate<8, 4>(T, T2, Cp, g, ph, ph2, workers, ml, diag_out);
    else emulate<8, 2>(T, T2, Cp, g, ph, ph2, workers, ml, diag_out);
    return 0;
}

extern "C" int igg_cpu_temporal2d_midlines(const double* T, const double* Cp, int nx, int ny, double lam, double dt,
                                           double _dx, double _dy, double* const* peer_mid) {
    if (!T || !Cp || !peer_mid || nx < 3 || ny < 3) return IGG_ERR_INVALID;
    Phys ph;
    ph.lam = lam;
    ph.dt = dt;
    ph._dx = _dx;
    ph._dy = _dy;
    for (int face = 0; face < 4; ++face) {
        if (!peer_mid[face]) continue;
        const int n = face_length(face, nx, ny);
        double* dst = peer_mid[face] + mid_line_offset(mid_line_of_face(face), nx, ny);
        for (int i = 1; i <= n - 2; ++i) dst[i] = mid_line_value(T, Cp, nx, ny, face, i, ph);
    }
    return 0;
}
