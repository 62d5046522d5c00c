// Two time steps per pass over memory ("temporal blocking") of the 2-D diffusion step: the part that is
// shared, source line for source line, between the sm_100a kernel (temporal2d.cu) and the host emulator
// (temporal2d_host.cpp) that the CPU tests run.
//
// The reference advances one step per kernel launch and per trip through memory
// (scripts/diffusion_2D_perf.jl:47-52: diffusion_step! -> swap -> update_halo!).  Here one work item loads a
// strip of T once, computes step k+1 into shared memory and step k+2 from there, so a pair of steps costs
// the memory traffic of one.  There is no counterpart of this file in the reference.
//
// Geometry.  A work item (ci, j) produces the cells
//     x in [1 + ci*scx, 1 + ci*scx + rows_out)        rows_out = min(scx, nx-1 - x0)
//     y in [j*BYO, j*BYO + BYO)  intersected with [1, ny-2]
// from a box of BW = BYO + 4 columns starting at the EVEN column yb = j*BYO - 2 (TMA boxes must start on a
// 16-byte boundary) and the rows xin0 = x0 - 2 ... streamed in blocks of SRB rows.  Block b holds the rows
// xin0 + b*SRB ... + SRB-1.  For every block the consumers
//   step 1: compute the step-(k+1) value ("mid") of every cell of block b into the M ring (the row above the
//           block's first row is the last row of block b-1, the row below its last row the first of b+1);
//   step 2: after a barrier, compute step k+2 for the rows xb-1 ... xb+SRB-2 (one row behind, so that the
//           mid rows they need -- xb-2 ... xb+SRB-1 -- are the last two of mid block b-1 and block b).
// Box columns 0 and BW-1 and the first/last row of a strip's blocks get no valid mid value; no valid output
// cell depends on them.  Cells outside the array read as 0 (TMA zero fill) and are never part of a valid
// output's dependency cone.
//
// Mid values on the array boundary: the global boundary keeps its value (Dirichlet: mid = T); a halo cell
// whose neighbour exists takes the neighbour's step-(k+1) edge value from the "mid lines" that the neighbour
// stored into this rank's buffer before the pass (MidLines below).
#pragma once
#include <stddef.h>

#if defined(__CUDACC__)
#define T2_HD __host__ __device__ __forceinline__
#else
#define T2_HD inline
#endif

namespace igg_t2 {

constexpr int BW = 128;            // box width in doubles (1024 B per row)
constexpr int BYO = BW - 4;        // output columns per work item
constexpr int NPAIR = BW / 2;      // column pairs = consumer threads per row group
// rows per block: template parameter S of the kernels (16: one CTA of 16 warps per SM; 8: half-size rings so that
// two CTAs share an SM and one computes while the other sits in its per-block barrier); Geom.srb at run time
constexpr int SNT = 4;             // T ring: b-1, b, b+1 in use + one in flight
constexpr int SNC = 3;             // Cp ring: b-1, b in use + one in flight
constexpr int SNM = 3;             // mid ring: b-1, b in use + the one step 1 of b+1 writes
constexpr int blk_doubles(int srb) { return srb * BW; }          // doubles per block (16 rows: 16 KB)
// The rings sit back to back (T, Cp, mid) with RING_PAD doubles of slack before the first and after the last
// block: the left/right neighbour loads of box columns 0 and BW-1 are unconditional and may touch one element
// outside a block.
constexpr int RING_PAD = 16;
constexpr int ring_doubles(int srb) { return (SNT + SNC + SNM) * blk_doubles(srb); }

#if defined(__CUDACC__)
typedef double2 d2;
#else
struct alignas(16) d2 {
    double x, y;
};
#endif

T2_HD d2 ld2(const double* p) { return *reinterpret_cast<const d2*>(p); }
T2_HD void st2(double* p, double a, double b) {
    d2 v;
    v.x = a;
    v.y = b;
    *reinterpret_cast<d2*>(p) = v;
}

struct Phys {
    double lam, dt, _dx, _dy;
};

// The arithmetic of reference scripts/diffusion_2D_perf.jl:8-10, operation for operation
// (c = centre, xm/xp = row above/below, ym/yp = column left/right).
T2_HD double update_cell(double c, double xm, double xp, double ym, double yp, double cp, const Phys& ph) {
    const double qxr = -ph.lam * (xp - c) * ph._dx;
    const double qxl = -ph.lam * (c - xm) * ph._dx;
    const double qyt = -ph.lam * (yp - c) * ph._dy;
    const double qyb = -ph.lam * (c - ym) * ph._dy;
    return c + ph.dt * (cp * (-(qxr - qxl) * ph._dx - (qyt - qyb) * ph._dy));
}

// Step-(k+1) values of the halo cells that have a neighbour: four lines in this rank's memory, written by
// the neighbours.  NULL = no neighbour on that side (the halo cell is a global boundary cell).
struct MidLines {
    const double* xlo;   // [ny]  mid value of cell (0,    y)
    const double* xhi;   // [ny]  mid value of cell (nx-1, y)
    const double* ylo;   // [nx]  mid value of cell (x, 0)
    const double* yhi;   // [nx]  mid value of cell (x, ny-1)
};

T2_HD double ld_mid(const double* p) {
#if defined(__CUDA_ARCH__)
    return __ldcg(p);    // written by another GPU: never through L1
#else
    return *p;
#endif
}

// Mid value of a cell that is not updated (outside [1,nx-2] x [1,ny-2]); `keep` is its current value.
T2_HD double boundary_mid(const MidLines& ml, int x, int y, int nx, int ny, double keep) {
    const bool yin = y >= 1 && y <= ny - 2, xin = x >= 1 && x <= nx - 2;
    if (x == 0 && yin && ml.xlo) return ld_mid(ml.xlo + y);
    if (x == nx - 1 && yin && ml.xhi) return ld_mid(ml.xhi + y);
    if (y == 0 && xin && ml.ylo) return ld_mid(ml.ylo + x);
    if (y == ny - 1 && xin && ml.yhi) return ld_mid(ml.yhi + x);
    return keep;
}

// ---- work items --------------------------------------------------------------------------------------
struct Geom {
    int nx, ny;
    int srb;        // rows per block (8 or 16)
    int scx;        // output rows per strip, (scx + 4) % srb == 0
    int nti, ntj;   // strips, tile columns
    int frame_first;  // 1: items on the array edge come early (they push to the neighbours)
    int lead;         // with frame_first: this many interior items are handed out BEFORE the edge items (one per
                      // CTA), so that an edge item's wait for the neighbours' mid lines is hidden behind an item
                      // that needs nothing from anybody; 0 = edge items first
};

inline Geom make_geom(int nx, int ny, int srb, int scx, int frame_first) {
    Geom g;
    g.nx = nx;
    g.ny = ny;
    g.srb = srb;
    g.scx = scx;
    g.nti = (nx - 2 + scx - 1) / scx;
    g.ntj = (ny - 2) / BYO + 1;
    g.frame_first = frame_first;
    g.lead = 0;
    return g;
}

T2_HD int num_items(const Geom& g) { return g.nti * g.ntj; }

// item id -> (ci, j).  frame_first order: [`lead` interior items,] strip 0, last strip, tile column 0, last tile
// column, the (remaining) interior items.  `lead` must not exceed the number of interior items (set_lead clamps).
T2_HD void set_lead(Geom& g, int want) {
    const int n_int = (g.nti >= 3 && g.ntj >= 3) ? (g.nti - 2) * (g.ntj - 2) : 0;
    g.lead = (g.frame_first && want > 0) ? (want < n_int ? want : n_int) : 0;
}

T2_HD void item_of(const Geom& g, int t, int& ci, int& j) {
    if (!g.frame_first || g.nti < 3 || g.ntj < 3) {
        ci = t / g.ntj;
        j = t - ci * g.ntj;
        return;
    }
    const int w0 = g.ntj - 2;
    if (t < g.lead) {                      // leading interior items
        ci = 1 + t / w0;
        j = 1 + (t - (ci - 1) * w0);
        return;
    }
    const int nframe = 2 * g.ntj + 2 * (g.nti - 2);
    const int lead = g.lead;
    t -= lead;
    if (t >= nframe) {                     // the remaining interior items
        t = t - nframe + lead;
        ci = 1 + t / w0;
        j = 1 + (t - (ci - 1) * w0);
        return;
    }
    if (t < g.ntj) { ci = 0; j = t; return; }
    t -= g.ntj;
    if (t < g.ntj) { ci = g.nti - 1; j = t; return; }
    t -= g.ntj;
    const int mid_rows = g.nti - 2;
    if (t < mid_rows) { ci = 1 + t; j = 0; return; }
    t -= mid_rows;
    if (t < mid_rows) { ci = 1 + t; j = g.ntj - 1; return; }
    t -= mid_rows;
    const int w = g.ntj - 2;
    ci = 1 + t / w;
    j = 1 + (t - (ci - 1) * w);
}

struct Item {
    int ci, j;
    int x0, rows_out;   // first output row, number of output rows
    int xin0, yb;       // global row of block 0 row 0, global column of box column 0
    int nbi;            // blocks
};

T2_HD Item make_item(const Geom& g, int t) {
    Item it;
    item_of(g, t, it.ci, it.j);
    it.x0 = 1 + it.ci * g.scx;
    const int left = g.nx - 1 - it.x0;
    it.rows_out = left < g.scx ? left : g.scx;
    it.xin0 = it.x0 - 2;
    it.yb = it.j * BYO - 2;
    it.nbi = (it.rows_out + 4 + g.srb - 1) / g.srb;
    return it;
}

// does every cell of the item's blocks lie in the updated region [1,nx-2] x [1,ny-2]?
T2_HD bool item_is_inner(const Geom& g, const Item& it) {
    return it.xin0 >= 1 && it.xin0 + it.nbi * g.srb - 1 <= g.nx - 2 && it.yb >= 1 && it.yb + BW - 1 <= g.ny - 2;
}

// ---- ring slots of block b of an item --------------------------------------------------------------------
// tq / cq / mq count the blocks this CTA has streamed through the T / Cp / mid rings before the item.
struct BlockPtrs {
    const double *prevT, *curT, *nextT;   // T blocks b-1, b, b+1 (cur when they do not exist: never used then)
    const double *prevC, *curC;           // Cp blocks b-1, b
    const double* prevM;                  // mid block b-1 (garbage for b == 0: masked)
    double* curM;                         // mid block b, written by step 1
};

T2_HD BlockPtrs block_ptrs(double* ringT, double* ringC, double* ringM, int blk, unsigned tq, unsigned cq,
                           unsigned mq, int b, int nbi) {
    const unsigned qb = tq + b, qc = cq + b, qm = mq + b;
    BlockPtrs p;
    p.curT = ringT + (qb % SNT) * blk;
    p.prevT = b > 0 ? ringT + ((qb + SNT - 1) % SNT) * blk : p.curT;
    p.nextT = b + 1 < nbi ? ringT + ((qb + 1) % SNT) * blk : p.curT;
    p.curC = ringC + (qc % SNC) * blk;
    p.prevC = b > 0 ? ringC + ((qc + SNC - 1) % SNC) * blk : p.curC;
    p.curM = ringM + (qm % SNM) * blk;
    p.prevM = ringM + ((qm + SNM - 1) % SNM) * blk;
    return p;
}

// global row of output r of thread row-group rg in step 2 of the block whose first row is xb (one row behind)
T2_HD int step2_row(int xb, int rg, int rpt, int r) { return xb - 1 + rg * rpt + r; }

// ---- step 1: mid values of block b --------------------------------------------------------------------
// Thread (cp, rg) owns box columns 2cp, 2cp+1 and the block rows rg*RPT ... rg*RPT+RPT-1.
// prevT / nextT: blocks b-1 / b+1 (any readable block when they do not exist: the rows that would need them
// are never used).  xb = global row of the block's row 0.
// The mid values are also returned (mA/mB, row r of the thread) for the optional step-(k+1) diagnostic.
template <int SRB, int RPT>
T2_HD void step1_rows(const double* prevT, const double* curT, const double* nextT, const double* curC,
                      double* curM, int cp, int rg, int xb, int yb, int nx, int ny, const Phys& ph,
                      const MidLines& ml, bool inner, double (&oA)[RPT], double (&oB)[RPT]) {
    const int col = 2 * cp;
    const int l0 = rg * RPT;
    const double* r_m1 = l0 > 0 ? curT + (l0 - 1) * BW : prevT + (SRB - 1) * BW;
    d2 a = ld2(r_m1 + col);
    d2 bb = ld2(curT + l0 * BW + col);
    const int y = yb + col;
    const bool vy0 = y >= 1 && y <= ny - 2;
    const bool vy1 = y + 1 >= 1 && y + 1 <= ny - 2;
#pragma unroll
    for (int r = 0; r < RPT; ++r) {
        const int l = l0 + r;
        const double* rc = curT + l * BW;
        const double* rn = (l + 1 < SRB) ? rc + BW : nextT;
        const d2 c = ld2(rn + col);
        const double bl = rc[col - 1];     // cp == 0 / NPAIR-1 read one element outside the row: in-bounds thanks
        const double br = rc[col + 2];     // to RING_PAD, and only garbage columns 0 / BW-1 depend on it
        const d2 cc = ld2(curC + l * BW + col);
        const int x = xb + l;
        const bool vx = x >= 1 && x <= nx - 2;
        double mA = update_cell(bb.x, a.x, c.x, bl, bb.y, cc.x, ph);
        double mB = update_cell(bb.y, a.y, c.y, bb.x, br, cc.y, ph);
        if (!inner) {      // items whose box lies inside [1,nx-2] x [1,ny-2] (nearly all) skip the boundary rules
            if (!(vx && vy0)) mA = boundary_mid(ml, x, y, nx, ny, bb.x);
            if (!(vx && vy1)) mB = boundary_mid(ml, x, y + 1, nx, ny, bb.y);
        }
        st2(curM + l * BW + col, mA, mB);
        oA[r] = mA;
        oB[r] = mB;
        a = bb;
        bb = c;
    }
}

// Which of the thread's step-1 cells belong to the item's OWN output region (every interior cell of the array
// is in exactly one item's region, while mid values are computed redundantly near item borders): bit r of the
// masks = row r.  Used to attribute the step-(k+1) diagnostic without double counting.
template <int RPT>
T2_HD void step1_own_masks(int cp, int rg, int xb, int yb, int x_lo, int x_hi, int ny, unsigned& wA, unsigned& wB) {
    const int y = yb + 2 * cp;
    const bool pair_ok = cp >= 1 && cp <= NPAIR - 2;
    const bool vy0 = pair_ok && y >= 1 && y <= ny - 2;
    const bool vy1 = pair_ok && y + 1 >= 1 && y + 1 <= ny - 2;
    wA = 0u;
    wB = 0u;
#pragma unroll
    for (int r = 0; r < RPT; ++r) {
        const int x = xb + rg * RPT + r;
        const bool vx = x >= x_lo && x <= x_hi;
        if (vx && vy0) wA |= 1u << r;
        if (vx && vy1) wB |= 1u << r;
    }
}

// ---- step 2: outputs of the rows xb-1 ... xb+SRB-2 -------------------------------------------------------
// Thread (cp, rg) owns the rows xb - 1 + rg*RPT + r.  prevM / prevC: mid and Cp of block b-1 (any readable
// block for b == 0; those rows are masked).  [x_lo, x_hi]: the item's valid output rows (already clipped to
// [1, nx-2]).  Returns the two values per row and the write masks (bit r = row r).
template <int SRB, int RPT>
T2_HD void step2_rows(const double* prevM, const double* curM, const double* prevC, const double* curC, int cp,
                      int rg, int xb, int yb, int x_lo, int x_hi, int ny, const Phys& ph, double (&oA)[RPT],
                      double (&oB)[RPT], unsigned& wA, unsigned& wB) {
    const int col = 2 * cp;
    const int lm0 = rg * RPT - 1;    // mid row (relative to block b) of this thread's first output row
    // row l of the two-block window: l >= 0 in block b, l < 0 in block b-1
    const double* m_m1 = lm0 - 1 >= 0 ? curM + (lm0 - 1) * BW : prevM + (SRB + lm0 - 1) * BW;
    const double* m_0 = lm0 >= 0 ? curM + lm0 * BW : prevM + (SRB + lm0) * BW;
    d2 a = ld2(m_m1 + col);
    d2 bb = ld2(m_0 + col);
    const int y = yb + col;
    const bool pair_ok = cp >= 1 && cp <= NPAIR - 2;
    const bool vy0 = pair_ok && y >= 1 && y <= ny - 2;
    const bool vy1 = pair_ok && y + 1 >= 1 && y + 1 <= ny - 2;
    wA = 0u;
    wB = 0u;
#pragma unroll
    for (int r = 0; r < RPT; ++r) {
        const int lm = lm0 + r;
        const double* rc = lm >= 0 ? curM + lm * BW : prevM + (SRB + lm) * BW;
        const double* rn = curM + (lm + 1) * BW;                  // lm + 1 is in [0, SRB-1]
        const double* cr = lm >= 0 ? curC + lm * BW : prevC + (SRB + lm) * BW;
        const d2 c = ld2(rn + col);
        const double bl = rc[col - 1];
        const double br = rc[col + 2];
        const d2 cc = ld2(cr + col);
        const int x = xb + lm;
        const bool vx = x >= x_lo && x <= x_hi;
        oA[r] = update_cell(bb.x, a.x, c.x, bl, bb.y, cc.x, ph);
        oB[r] = update_cell(bb.y, a.y, c.y, bb.x, br, cc.y, ph);
        if (vx && vy0) wA |= 1u << r;
        if (vx && vy1) wB |= 1u << r;
        a = bb;
        bb = c;
    }
}

// ---- the step-(k+1) edge lines a rank sends to its neighbours before the pass ----------------------------
// face 0: my row 1      -> x-low  neighbour's xhi line      face 2: my column 1      -> y-low  neighbour's yhi
// face 1: my row nx-2   -> x-high neighbour's xlo line      face 3: my column ny-2   -> y-high neighbour's ylo
// `i` runs along the face (y for faces 0/1, x for faces 2/3) over the updated cells 1 ... n-2.
// A rank's mid-line buffer is one array of 2*ny + 2*nx doubles: [xlo | xhi | ylo | yhi].
T2_HD size_t mid_line_offset(int line /* 0 xlo, 1 xhi, 2 ylo, 3 yhi */, int nx, int ny) {
    (void)nx;
    return line == 0 ? 0 : (line == 1 ? (size_t)ny : (line == 2 ? 2 * (size_t)ny : 2 * (size_t)ny + (size_t)nx));
}
// line of the RECEIVER that face f of the sender fills
T2_HD int mid_line_of_face(int face) { return face == 0 ? 1 : (face == 1 ? 0 : (face == 2 ? 3 : 2)); }
T2_HD int face_length(int face, int nx, int ny) { return face < 2 ? ny : nx; }

T2_HD double ld_field(const double* p) {
#if defined(__CUDA_ARCH__)
    return __ldcg(p);    // halo cells are written by other GPUs: never through L1
#else
    return *p;
#endif
}

T2_HD double mid_line_value(const double* T, const double* Cp, int nx, int ny, int face, int i, const Phys& ph) {
    const int x = face == 0 ? 1 : (face == 1 ? nx - 2 : i);
    const int y = face == 2 ? 1 : (face == 3 ? ny - 2 : i);
    const size_t k = (size_t)x * ny + y;
    return update_cell(ld_field(T + k), ld_field(T + k - ny), ld_field(T + k + ny), ld_field(T + k - 1),
                       ld_field(T + k + 1), ld_field(Cp + k), ph);
}

}  // namespace igg_t2
