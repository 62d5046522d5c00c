/* igg_b200.h -- C ABI of the native library (libigg_b200.so), sm_100a only.
 *
 * Conventions
 *   - every function returns 0 on success, a positive cudaError_t value when a CUDA call failed,
 *     or a negative IGG_ERR_* code; igg_last_error() then holds a readable message (per thread);
 *   - `stream` is a cudaStream_t passed as void* (0 = default stream); all work is asynchronous
 *     with respect to the host unless stated otherwise;
 *   - fields are row-major fp64 arrays T[ix][iy]([iz]) with the LAST index contiguous; sizes are the
 *     local sizes including the one-cell halo; indices are 0-based.
 *
 * The reference (williamfgc/ROCm-MPI) contains no native code; each declaration cites the
 * reference lines whose behaviour it provides.
 */
#ifndef IGG_B200_H
#define IGG_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define IGG_ERR_INVALID (-1)      /* bad argument                                   */
#define IGG_ERR_UNSUPPORTED (-2)  /* shape/alignment not supported by this variant  */
#define IGG_ERR_TIMEOUT (-3)      /* a cross-GPU flag wait gave up                  */

#define IGG_MAX_NEIGHBORS 27      /* 3x3x3 offsets, centre unused                   */
#define IGG_FLAG_BANKS 32         /* independent flag channels per rank             */
#define IGG_FLAG_BANK_STRIDE 64   /* uint64 slots per bank                          */

/* ---- library / device ---------------------------------------------------------------------- */
int igg_version(void);
const char* igg_last_error(void);
int igg_device_info(int* sm_count, int* cc_major, int* cc_minor, size_t* smem_optin_bytes);
/* number of kernels this library has launched in this process (bench.py's "gpu_launches") */
unsigned long long igg_launch_count(void);

/* ---- 2-D diffusion, fused form (reference scripts/diffusion_2D_perf.jl:3-13,
 *      split form scripts/diffusion_2D_perf_hide.jl:15-29) -------------------------------------
 * T2[ix][iy] = T + dt*(Cp*(-(qxR-qxL)*_dx - (qyT-qyB)*_dy)) on the interior; boundary cells of T2
 * are never written.
 *   mode 0: whole interior (boundary-frame tiles are scheduled first);
 *   mode 1: only the frame of width (bwx,bwy) -- istep==1 of the reference split kernel;
 *   mode 2: only the cells inside the frame  -- istep==2.
 *   variant 0: direct global loads ("recompiled baseline" flavour);
 *   variant 1: persistent kernel, TMA-staged tiles + mbarrier pipeline (needs ny % 2 == 0).
 * `push` may be NULL (no communication).  With a push descriptor the kernel also stores every
 * freshly computed edge cell into the neighbours' halo cells over peer memory and signals them;
 * tiles that read a halo plane first wait for that neighbour's signal of the previous epoch.
 */
typedef struct igg_push2d {
    /* indexed [(ox+1)*3 + (oy+1)]: neighbour at Cartesian offset (ox,oy); NULL = no neighbour */
    double* peer_dst[9];            /* base of the neighbour's destination array (same shape)   */
    uint64_t* peer_flag[9];         /* slot in the NEIGHBOUR's flag pad that this rank signals   */
    uint64_t* my_flag[9];           /* slot in MY flag pad that the neighbour at (ox,oy) signals */
    uint64_t* my_count[9];          /* local tile-arrival counters (monotonic)                   */
    uint64_t* err_word;             /* set (bit 63) if a wait timed out                          */
    uint64_t epoch;                 /* this step's epoch; halo reads wait for epoch-1            */
} igg_push2d;

/* Optional extras of a step (pass NULL for none):
 *   dev_params   device array {lam, dt, _dx, _dy} that overrides the by-value scalars, so a host control
 *                loop can feed per-step parameters through an asynchronous H2D copy;
 *   diag_out     device array of two doubles that receives {max, sum} over the cells written by this
 *                launch (the halo-free interior in mode 0) -- the `maximum(T_v)` diagnostic of reference
 *                scripts/diffusion_2D_perf_hide.jl:115 computed inside the stencil kernel, no extra pass;
 *                variant 1 only.  `diag_scratch`: igg_reduce_scratch_bytes(0) bytes, zero-filled once. */
typedef struct igg_step_opts {
    const double* dev_params;
    double* diag_out;
    void* diag_scratch;
} igg_step_opts;
int igg_diffusion2d_step(const double* T, double* T2, const double* Cp, int nx, int ny, double lam,
                         double dt, double _dx, double _dy, const igg_step_opts* opts, int mode, int bwx,
                         int bwy, int variant, const igg_push2d* push, void* stream);

/* One end-to-end step of a host control loop: H2D copy of {lam, dt, _dx, _dy} (32 B) from pinned host
 * memory, fused step (TMA variant, mode 0) reading them from `params_dev`, in-kernel {max, sum} diagnostic,
 * D2H copy of the diagnostic (16 B) into pinned host memory, stream synchronised on return
 * (the per-step monitoring of reference scripts/diffusion_2D_perf_hide.jl:115, done every step). */
int igg_diffusion2d_step_controlled(const double* T, double* T2, const double* Cp, int nx, int ny,
                                    const double* params_host_pinned, double* params_dev, double* diag_dev,
                                    void* diag_scratch, double* diag_host_pinned, int bwx, int bwy,
                                    const igg_push2d* push, void* stream);

/* The same step without the wait: queues H2D, step and D2H on `stream` and records `done_event` (an event from
 * igg_event_create).  A control loop that submits step k+1 before it waits for step k (igg_event_wait) never
 * leaves the GPU idle -- the reference's loop does not block the device per step either
 * (scripts/diffusion_2D_perf.jl:47-52).  The caller owns the slots: `params_host_pinned` must stay untouched
 * until the event of its step has completed, so a depth-2 loop alternates two pinned parameter slots and two
 * pinned diagnostic slots; `params_dev` / `diag_dev` are stream-ordered and can be shared. */
int igg_diffusion2d_step_controlled_submit(const double* T, double* T2, const double* Cp, int nx, int ny,
                                           const double* params_host_pinned, double* params_dev, double* diag_dev,
                                           void* diag_scratch, double* diag_host_pinned, int bwx, int bwy,
                                           const igg_push2d* push, void* done_event, void* stream);
/* Completion events for the submit forms (cudaEvent_t without timing).  igg_event_wait: spin != 0 polls
 * (lowest latency, one host core busy), spin == 0 blocks. */
int igg_event_create(void** event_out);
int igg_event_destroy(void* event);
int igg_event_wait(void* event, int spin);

/* `nsteps` back-to-back launches of igg_diffusion2d_step (mode 0) from a native loop; step s reads T
 * (s even) or T2 (s odd).  `push_even` pushes into the neighbours' T2 and carries the epoch of step 0,
 * `push_odd` pushes into the neighbours' T; both NULL = no communication.
 * (reference loop: scripts/diffusion_2D_perf.jl:47-52) */
int igg_diffusion2d_run(double* T, double* T2, const double* Cp, int nx, int ny, double lam, double dt, double _dx,
                        double _dy, int bwx, int bwy, int variant, int nsteps, const igg_push2d* push_even,
                        const igg_push2d* push_odd, void* stream);

/* `nsteps` fused steps in ONE cooperative launch, for local grids so small that a step is pure launch
 * latency (reference loop: scripts/diffusion_2D_kp.jl:87-92 at 128 x 128; no reference counterpart for the
 * single-launch form).  Step s reads T (s even) or T2 (s odd) and writes the other; after the call the
 * result is in T2 if nsteps is odd, else in T.  `push_even` (may be NULL) describes the pushes of even
 * steps (destinations: the neighbours' T2) with `epoch` = epoch of the first step; `peer_dst_odd[9]` are
 * the destinations of odd steps (the neighbours' T).  `grid_barrier`: device counter private to this rank,
 * `grid_barrier_base` its current value; the launch adds nsteps * (*grid_size_out) to it. */
int igg_diffusion2d_multistep(double* T, double* T2, const double* Cp, int nx, int ny, double lam, double dt,
                              double _dx, double _dy, int nsteps, const igg_push2d* push_even,
                              double* const* peer_dst_odd, unsigned long long* grid_barrier,
                              unsigned long long grid_barrier_base, int* grid_size_out, void* stream);
/* The same with the step arithmetic selectable: form 0 = fused perf form (scripts/diffusion_2D_perf.jl:8-10), form 1 =
 * the kp application's three kernels folded into one expression with their roundings kept (Flux! / Residual! /
 * Update!, scripts/diffusion_2D_kp.jl:20-51: rounded flux products, DIVISION by Cp) -- the fast path of the kp app. */
int igg_diffusion2d_multistep_ex(double* T, double* T2, const double* Cp, int nx, int ny, double lam, double dt,
                                 double _dx, double _dy, int nsteps, int form, const igg_push2d* push_even,
                                 double* const* peer_dst_odd, unsigned long long* grid_barrier,
                                 unsigned long long grid_barrier_base, int* grid_size_out, void* stream);

/* `nsteps` steps with the whole local field resident in the shared memory of ONE thread-block cluster
 * (<= 16 CTAs): rows shared between CTAs are read through distributed shared memory, steps are separated by
 * cluster barriers, domain-edge cells are pushed to the neighbours' global arrays and announced with per-CTA
 * epoch flags.  No reference counterpart (its small runs, scripts/diffusion_2D_kp.jl:87-92, pay 3 launches +
 * host waits + MPI per step).  IGG_ERR_UNSUPPORTED if the field does not fit (use igg_diffusion2d_multistep).
 * Result in T2 if nsteps is odd, else in T.  Tables are indexed [(ox+1)*3 + (oy+1)]; `peer_flags[o]` is the
 * base of the flag region (>= 9*16+1 uint64, zero-initialised, fabric memory) of the neighbour at offset o,
 * `my_flags` mine; `epoch0` is the epoch of step 0 and must grow by nsteps from call to call.  Pass
 * peer_flags == NULL (and peer_land == NULL) for a run without communication.
 * Low-latency mode (`peer_land` != NULL, takes precedence): instead of storing into the neighbours' arrays and
 * raising a flag after a system fence, every pushed value travels with its epoch tag as two 8-byte words into
 * a landing buffer of the neighbour (`peer_land[o]` = base of the buffer of the neighbour at offset o,
 * `my_land` mine; 2 * 9 * land_len * 2 uint64 each, zero-initialised fabric memory, land_len >= max(nx, ny)),
 * and the receiver polls each halo cell's own slot: one NVLink one-way latency per step, no fence round trip.
 * `my_flags` is still needed for the error word. */
int igg_diffusion2d_resident(double* T, double* T2, const double* Cp, int nx, int ny, double lam, double dt,
                             double _dx, double _dy, int nsteps, double* const* peer_T, double* const* peer_T2,
                             uint64_t* const* peer_flags, uint64_t* my_flags, uint64_t epoch0,
                             uint64_t* const* peer_land, uint64_t* my_land, int land_len,
                             int* cluster_size_out, void* stream);
/* form as in igg_diffusion2d_multistep_ex (1 = kp arithmetic) */
int igg_diffusion2d_resident_ex(double* T, double* T2, const double* Cp, int nx, int ny, double lam, double dt,
                                double _dx, double _dy, int nsteps, int form, double* const* peer_T,
                                double* const* peer_T2, uint64_t* const* peer_flags, uint64_t* my_flags,
                                uint64_t epoch0, uint64_t* const* peer_land, uint64_t* my_land, int land_len,
                                int* cluster_size_out, void* stream);

/* ---- 2-D diffusion, two steps per pass over memory (no reference counterpart: its loop,
 *      scripts/diffusion_2D_perf.jl:47-52, makes one trip through memory per step) -----------------
 * T2 <- two steps applied to T; step k+1 only ever exists in shared memory.  Boundary cells of T2 are never
 * written, so T2 must already hold T's boundary values (true for any T2 that started as a copy of T).
 * Needs even ny and 16-byte aligned fields (IGG_ERR_UNSUPPORTED otherwise).  `dev_params`: optional device
 * array {lam, dt, _dx, _dy}.
 * Multi-GPU: halo width stays 1, so the neighbours' step-(k+1) edge values are needed half way.  Per pass:
 *   igg_diffusion2d_midlines(T, ...)   waits for the neighbours' pushes of epoch-1, computes step k+1 on this
 *                                       rank's four send lines, stores them into the neighbours' mid-line
 *                                       buffers and releases the mid flag `epoch`;
 *   igg_diffusion2d_step2(T, T2, ...)   items that read a mid line acquire the mid flag `epoch` first; the
 *                                       step-(k+2) edge cells are pushed into the neighbours' T2 and the DATA
 *                                       flag `epoch` is released as igg_diffusion2d_step does.
 * Both take the SAME push descriptor (one epoch per pass).  Sides are indexed x-low, x-high, y-low, y-high.
 * STATUS: the benchmarked path (GPU tests: tests/test_temporal_gpu.py; multi-GPU: bench.py's verify block); the host
 * emulator igg_cpu_temporal2d_* executes the generic path's index code. */
typedef struct igg_mid2d {
    const double* my_mid;           /* my mid-line buffer, 2*ny + 2*nx doubles: [xlo | xhi | ylo | yhi]      */
    double* peer_mid[4];            /* mid-line buffer of the neighbour on side s (peer-mapped); NULL = none */
    uint64_t* peer_flag[4];         /* mid flag in the pad of the neighbour on side s that this rank releases */
    uint64_t* my_flag[4];           /* mid flag in MY pad that the neighbour on side s releases              */
    uint64_t* my_count[4];          /* private CTA arrival counters (monotonic)                              */
} igg_mid2d;
#define IGG_SLOT_MID 56             /* flag pad: mid flags of sides 0..3; counter pad: their CTA counters     */
/* `opts` (may be NULL): dev_params = device array of EIGHT doubles, {lam, dt, _dx, _dy} of step k+1 then of step
 * k+2; diag_out = device array of FOUR doubles {max, sum} of step k+1 and {max, sum} of step k+2 over the
 * halo-free interior; diag_scratch = igg_temporal2d_diag_scratch_bytes() bytes, zero-filled once. */
int igg_diffusion2d_step2(const double* T, double* T2, const double* Cp, int nx, int ny, double lam, double dt,
                          double _dx, double _dy, const igg_step_opts* opts, const igg_push2d* push,
                          const igg_mid2d* mid, void* stream);
size_t igg_temporal2d_diag_scratch_bytes(void);
/* One end-to-end PAIR of steps of a host control loop: H2D copy of both steps' parameters (64 B, pinned), mid
 * lines, the pass with both diagnostics, D2H copy of the four diagnostic values (32 B, pinned); the stream has
 * been waited for on return.  The host decides every two steps instead of every step. */
int igg_diffusion2d_step2_controlled(const double* T, double* T2, const double* Cp, int nx, int ny,
                                     const double* params_host_pinned, double* params_dev, double* diag_dev,
                                     void* diag_scratch, double* diag_host_pinned, const igg_push2d* push,
                                     const igg_mid2d* mid, void* stream);
/* Submit form of the pair (see igg_diffusion2d_step_controlled_submit): records `done_event`, does not wait. */
int igg_diffusion2d_step2_controlled_submit(const double* T, double* T2, const double* Cp, int nx, int ny,
                                            const double* params_host_pinned, double* params_dev, double* diag_dev,
                                            void* diag_scratch, double* diag_host_pinned, const igg_push2d* push,
                                            const igg_mid2d* mid, void* done_event, void* stream);
int igg_diffusion2d_midlines(const double* T, const double* Cp, int nx, int ny, double lam, double dt, double _dx,
                             double _dy, const double* dev_params, const igg_push2d* push, const igg_mid2d* mid,
                             void* stream);
/* output rows per work item of igg_diffusion2d_step2 (IGG_T2_ROWS, default 252) */
int igg_temporal2d_default_rows(void);
/* Host emulation of igg_diffusion2d_step2 that executes the kernel's own index code (csrc/temporal2d_core.h)
 * thread by thread: `workers` plays gridDim.x, `srb` the rows per block (8 or 16; IGG_T2_SRB), `rpt` the rows per
 * thread (2 or 4; IGG_T2_RPT), `scx` the rows per work item ((scx + 4) % srb == 0; IGG_T2_ROWS).  `mid_lines`: this rank's mid-line buffer or NULL; `side_mask` bit s = a neighbour
 * exists on side s.  igg_cpu_temporal2d_midlines fills the neighbours' buffers (`peer_mid[s]`, NULL = none). */
int igg_cpu_temporal2d_emulate(const double* T, double* T2, const double* Cp, int nx, int ny, double lam,
                               double dt, double _dx, double _dy, int srb, int scx, int rpt, int workers,
                               const double* mid_lines, int side_mask, const double* phys2 /* step k+2, or NULL */,
                               double* diag_out /* {max1, sum1, max2, sum2} or NULL */);
int igg_cpu_temporal2d_midlines(const double* T, const double* Cp, int nx, int ny, double lam, double dt,
                                double _dx, double _dy, double* const* peer_mid);

/* ---- 2-D diffusion, three-kernel form (reference scripts/diffusion_2D_kp.jl:16-54) ------------ */
int igg_flux2d(double* qx, double* qy, const double* T, int nx, int ny, double lam, double _dx,
               double _dy, void* stream);
int igg_residual2d(double* dTdt, const double* qx, const double* qy, const double* Cp, int nx, int ny,
                   double _dx, double _dy, void* stream);
/* Update! fused with the edge push (push may be NULL): T[inner] += dt*dTdt */
int igg_update2d(double* T, const double* dTdt, int nx, int ny, double dt, const igg_push2d* push,
                 void* stream);

/* ---- 3-D diffusion (7-point; IGG capability, the reference itself is 2-D only) ---------------- */
typedef struct igg_push3d {
    double* peer_dst[27];
    uint64_t* peer_flag[27];
    uint64_t* my_flag[27];
    uint64_t* my_count[27];
    uint64_t* err_word;
    uint64_t epoch;
} igg_push3d;
int igg_diffusion3d_step(const double* T, double* T2, const double* Cp, int nx, int ny, int nz,
                         double lam, double dt, double _dx, double _dy, double _dz, int mode, int bwx,
                         int bwy, int bwz, int variant, const igg_push3d* push, void* stream);

/* ---- initial condition (reference scripts/diffusion_2D_ap.jl:28) ------------------------------- */
int igg_gaussian_ic2d(double* T, int nx, int ny, double x0, double y0, double dx, double dy, double lx,
                      double ly, void* stream);
int igg_gaussian_ic3d(double* T, int nx, int ny, int nz, double x0, double y0, double z0, double dx,
                      double dy, double dz, double lx, double ly, double lz, void* stream);

/* ---- halo planes: pack / unpack (what IGG does inside update_halo!) ------------- */
/* A has shape (n0,n1,n2) (use 1 for unused trailing dims); plane `idx` of dimension `dim`. */
int igg_pack_plane(const double* A, int n0, int n1, int n2, int dim, int idx, double* buf, void* stream);
int igg_unpack_plane(double* A, int n0, int n1, int n2, int dim, int idx, const double* buf, void* stream);
/* interior extraction A[1:-1,...] -> dense buffer (reference scripts/diffusion_2D_ap.jl:45) */
int igg_extract_inner(const double* A, int n0, int n1, int n2, double* out, void* stream);
/* gather! without host staging (reference scripts/diffusion_2D_ap.jl:45-46: Array(T[2:end-1,2:end-1]) then gather!):
 * the halo-free interior of A is stored as the block with Cartesian coordinates (c0,c1,c2) of the global array G,
 * whose shape is (m0*d0, m1*d1, m2*d2) with m = the interior shape and (d0,d1,d2) the process grid.  G is normally a
 * peer-mapped pointer into the ROOT rank's memory: every rank's tile crosses NVLink once, as 16-byte stores. */
int igg_gather_inner(const double* A, int n0, int n1, int n2, double* G, int d0, int d1, int d2, int c0, int c1,
                     int c2, void* stream);

/* ---- peer-memory fabric (replaces GPU-aware MPI, reference scripts/rocmaware_test_selectdevice.jl:15-22)
 * Allocations are plain cudaMalloc blocks that can be exported with CUDA IPC and mapped by the
 * other ranks of the node. */
int igg_fabric_alloc(size_t bytes, void** ptr);            /* zero-initialised, 512-byte aligned   */
int igg_fabric_free(void* ptr);
int igg_ipc_export(const void* ptr, void* handle64, uint64_t* offset); /* 64-byte handle + offset   */
int igg_ipc_open(const void* handle64, void** base);       /* maps a peer allocation               */
int igg_ipc_close(void* base);
int igg_can_access_peer(int dev, int peer_dev, int* can);

/* Flag pad layout.  Every rank owns a pad of IGG_FLAG_BANKS banks x IGG_FLAG_BANK_STRIDE uint64
 * slots in fabric memory (peers signal into it) and a same-sized private pad of counters.
 * Within a bank, slots are indexed by the neighbour offset o = ((ox+1)*3+(oy+1))*3+(oz+1) (2-D code
 * uses oz = 0):
 *   [IGG_SLOT_DATA  + o]  epoch of the last completed push FROM the neighbour at offset o
 *   [IGG_SLOT_READY + o]  epoch up to which the neighbour at offset o is ready to RECEIVE
 *   [IGG_SLOT_ERR]        error word (bit 63 set = a wait timed out)                                  */
#define IGG_SLOT_DATA 0
#define IGG_SLOT_READY 27
#define IGG_SLOT_ERR 54
#define IGG_SLOT_CTA_COUNT 55     /* counter pad only: CTA arrival counter of igg_halo_exchange_peer */

/* update_halo over peer memory in ONE launch.  The dimension-sequential exchange of
 * ImplicitGlobalGrid (x, then y, then z, each with full-extent planes) is
 * algebraically a set of independent box copies to the up-to-26 neighbours, every box taken
 * from cells that the exchange itself never writes; so the kernel does: ready handshake ->
 * all boxes stored into the neighbours' arrays -> release the epoch -> wait for the neighbours'.
 * `xfers` lives in device memory. */
typedef struct igg_halo_xfer {
    const void* src;                /* my array                                                   */
    void* dst;                      /* the neighbour's array (peer-mapped; same shape)            */
    int n[3];                       /* array shape (1 for unused trailing dims)                   */
    int src_start[3];
    int dst_start[3];
    int extent[3];
    int elem_bytes;                 /* 8, 4, 2 or 1 (any field dtype of that size)                */
    int reserved;
} igg_halo_xfer;
typedef struct igg_halo_sync {
    uint64_t* peer_bank[27];        /* flag bank base in the pad of the neighbour at offset o (NULL = none) */
    uint64_t* my_bank;              /* my flag bank base                                           */
    uint64_t* my_count_bank;        /* my counter bank base                                        */
    uint64_t epoch;
} igg_halo_sync;
int igg_halo_exchange_peer(const igg_halo_xfer* xfers_dev, int nxfers, const igg_halo_sync* sync, void* stream);
/* same, with the number of CTAs chosen by the caller from the bytes it moves (0 = default); at most
 * 27 * IGG_MAX_HALO_FIELDS boxes per call */
#define IGG_MAX_HALO_FIELDS 8
int igg_halo_exchange_peer_ex(const igg_halo_xfer* xfers_dev, int nxfers, const igg_halo_sync* sync, int ctas,
                              void* stream);

/* neighbour barrier on a bank: signal `epoch` into DATA slots of every neighbour, wait for theirs */
int igg_neighbor_barrier(const igg_halo_sync* sync, void* stream);
/* wait until the DATA slot of every neighbour in `sync` is >= sync->epoch (all pushes of that epoch
 * have landed in my halo cells) */
int igg_wait_epoch(const igg_halo_sync* sync, void* stream);
/* peer-store copy kernel used by the ring test / bandwidth sweep: dst (peer) <- src (local), then
 * *peer_flag = flag_value (release, system scope).  `my_count` is a private zero-initialised counter. */
int igg_peer_copy(void* dst, const void* src, size_t bytes, uint64_t* peer_flag, uint64_t flag_value,
                  uint64_t* my_count, void* stream);
/* The device-buffer Sendrecv of reference scripts/rocmaware_test_selectdevice.jl:22 as ONE launch: store `src` into the
 * peer's `dst`, release `flag_value` to `peer_flag` (in the receiver's memory), and do not finish before `my_flag` (in
 * mine, released by whoever sends to me) has reached `flag_value` too. */
int igg_peer_sendrecv(void* dst, const void* src, size_t bytes, uint64_t* peer_flag, const uint64_t* my_flag,
                      uint64_t flag_value, uint64_t* my_count, uint64_t* err_word, void* stream);
/* wait until *my_flag >= value */
int igg_wait_value(uint64_t* my_flag, uint64_t value, uint64_t* err_word, void* stream);

/* ---- diagnostics / output ---------------------------------------------------------------------- */
/* max and sum of n fp64 values (reference scripts/diffusion_2D_perf_hide.jl:115 `@show maximum(T_v)`).
 * `scratch`: igg_reduce_scratch_bytes(n) bytes of device memory, 16-byte aligned, zero-filled once
 * before its first use (the kernel leaves it ready for the next call); calls sharing one scratch
 * must not overlap on the GPU.  `out`: two device doubles {max, sum}. */
size_t igg_reduce_scratch_bytes(size_t n);
int igg_reduce_max_sum(const double* x, size_t n, void* scratch, double* out, void* stream);
/* The same reduction for callers that cannot allocate device memory (no counterpart in the
 * reference: it never reduces on the device without allocating).  Meant for hooks that run after the
 * application has taken the whole GPU for its fields, for hosts whose memory policy keeps allocation
 * to their own pool, and for time loops that cannot take the device-wide stall of cudaMalloc/cudaFree.
 * The scratch is a 16,640-byte array built into the library (held by every GPU that loads it, until
 * exit); inputs and outputs are still yours.
 *   - n <= IGG_REDUCE_STATIC_MAX_N, else IGG_ERR_UNSUPPORTED and a message naming the limit;
 *   - no cudaMalloc/cudaFree or allocating library call is made;
 *   - the first call on each device looks up the array's device address (host-side set-up): make it
 *     during start-up, before any stream capture;
 *   - all calls on a device share the array and it keeps the last call's contents (the ticket is
 *     re-zeroed on `stream` before each launch); calls that can overlap on the GPU race -- order them. */
#define IGG_REDUCE_STATIC_MAX_N (((size_t)1) << 29)
int igg_reduce_max_sum_static(const double* x, size_t n, double* out, void* stream);
/* field -> RGB8 through a 256x3 look-up table (reference scripts/diffusion_2D_ap.jl:47 heatmap) */
int igg_colormap(const double* field, size_t n, double vmin, double vmax, const uint8_t* lut_dev,
                 uint8_t* rgb, void* stream);

/* ---- CPU restatement (oracle + timed CPU baseline; the reference has no CPU path) --------------- */
int igg_cpu_diffusion2d_step(const double* T, double* T2, const double* Cp, int nx, int ny, double lam,
                             double dt, double _dx, double _dy);
int igg_cpu_diffusion2d_run(double* T, double* T2, const double* Cp, int nx, int ny, double lam, double dt,
                            double _dx, double _dy, int nt);
/* three-kernel form on the CPU (reference scripts/diffusion_2D_kp.jl:16-54); qx (nx-1)x(ny-2), qy (nx-2)x(ny-1),
 * dTdt (nx-2)x(ny-2), row-major */
int igg_cpu_flux2d(double* qx, double* qy, const double* T, int nx, int ny, double lam, double _dx, double _dy);
int igg_cpu_residual2d(double* dTdt, const double* qx, const double* qy, const double* Cp, int nx, int ny,
                       double _dx, double _dy);
int igg_cpu_update2d(double* T, const double* dTdt, int nx, int ny, double dt);
/* 3-D fused step on the CPU (what igg_diffusion3d_step computes in mode 0) */
int igg_cpu_diffusion3d_step(const double* T, double* T2, const double* Cp, int nx, int ny, int nz, double lam,
                             double dt, double _dx, double _dy, double _dz);

#ifdef __cplusplus
}
#endif
#endif /* IGG_B200_H */
