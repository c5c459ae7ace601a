/*
 * fluidb200.h -- C ABI of the B200-native lid-driven-cavity projection solver.
 *
 * This is the drop-in boundary for the hot path of danieljvickers/fluid_simulation's
 * fluid_cpp library (tentative velocity -> Jacobi pressure-Poisson sweeps -> velocity
 * correction). The reference has no C ABI of its own (no extern "C", pybind or ctypes
 * anywhere in its tree); the entry points below are what a binding for this path would
 * bind, one for each call the reference's C++ classes make. Each declaration cites the
 * reference interface it replaces. The C++ class templates that keep the reference's
 * names and signatures (NavierStokesSolver / ParallelNavierStokes / TiledNavierStokes)
 * are header-only shims over this file: include/fluid_cpp/.
 *
 * Conventions
 *  - plain pointers and sizes only; no C++ or torch types cross this boundary;
 *  - every function returns 0 on success or a negative FLUIDB200_E* code, and records
 *    a message retrievable with fluidb200_last_error() (thread-local). The reference
 *    signals no errors at all (its CUDA calls are unchecked, its int returns are
 *    always 0), so the C++ shims swallow the code and keep the reference's returns;
 *  - dtype: 0 = float (fp32), 1 = double (fp64)  -- the reference instantiates exactly
 *    these two (fluid_cpp/cuda/ParallelNavierStokes.cuh:50-51);
 *  - HOST field arrays are dense X*Y, index i = x*Y + y (fluid_cpp/tests/test_main.cpp:76;
 *    the device order of fluid_cpp/cuda/ParallelNavierStokes.cu:22);
 *  - DEVICE field arrays are structure-of-arrays, one array per field, row x starts at
 *    base + x*pitch elements, y contiguous; pitch >= Y and pitch % 32 == 0; the bytes of
 *    a row beyond Y are padding the kernels may overwrite;
 *  - scalar parameters travel as double and must hold exactly the T value the caller
 *    has (float -> double is exact), so no precision is lost at the boundary;
 *  - `stream` is a cudaStream_t passed as void* (NULL = the legacy default stream).
 *    All launches are asynchronous, like the reference's solve()
 *    (fluid_cpp/cuda/ParallelNavierStokes.cu:269-288 returns without synchronising).
 */
#ifndef FLUIDB200_H
#define FLUIDB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FLUIDB200_F32 0
#define FLUIDB200_F64 1

#define FLUIDB200_OK 0
#define FLUIDB200_EINVAL (-1)  /* bad argument (sizes, alignment, ranges, NULL)          */
#define FLUIDB200_ECUDA (-2)   /* a CUDA runtime call or kernel launch failed             */
#define FLUIDB200_ENOMEM (-3)  /* host or device allocation failed                        */
#define FLUIDB200_ESTATE (-4)  /* call made in the wrong state (e.g. no device buffers)   */

/* Largest number of Jacobi sweeps one kernel launch can block in time. */
#define FLUIDB200_MAX_SWEEPS_PER_LAUNCH 8

/* Field selectors for download/upload/pointer queries. */
enum {
    FLUIDB200_FIELD_U = 0,      /* NavierStokesCell::u               (NavierStokesCell.h:12) */
    FLUIDB200_FIELD_V = 1,      /* NavierStokesCell::v               (:13)                   */
    FLUIDB200_FIELD_P = 2,      /* NavierStokesCell::p               (:14)                   */
    FLUIDB200_FIELD_U_NEXT = 3, /* NavierStokesCell::u_next          (:28)                   */
    FLUIDB200_FIELD_V_NEXT = 4, /* NavierStokesCell::v_next          (:29)                   */
    FLUIDB200_FIELD_P_NEXT = 5, /* NavierStokesCell::p_next          (:30)                   */
    FLUIDB200_FIELD_RHS = 6,    /* NavierStokesCell::right_hand_size (:25)                   */
    FLUIDB200_NUM_FIELDS = 7
};

/* ------------------------------------------------------------------------------------
 * Boundary description of ONE field on the four edges of the global box.
 *
 * Replaces the per-cell `*_boundary` / `*_boundary_set` members the reference scans on
 * every sweep (NavierStokesCell.h:36-41; SerialNavierStokes.cpp:152-189 and :222-295).
 * Per edge a mode says how to read it:
 *   FLUIDB200_EDGE_UNSET  no cell of the edge has its flag set (pure copy rule),
 *   FLUIDB200_EDGE_CONST  every cell of the edge is set to the same value `cval`,
 *   FLUIDB200_EDGE_ARRAY  per-cell `set` (bytes) and `val` (T) DEVICE arrays.
 * Edge order: 0 = bottom (y=0,   indexed by x, length X, owns both bottom corners),
 *             1 = top    (y=Y-1, indexed by x, length X, owns both top corners),
 *             2 = left   (x=0,   indexed by y, length Y, entries 1..Y-2 used),
 *             3 = right  (x=X-1, indexed by y, length Y, entries 1..Y-2 used).
 * The corner ownership mirrors the reference's loop order (y-edges for all x first,
 * then x-edges for y in [1,Y-2]).
 * ---------------------------------------------------------------------------------- */
#define FLUIDB200_EDGE_UNSET 0
#define FLUIDB200_EDGE_CONST 1
#define FLUIDB200_EDGE_ARRAY 2
#define FLUIDB200_EDGE_BOTTOM 0
#define FLUIDB200_EDGE_TOP 1
#define FLUIDB200_EDGE_LEFT 2
#define FLUIDB200_EDGE_RIGHT 3

typedef struct fluidb200_edges {
    int mode[4];
    double cval[4];           /* used when mode == CONST (exact T value)                  */
    const void* val[4];       /* device T[len]; used when mode == ARRAY                   */
    const uint8_t* set[4];    /* device uint8[len]; used when mode == ARRAY               */
} fluidb200_edges;

/* Optional interior-flag mask of ONE field (the "general BC" path): dense device arrays
 * with the field's own pitch. NULL `set` = no interior cell carries a flag, which is the
 * lid-driven cavity and the fast path (SerialNavierStokes.cpp:154-160 / :224-233 then do
 * nothing). */
typedef struct fluidb200_mask {
    const uint8_t* set;       /* device uint8[rows*pitch] or NULL (4-byte aligned)        */
    const void* val;          /* device T[rows*pitch]     or NULL (16-byte aligned)       */
    /* Optional (pressure mask only, may be NULL): a summary of `set`, one byte per tile of
     * FLUIDB200_MASK_TILE_ROWS x FLUIDB200_MASK_TILE_COLS cells (local rows, columns from 0),
     * tile (tr, tc) at tiles[tr * ceil(Y / TILE_COLS) + tc], non-zero when the tile holds a
     * flagged INTERIOR cell (a zero byte promises there is none). With it fluidb200_poisson_sweeps
     * runs the flag-free parts of the grid through its lean kernels; without it every task of a
     * masked launch takes the general kernel. The solver handle builds it on upload. */
    const uint8_t* tiles;
} fluidb200_mask;
#define FLUIDB200_MASK_TILE_ROWS 16
#define FLUIDB200_MASK_TILE_COLS 64

/* Geometry of the rows held in one device array set (a whole grid, or one x-slab of it).
 * Local row 0 is global row `row0`; the arrays hold `rows` rows. */
typedef struct fluidb200_slab {
    int dtype;                /* FLUIDB200_F32 / FLUIDB200_F64                            */
    int X, Y;                 /* GLOBAL box_dimension_x / _y (NavierStokesSolver.h:11-12) */
    int64_t pitch;            /* elements between consecutive rows                        */
    int row0;                 /* global x of local row 0                                  */
    int rows;                 /* rows present locally                                     */
    double hx, hy;            /* element_length_x/_y as exact T values (NavierStokesSolver.cpp:57-58) */
} fluidb200_slab;

/* ====================================================================================
 * Stateless kernel launches on caller-owned device memory
 * ==================================================================================*/

/* Fused tentative velocity + divergence right-hand side, one HBM pass.
 * Replaces ParallelNavierStokes::unifiedApproximateTimeStep + unifiedComputeRightHand
 * (fluid_cpp/cuda/ParallelNavierStokes.cu:113-170, host wrappers :312-322), i.e.
 * SerialNavierStokes.cpp:298-335.
 * Reads u, v; writes u_next, v_next on interior cells of global rows [star_lo, star_hi) and
 * rhs on interior cells of global rows [rhs_lo, rhs_hi). Needs u, v on rows
 * [min(star_lo, rhs_lo-1)-1, max(star_hi, rhs_hi+1)+1) clipped to the box, and the stored edge
 * values of u_next / v_next (the reference never writes them).
 * dt, nu, rho: time_step, kinematic_viscosity, density (NavierStokesSolver.h:22-24). */
int fluidb200_predictor(const fluidb200_slab* g, const void* u, const void* v, void* u_next,
                        void* v_next, void* rhs, double dt, double nu, double rho, int star_lo,
                        int star_hi, int rhs_lo, int rhs_hi, void* stream);

/* `sweeps` (1..FLUIDB200_MAX_SWEEPS_PER_LAUNCH) Jacobi sweeps in ONE launch, each with its
 * boundary rows applied, result written to p_out (ping-pong: p_out must not alias p_in).
 * Replaces `sweeps` iterations of computePoissonStepApproximation + enforcePressureBoundary
 * Conditions + updatePressure (fluid_cpp/cuda/ParallelNavierStokes.cu:172-192, :14-56, loop at
 * :277-282), i.e. SerialNavierStokes.cpp:140-198.
 * Writes global rows [out_lo, out_hi) of p_out (all Y columns, edge rows included when the
 * range touches them). Needs p_in on rows [out_lo-sweeps, out_hi+sweeps) and rhs on rows
 * [out_lo-sweeps+1, out_hi+sweeps-1), both clipped to the box; the range must be present
 * locally. `mask` (may be NULL) enables interior pressure flags (SerialNavierStokes.cpp:154-160),
 * applied at every one of the `sweeps` levels. A masked launch runs the general kernel on every
 * task unless mask->tiles is given (see fluidb200_mask).
 * Large launches are split over the caller's stream and a high-priority helper stream the library
 * keeps per (device, caller stream); the first such launch from a stream creates it (one stream,
 * two events), so make that first call outside any stream capture. */
int fluidb200_poisson_sweeps(const fluidb200_slab* g, const void* p_in, void* p_out,
                             const void* rhs, const fluidb200_edges* p_edges,
                             const fluidb200_mask* mask, int sweeps, int out_lo, int out_hi,
                             void* stream);

/* Fused pressure-gradient correction + velocity boundary rows.
 * Replaces ParallelNavierStokes::unifiedVelocityCorrection + enforceVelocityBoundaryConditions
 * (fluid_cpp/cuda/ParallelNavierStokes.cu:194-217 and :58-111), i.e.
 * SerialNavierStokes.cpp:338-350 and :222-295.
 * Reads p, u_next, v_next; writes u, v on global rows [out_lo, out_hi) (all columns).
 * Needs p on rows [out_lo-1, out_hi+1) clipped to the box. */
int fluidb200_corrector(const fluidb200_slab* g, const void* p, const void* u_next,
                        const void* v_next, void* u, void* v, const fluidb200_edges* u_edges,
                        const fluidb200_edges* v_edges, const fluidb200_mask* u_mask,
                        const fluidb200_mask* v_mask, double dt, double rho, int out_lo,
                        int out_hi, void* stream);

/* Jacobi sweeps IN PLACE on the caller's pressure field without allocating anything.
 * No counterpart in the reference (its solver owns all of its memory and allocates it in the
 * constructor); added for callers that cannot allocate device memory themselves and cannot let a
 * library do it for them: code that runs after a memory pool has been frozen or filled, loops
 * with a latency budget that a trip into the driver's allocator would break (cudaMalloc may
 * block and synchronises the device), and hosts that pass a plug-in its buffers and forbid it
 * any allocation of its own.
 *
 * Runs `sweeps` (any count >= 0) sweeps of fluidb200_poisson_sweeps' arithmetic on the whole grid
 * (g->row0 must be 0 and g->rows == g->X) and leaves the result in `p`; `rhs` and `p_edges` as for
 * fluidb200_poisson_sweeps; no interior mask. The second buffer Jacobi needs is a fixed array of
 * FLUIDB200_NOALLOC_MAX_BYTES inside the library image: every GPU that loads the library holds
 * those bytes until the process exits, whether or not this function is ever called.
 *   - Limit: g->rows * g->pitch * sizeof(T) must not exceed FLUIDB200_NOALLOC_MAX_BYTES (16 MiB: a
 *     2048 x 2048 fp32 or 2048 x 1024 fp64 grid). Larger grids are refused with FLUIDB200_EINVAL
 *     and a message that says so; use fluidb200_poisson_sweeps with a buffer of your own.
 *   - The function calls no cudaMalloc / cudaFree and no routine that allocates, and creates no
 *     stream or event: every launch takes the single-kernel form (never the split form of
 *     fluidb200_poisson_sweeps, whose first use on a stream creates a helper stream). The first call
 *     on each device looks the array's device address up (one cudaGetSymbolAddress, with that
 *     device current) and caches it; make that first call outside any stream capture.
 *   - One array serves every call on a device and is not cleared between calls. Nothing needs to
 *     start at zero: each launch writes every cell that a later launch of the same call reads for
 *     a result, so leftovers of an earlier call (even of another grid size) cannot reach `p`.
 *   - Because the array is shared, two calls whose kernels may run at the same time on one GPU
 *     (different streams, different host threads) race on it. Serialise them.
 * Asynchronous on `stream`, like the other launches. */
#define FLUIDB200_NOALLOC_MAX_BYTES (16u << 20)
int fluidb200_poisson_solve_noalloc(const fluidb200_slab* g, void* p, const void* rhs,
                                    const fluidb200_edges* p_edges, int sweeps, void* stream);

/* Device-side initialiser: zero state plus the lid-driven-cavity boundary values of
 * fluid_cpp/main.cpp:43-55 (lid u = lid_velocity for all x at y=Y-1 incl. corners, everything
 * else 0), without any host array. Fills the local rows of u, v, p, u_next, v_next, p_next, rhs
 * (any may be NULL). Needed because the reference's host AoS grid (92 B/cell,
 * NavierStokesSolver.cpp:15-18) cannot hold the 16384^2+ configurations. */
int fluidb200_init_cavity(const fluidb200_slab* g, void* u, void* v, void* p, void* u_next,
                          void* v_next, void* p_next, void* rhs, double lid_velocity,
                          void* stream);

/* Edge description of the cavity of fluid_cpp/main.cpp:43-55 for field 'u', 'v' or 'p'
 * (no device arrays needed: u/v all four edges CONST, p top CONST 0 and the rest UNSET). */
int fluidb200_cavity_edges(char field, double lid_velocity, fluidb200_edges* out);

/* ====================================================================================
 * Stateful solver handle: owns the device SoA state of one whole grid on the current
 * device. One-to-one with the reference's ParallelNavierStokes<T> object.
 * ==================================================================================*/
typedef struct fluidb200_solver fluidb200_solver;

/* ParallelNavierStokes(int box_dim_x, int box_dim_y, T domain_size_x, T domain_size_y)
 * (fluid_cpp/cuda/ParallelNavierStokes.cu:223-237): allocates device state, zero-filled,
 * no boundary flags -- what the reference uploads from its pristine host grid. */
int fluidb200_create(fluidb200_solver** out, int dtype, int X, int Y, double domain_x,
                     double domain_y);
/* ~ParallelNavierStokes (:239-242); frees everything (the reference leaks five arrays). */
int fluidb200_destroy(fluidb200_solver* s);

/* The public data members read at every phase (NavierStokesSolver.h:22-26). May be called
 * between solve() calls, like assigning the fields. */
int fluidb200_set_params(fluidb200_solver* s, double time_step, double kinematic_viscosity,
                         double density, int num_poisson_iterations, int num_iterations);
/* setDomainSize (NavierStokesSolver.cpp:53-61): recomputes element lengths in T. */
int fluidb200_set_domain(fluidb200_solver* s, double domain_x, double domain_y);

/* migrateHostToDevice (fluid_cpp/cuda/ParallelNavierStokes.cu:244-249). `fields[i]` is a dense
 * host array for FLUIDB200_FIELD_i or NULL to leave that device field untouched.
 * `bc_val[f]` / `bc_set[f]` (f = 0:u, 1:v, 2:p) are dense host arrays of boundary values and
 * flags (NavierStokesCell.h:36-41), or both NULL to keep the current boundary description. */
int fluidb200_upload(fluidb200_solver* s, const void* const fields[FLUIDB200_NUM_FIELDS],
                     const void* const bc_val[3], const uint8_t* const bc_set[3]);
/* Boundary set-up of fluid_cpp/main.cpp:43-55 applied on the device (state zeroed). */
int fluidb200_upload_cavity(fluidb200_solver* s, double lid_velocity);

/* solve() (fluid_cpp/cuda/ParallelNavierStokes.cu:269-288): num_iterations timesteps from the
 * current device state; asynchronous on the solver's stream. */
int fluidb200_solve(fluidb200_solver* s);
/* Individual phases of one timestep (the protected wrappers at :297-334), for tests/tools. */
int fluidb200_step_predictor(fluidb200_solver* s);
int fluidb200_step_poisson(fluidb200_solver* s, int sweeps);
int fluidb200_step_corrector(fluidb200_solver* s);
/* What the reference's callers do with cudaDeviceSynchronize() (fluid_cpp/cuda_main.cu:98). */
int fluidb200_synchronize(fluidb200_solver* s);

/* migrateDeviceToHost + get*Values in one step (ParallelNavierStokes.cu:252-257,
 * NavierStokesSolver.cpp:87-145): dense host array, i = x*Y + y. Blocks until done. */
int fluidb200_download(fluidb200_solver* s, int field, void* out);

/* migrateSolve (fluid_cpp/cuda/ParallelNavierStokes.cu:290-295) on caller-owned HOST arrays: upload u, v, p
 * (host_in[0..2], dense X*Y, i = x*Y + y), run solve(), write the new u, v, p to host_out[0..2] (any may be NULL).
 * The other fields and the boundary description stay as they are on the device. Blocks until host_out is complete.
 * With num_iterations == 1 on a tall grid without interior flags the three legs are PIPELINED over row chunks
 * (upload of chunk c+1, the timestep on chunk c, download of chunk c-1 on three streams, each chunk computed with
 * K+2 recomputed halo rows), so the link is busy in both directions at once; pinned host memory is needed for
 * that overlap to happen (pageable memory works, serially). Results are bit-identical to
 * fluidb200_upload + fluidb200_solve + fluidb200_download. */
int fluidb200_migrate_solve(fluidb200_solver* s, const void* const host_in[3], void* const host_out[3]);

/* Introspection for tools and benchmarks. */
int fluidb200_get_slab(const fluidb200_solver* s, fluidb200_slab* out);
void* fluidb200_field_ptr(fluidb200_solver* s, int field);        /* device pointer            */
int fluidb200_set_stream(fluidb200_solver* s, void* stream);       /* default: own stream       */
int fluidb200_set_sweeps_per_launch(fluidb200_solver* s, int k);   /* temporal-blocking depth   */
int fluidb200_set_use_graph(fluidb200_solver* s, int enable);      /* replay steps as CUDA graph */
/* solve() on a small grid (up to about 1.2 M cells fp32) whose boundaries are the cavity's kind (every
 * velocity edge one constant, every pressure edge unset or one constant, no interior flags) runs all
 * num_iterations timesteps in ONE cooperative launch with the pressure held in registers, instead of
 * the 4 + 3K launches per step of fluid_cpp/cuda/ParallelNavierStokes.cu:269-288. Same bits either way.
 * enable: 1 / 0, or -1 for the default (on; FLUIDB200_RESIDENT=0 turns it off). */
int fluidb200_set_resident(fluidb200_solver* s, int enable);
int64_t fluidb200_resident_launch_count(const fluidb200_solver* s); /* solve() calls that took that path */
/* Measurement aid: SM cycles one tile spent per phase of the last such launch when FLUIDB200_RESIDENT_PROF=1
 * (out[0..5]: halo load + barrier, tentative velocity, right-hand side, sweeps, corrector, write-back; a library built
 * with -DFLUIDB200_RESIDENT_SUBPROF=1 also splits the sweeps into out[8..12]: waiting for the neighbours' rows, first + last
 * row and their publication, rows between, fix-ups + columns, CTA barrier, and counts polling rounds in out[13]). */
int fluidb200_resident_profile(fluidb200_solver* s, unsigned long long out[24]);
/* Kernels launched by this solver since creation (for bench.py's gpu_launches). */
int64_t fluidb200_launch_count(const fluidb200_solver* s);         /* kernels launched so far (a split Jacobi launch = 3) */

/* ====================================================================================
 * x-slab halo links between GPUs (no counterpart in the reference, which knows one device): the
 * plumbing the multi-GPU driver (fluid_simulation_b200/distributed.py) uses instead of a send/recv
 * pair per sweep launch. A rank PUSHES the rows its neighbour needs straight into the neighbour's
 * slab over NVLink (a plain copy into peer memory mapped through CUDA IPC, or reached by peer access
 * inside one process) and then sets a 32-bit flag in the neighbour's memory to a sequence number; the
 * neighbour's stream waits for that number. Everything is stream-ordered: no host thread waits.
 * ==================================================================================*/
/* Device memory another process can map: a plain cudaMalloc, zero-filled (pool allocators hand out
 * sub-ranges that CUDA IPC cannot export). */
int fluidb200_device_alloc(void** ptr, size_t bytes);
int fluidb200_device_free(void* ptr);
int fluidb200_ipc_export(const void* ptr, unsigned char handle[64]);   /* cudaIpcGetMemHandle             */
int fluidb200_ipc_open(const unsigned char handle[64], void** ptr);    /* cudaIpcOpenMemHandle, in the peer */
int fluidb200_ipc_close(void* ptr);
int fluidb200_peer_enable(int peer_device);                            /* same process, other device       */
/* On `stream`: copy `bytes` from src (this GPU) to dst (the peer's memory), then *flag = seq (flag in the
 * peer's memory). A push of several arrays: fluidb200_halo_copy for all but the last. */
int fluidb200_halo_push(void* dst, const void* src, size_t bytes, void* flag, uint32_t seq, void* stream);
int fluidb200_halo_copy(void* dst, const void* src, size_t bytes, void* stream);
/* On `stream`: wait until (int32)(*flag - seq) >= 0 (flag in THIS GPU's memory); cuStreamWaitValue32, no SM used. */
int fluidb200_halo_wait(void* flag, uint32_t seq, void* stream);

/* ====================================================================================
 * Multi-GPU solver handle: ONE process, several devices, the grid cut into x-slabs (one per entry of
 * `devices`; NULL = devices 0 .. ndev-1; an entry may repeat, which puts several slabs on one GPU --
 * useful for tests). The reference has no multi-device path (no cudaSetDevice anywhere in its tree); the
 * calls mirror the single-device handle above one to one, so ParallelNavierStokes<T>'s methods map onto
 * them the same way. Cavity-style boundaries (fluidb200_multi_upload_cavity).
 * Results are bit-identical to the single-device handle: Jacobi is order-free and the kernels are the same.
 * Halos travel by fluidb200_halo_push over NVLink peer access; every call is made from one host thread,
 * solve() itself issues the devices' work from one short-lived thread per slab and returns without waiting.
 * ==================================================================================*/
typedef struct fluidb200_multi fluidb200_multi;
int fluidb200_multi_create(fluidb200_multi** out, int dtype, int X, int Y, double domain_x, double domain_y,
                           int ndev, const int* devices);
int fluidb200_multi_destroy(fluidb200_multi* m);
int fluidb200_multi_set_params(fluidb200_multi* m, double time_step, double kinematic_viscosity, double density,
                               int num_poisson_iterations, int num_iterations);
int fluidb200_multi_upload_cavity(fluidb200_multi* m, double lid_velocity);                 /* fluid_cpp/main.cpp:43-55 */
int fluidb200_multi_upload(fluidb200_multi* m, const void* const fields[FLUIDB200_NUM_FIELDS]); /* migrateHostToDevice: dense X*Y arrays or NULL */
int fluidb200_multi_solve(fluidb200_multi* m);                                              /* solve(), asynchronous */
int fluidb200_multi_synchronize(fluidb200_multi* m);
int fluidb200_multi_download(fluidb200_multi* m, int field, void* out);                     /* migrateDeviceToHost + getter, blocking */
int64_t fluidb200_multi_launch_count(const fluidb200_multi* m);
int fluidb200_multi_num_slabs(const fluidb200_multi* m);

/* ====================================================================================
 * Library-level
 * ==================================================================================*/
/* Select the CUDA device the calling thread's later fluidb200_* calls use (cudaSetDevice inside
 * this library's own runtime instance). The reference never selects a device (single GPU). */
int fluidb200_set_device(int device);
const char* fluidb200_last_error(void);
/* Record `message` as the calling thread's last error and return `code` (used by the parts of the
 * library that live in other translation units; harmless for callers). */
int fluidb200_report_error(int code, const char* message);
const char* fluidb200_version(void);
/* element_length = domain / (n - 1) evaluated in T (NavierStokesSolver.cpp:57-58). */
double fluidb200_element_length(int dtype, double domain, int n);

#ifdef __cplusplus
}
#endif
#endif /* FLUIDB200_H */
