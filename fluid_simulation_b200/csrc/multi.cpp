// multi.cpp -- the x-slab multi-GPU solver behind the C ABI (fluidb200_multi_* in include/fluidb200.h):
// ONE process drives several devices, the way SURVEY.md section 8(e) sketches it, so that C++ callers of the
// reference's ParallelNavierStokes<T> interface can use more than one GPU without Python.
//
// Same decomposition and schedule as fluid_simulation_b200/distributed.py (one process per GPU): slab d owns
// rows [x0, x1) and stores kmax+1 halo rows per interior side; per timestep one u, v halo exchange and one p
// exchange per Jacobi launch. Halos travel by one-sided pushes over NVLink (fluidb200_halo_push: peer copy +
// flag; the receiver's stream waits with fluidb200_halo_wait), the two k-row edge strips of a launch are swept
// first on high-priority streams and pushed while the interior is swept. Everything is stream-ordered; one
// host thread per slab only ISSUES work. Built on the public stateless entry points, so the arithmetic -- and
// therefore the bits -- are those of the single-GPU path.
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <new>
#include <string>
#include <thread>
#include <vector>

#include "../../include/fluidb200.h"
#include "backend.h"

namespace {

constexpr int kFields = FLUIDB200_NUM_FIELDS;

struct Slab {
    int device = 0;
    int x0 = 0, x1 = 0, lo = 0, hi = 0;   // owned rows, stored rows
    char* block = nullptr;                // 7 fields + two flag words
    size_t field_bytes = 0;
    int phys[kFields];                    // which physical array a field name refers to (p / p_next swap)
    void* main = nullptr;
    void* edge[2] = {nullptr, nullptr};   // one high-priority stream per side
    void* ev_start = nullptr;
    void* ev_side[2] = {nullptr, nullptr};
    fluidb200_slab g;
    long long kernels = 0;
    char* field(int f) const { return block + (size_t)phys[f] * field_bytes; }
    char* flag(int side) const { return block + (size_t)kFields * field_bytes + 128 * side; }
};

}  // namespace

struct fluidb200_multi {
    int dtype = 0, X = 0, Y = 0;
    size_t esize = 4;
    int64_t pitch = 0;
    double hx = 0, hy = 0;
    double dt = 0.001, nu = 0.1, rho = 1.0;
    int K = 50, S = 1000;
    int kmax = 8, H = 9;
    std::vector<Slab> slabs;
    fluidb200_edges edges[3];
    unsigned seq = 0;
    bool fresh = true;          // state was (re)written from the host: settle before the first pushes
};

namespace {

int mfail(int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    return fluidb200_report_error(code, buf);
}
#define M_CHECK(expr)            \
    do {                         \
        const int rc_ = (expr);  \
        if (rc_ != 0) return rc_; \
    } while (0)
#define M_BE(expr, what)                                                                  \
    do {                                                                                  \
        if ((expr) != 0) return mfail(FLUIDB200_ECUDA, "%s: %s", what, fbk::be_last_error_string()); \
    } while (0)

size_t row_bytes(const fluidb200_multi* m) { return (size_t)m->pitch * m->esize; }
char* rows_ptr(const fluidb200_multi* m, const Slab& s, int f, int g0) { return s.field(f) + (size_t)(g0 - s.lo) * row_bytes(m); }

// my `depth` outermost owned rows of each listed field -> both neighbours' halo rows, each side followed by its flag
int push(fluidb200_multi* m, int d, const int* fields, int nfields, int depth, unsigned seq, void* const streams[2]) {
    Slab& s = m->slabs[d];
    const size_t nbytes = (size_t)depth * row_bytes(m);
    for (int side = 0; side < 2; ++side) {
        const int nb = side == 0 ? d - 1 : d + 1;
        if (nb < 0 || nb >= (int)m->slabs.size()) continue;
        const Slab& t = m->slabs[nb];
        const int g0 = side == 0 ? s.x0 : s.x1 - depth;
        void* st = streams ? streams[side] : s.main;
        for (int i = 0; i < nfields; ++i) {
            // the neighbour keeps the same name -> array assignment: every slab swaps p and p_next in lock step
            char* dst = t.block + (size_t)s.phys[fields[i]] * t.field_bytes + (size_t)(g0 - t.lo) * row_bytes(m);
            const char* src = rows_ptr(m, s, fields[i], g0);
            if (i < nfields - 1) M_CHECK(fluidb200_halo_copy(dst, src, nbytes, st));
            else M_CHECK(fluidb200_halo_push(dst, src, nbytes, t.flag(side == 0 ? 1 : 0), seq, st));   // I am its hi / lo neighbour
        }
        ++s.kernels;
    }
    return 0;
}
int wait_pushed(fluidb200_multi* m, int d, unsigned seq) {
    Slab& s = m->slabs[d];
    for (int side = 0; side < 2; ++side) {
        const int nb = side == 0 ? d - 1 : d + 1;
        if (nb < 0 || nb >= (int)m->slabs.size()) continue;
        M_CHECK(fluidb200_halo_wait(s.flag(side), seq, s.main));
    }
    return 0;
}

// One slab's share of `steps` timesteps, issued from its own host thread. `seq0`: exchanges made before.
int issue_steps(fluidb200_multi* m, int d, int steps, unsigned seq0) {
    Slab& s = m->slabs[d];
    const int n = (int)m->slabs.size();
    M_CHECK(fluidb200_set_device(s.device));
    const long long k0 = fbk::be_kernel_count();
    const int X = m->X, k = m->kmax;
    const bool first = d == 0, last = d == n - 1;
    unsigned seq = seq0;
    const int uv[2] = {FLUIDB200_FIELD_U, FLUIDB200_FIELD_V};
    const int pn[1] = {FLUIDB200_FIELD_P_NEXT};
    for (int step = 0; step < steps; ++step) {
        if (n > 1) {
            ++seq;
            M_CHECK(push(m, d, uv, 2, m->H, seq, nullptr));
            M_CHECK(wait_pushed(m, d, seq));
        }
        // predictor: u_next, v_next on owned +- kmax rows, rhs on owned +- (kmax - 1) rows (recomputed, not sent)
        int star_lo = std::max(0, s.x0 - k), star_hi = std::min(X, s.x1 + k);
        int rhs_lo = std::max(0, s.x0 - k + 1), rhs_hi = std::min(X, s.x1 + k - 1);
        if (n == 1) { star_lo = rhs_lo = 0; star_hi = rhs_hi = X; }
        M_CHECK(fluidb200_predictor(&s.g, s.field(0), s.field(1), s.field(3), s.field(4), s.field(6), m->dt, m->nu, m->rho, star_lo,
                                    star_hi, rhs_lo, rhs_hi, s.main));
        const int nlaunch = (m->K + k - 1) / k;
        int left = m->K;
        for (int i = 0; i < nlaunch; ++i) {
            const int kk = (left + (nlaunch - i) - 1) / (nlaunch - i);
            left -= kk;
            if (n == 1) {
                M_CHECK(fluidb200_poisson_sweeps(&s.g, s.field(2), s.field(5), s.field(6), &m->edges[2], nullptr, kk, 0, X, s.main));
            } else {
                // the k-row strips the neighbours wait for, on their own high-priority streams; pushed from there
                M_BE(fbk::be_event_record(s.ev_start, s.main), "multi: event");
                int in_lo = s.x0, in_hi = s.x1;
                for (int side = 0; side < 2; ++side) {
                    if ((side == 0 && first) || (side == 1 && last)) continue;
                    const int a = side == 0 ? s.x0 : s.x1 - k, b = a + k;
                    if (side == 0) in_lo = b; else in_hi = a;
                    M_BE(fbk::be_stream_wait_event(s.edge[side], s.ev_start), "multi: event");
                    M_CHECK(fluidb200_poisson_sweeps(&s.g, s.field(2), s.field(5), s.field(6), &m->edges[2], nullptr, kk, a, b, s.edge[side]));
                }
                ++seq;
                M_CHECK(push(m, d, pn, 1, k, seq, s.edge));
                for (int side = 0; side < 2; ++side)
                    if (!((side == 0 && first) || (side == 1 && last))) M_BE(fbk::be_event_record(s.ev_side[side], s.edge[side]), "multi: event");
                M_CHECK(fluidb200_poisson_sweeps(&s.g, s.field(2), s.field(5), s.field(6), &m->edges[2], nullptr, kk, in_lo, in_hi, s.main));
                for (int side = 0; side < 2; ++side)
                    if (!((side == 0 && first) || (side == 1 && last))) M_BE(fbk::be_stream_wait_event(s.main, s.ev_side[side]), "multi: event");
                M_CHECK(wait_pushed(m, d, seq));
            }
            std::swap(s.phys[FLUIDB200_FIELD_P], s.phys[FLUIDB200_FIELD_P_NEXT]);   // ping-pong instead of p = p_next
        }
        const int c_lo = n == 1 ? 0 : s.x0, c_hi = n == 1 ? X : s.x1;
        M_CHECK(fluidb200_corrector(&s.g, s.field(2), s.field(3), s.field(4), s.field(0), s.field(1), &m->edges[0], &m->edges[1], nullptr,
                                    nullptr, m->dt, m->rho, c_lo, c_hi, s.main));
    }
    s.kernels += fbk::be_kernel_count() - k0;
    return 0;
}

}  // namespace

// =========================================================================================
// C ABI
// =========================================================================================
extern "C" int fluidb200_multi_destroy(fluidb200_multi* m) {
    if (!m) return 0;
    for (Slab& s : m->slabs) {
        fluidb200_set_device(s.device);
        if (s.main) fbk::be_stream_sync(s.main);
        for (void* e : {s.ev_start, s.ev_side[0], s.ev_side[1]})
            if (e) fbk::be_event_destroy(e);
        for (void* st : {s.edge[0], s.edge[1], s.main})
            if (st) fbk::be_stream_destroy(st);
        if (s.block) fbk::be_free(s.block);
    }
    delete m;
    return 0;
}

extern "C" int fluidb200_multi_create(fluidb200_multi** out, int dtype, int X, int Y, double domain_x, double domain_y, int ndev,
                                      const int* devices) {
    if (!out) return mfail(FLUIDB200_EINVAL, "multi_create: out is NULL");
    *out = nullptr;
    if (dtype != FLUIDB200_F32 && dtype != FLUIDB200_F64) return mfail(FLUIDB200_EINVAL, "multi_create: bad dtype %d", dtype);
    if (X < 3 || Y < 3 || ndev < 1) return mfail(FLUIDB200_EINVAL, "multi_create: box %dx%d on %d devices", X, Y, ndev);
    fluidb200_multi* m = new (std::nothrow) fluidb200_multi();
    if (!m) return mfail(FLUIDB200_ENOMEM, "multi_create: host allocation failed");
    m->dtype = dtype; m->X = X; m->Y = Y;
    m->esize = dtype == FLUIDB200_F32 ? 4 : 8;
    m->pitch = ((int64_t)Y + 31) / 32 * 32;
    m->hx = fluidb200_element_length(dtype, domain_x, X);
    m->hy = fluidb200_element_length(dtype, domain_y, Y);
    {   // blocking depth as for a single device of the slab's size (csrc/api.cpp, fluidb200_create)
        const long long cells = (long long)X * Y / ndev;
        int k = cells >= 60000000LL ? 8 : cells >= 30000000LL ? 7 : cells >= 8000000LL ? 5 : 3;
        if (dtype == FLUIDB200_F64) k = std::min(k, 6);
        const char* e = getenv("FLUIDB200_KMAX");
        if (e && *e) k = atoi(e);
        m->kmax = std::max(1, std::min(k, FLUIDB200_MAX_SWEEPS_PER_LAUNCH));
    }
    m->H = m->kmax + 1;
    if (ndev > 1 && X / ndev < 2 * m->H) {
        delete m;
        return mfail(FLUIDB200_EINVAL, "multi_create: slabs of %d rows are too thin for halo depth %d", X / ndev, m->kmax + 1);
    }
    fluidb200_cavity_edges('u', 1.0, &m->edges[0]);
    fluidb200_cavity_edges('v', 1.0, &m->edges[1]);
    fluidb200_cavity_edges('p', 1.0, &m->edges[2]);
    m->slabs.resize(ndev);
    int rc = 0;
    for (int d = 0; d < ndev && rc == 0; ++d) {
        Slab& s = m->slabs[d];
        s.device = devices ? devices[d] : d;
        s.x0 = (int)((long long)d * X / ndev);
        s.x1 = (int)((long long)(d + 1) * X / ndev);
        s.lo = std::max(0, s.x0 - m->H);
        s.hi = std::min(X, s.x1 + m->H);
        if (ndev == 1) { s.lo = 0; s.hi = X; }
        for (int f = 0; f < kFields; ++f) s.phys[f] = f;
        s.field_bytes = (size_t)(s.hi - s.lo) * m->pitch * m->esize;
        s.g.dtype = dtype; s.g.X = X; s.g.Y = Y; s.g.pitch = m->pitch; s.g.row0 = s.lo; s.g.rows = s.hi - s.lo; s.g.hx = m->hx; s.g.hy = m->hy;
        if ((rc = fluidb200_set_device(s.device))) break;
        const size_t total = kFields * s.field_bytes + 256;
        void* p = nullptr;
        if (fbk::be_malloc(&p, total) != 0) { rc = mfail(FLUIDB200_ENOMEM, "multi_create: %zu bytes on device %d: %s", total, s.device, fbk::be_last_error_string()); break; }
        s.block = (char*)p;
        if (fbk::be_stream_create(&s.main) || fbk::be_stream_create(&s.edge[0]) || fbk::be_stream_create(&s.edge[1]) ||
            fbk::be_event_create(&s.ev_start) || fbk::be_event_create(&s.ev_side[0]) || fbk::be_event_create(&s.ev_side[1]) ||
            fbk::be_memset_async(s.block, 0, total, s.main) || fbk::be_stream_sync(s.main)) {
            rc = mfail(FLUIDB200_ECUDA, "multi_create: streams on device %d: %s", s.device, fbk::be_last_error_string());
            break;
        }
    }
    // every device reaches its neighbours' memory (NVLink peer access; nothing to do for slabs sharing a device)
    for (int d = 0; d < ndev && rc == 0; ++d)
        for (int nb : {d - 1, d + 1})
            if (nb >= 0 && nb < ndev && m->slabs[nb].device != m->slabs[d].device) {
                if ((rc = fluidb200_set_device(m->slabs[d].device))) break;
                if ((rc = fluidb200_peer_enable(m->slabs[nb].device))) break;
            }
    if (rc) {
        fluidb200_multi_destroy(m);
        return rc;
    }
    *out = m;
    return 0;
}

extern "C" int fluidb200_multi_set_params(fluidb200_multi* m, double time_step, double kinematic_viscosity, double density,
                                          int num_poisson_iterations, int num_iterations) {
    if (!m) return mfail(FLUIDB200_EINVAL, "multi: handle is NULL");
    if (num_poisson_iterations < 0 || num_iterations < 0) return mfail(FLUIDB200_EINVAL, "multi: iteration counts must be >= 0");
    m->dt = time_step; m->nu = kinematic_viscosity; m->rho = density;
    m->K = num_poisson_iterations; m->S = num_iterations;
    return 0;
}

extern "C" int fluidb200_multi_synchronize(fluidb200_multi* m) {
    if (!m) return mfail(FLUIDB200_EINVAL, "multi: handle is NULL");
    for (Slab& s : m->slabs) {
        M_CHECK(fluidb200_set_device(s.device));
        for (void* st : {s.edge[0], s.edge[1], s.main}) M_BE(fbk::be_stream_sync(st), "multi_synchronize");
    }
    return 0;
}

// zero state + the cavity's boundary values (fluid_cpp/main.cpp:43-55) on every slab, set up on the devices
extern "C" int fluidb200_multi_upload_cavity(fluidb200_multi* m, double lid_velocity) {
    if (!m) return mfail(FLUIDB200_EINVAL, "multi: handle is NULL");
    fluidb200_cavity_edges('u', lid_velocity, &m->edges[0]);
    fluidb200_cavity_edges('v', lid_velocity, &m->edges[1]);
    fluidb200_cavity_edges('p', lid_velocity, &m->edges[2]);
    for (Slab& s : m->slabs) {
        M_CHECK(fluidb200_set_device(s.device));
        M_CHECK(fluidb200_init_cavity(&s.g, s.field(0), s.field(1), s.field(2), s.field(3), s.field(4), s.field(5), s.field(6), lid_velocity, s.main));
        ++s.kernels;
    }
    m->fresh = true;
    return 0;
}

// host -> devices: fields[i] (dense X*Y, i = x*Y + y) or NULL; every slab takes its stored rows (owned + halo)
extern "C" int fluidb200_multi_upload(fluidb200_multi* m, const void* const fields[FLUIDB200_NUM_FIELDS]) {
    if (!m || !fields) return mfail(FLUIDB200_EINVAL, "multi_upload: NULL argument");
    const size_t row = (size_t)m->Y * m->esize;
    for (Slab& s : m->slabs) {
        M_CHECK(fluidb200_set_device(s.device));
        for (int f = 0; f < kFields; ++f)
            if (fields[f])
                M_BE(fbk::be_memcpy2d_async(s.field(f), row_bytes(m), (const char*)fields[f] + (size_t)s.lo * row, row, row, (size_t)(s.hi - s.lo),
                                            fbk::H2D, s.main), "multi_upload");
    }
    m->fresh = true;
    return fluidb200_multi_synchronize(m);
}

// devices -> host: every slab contributes its owned rows. Blocks until done.
extern "C" int fluidb200_multi_download(fluidb200_multi* m, int field, void* out) {
    if (!m || !out) return mfail(FLUIDB200_EINVAL, "multi_download: NULL argument");
    if (field < 0 || field >= kFields) return mfail(FLUIDB200_EINVAL, "multi_download: bad field %d", field);
    const size_t row = (size_t)m->Y * m->esize;
    int src = field;
    if (field == FLUIDB200_FIELD_P_NEXT && m->seq > 0) src = FLUIDB200_FIELD_P;   // p_next == p after any sweep (see fluidb200_download)
    for (Slab& s : m->slabs) {
        M_CHECK(fluidb200_set_device(s.device));
        M_BE(fbk::be_memcpy2d_async((char*)out + (size_t)s.x0 * row, row, rows_ptr(m, s, src, s.x0), row_bytes(m), row, (size_t)(s.x1 - s.x0), fbk::D2H,
                                    s.main), "multi_download");
    }
    return fluidb200_multi_synchronize(m);
}

// solve(): num_iterations timesteps from the current device state, asynchronous on every device
extern "C" int fluidb200_multi_solve(fluidb200_multi* m) {
    if (!m) return mfail(FLUIDB200_EINVAL, "multi: handle is NULL");
    if (m->S == 0) return 0;
    if (!(m->dt != 0) || !(m->rho != 0)) return mfail(FLUIDB200_EINVAL, "multi: time_step and density must be non-zero");
    const int n = (int)m->slabs.size();
    if (m->fresh) {   // a neighbour's pushes may only arrive once every slab's set-up is done
        M_CHECK(fluidb200_multi_synchronize(m));
        m->fresh = false;
    }
    // With no sweep launch between them nothing orders one step's pushes behind the neighbour's reads of the previous
    // step: such solves (K = 0) are issued one step at a time with a host-side join between the steps.
    const int rounds = m->K == 0 && n > 1 ? m->S : 1, per_round = m->K == 0 && n > 1 ? 1 : m->S;
    const int exchanges_per_step = n > 1 ? 1 + (m->K + m->kmax - 1) / m->kmax : 0;
    for (int r = 0; r < rounds; ++r) {
        std::vector<int> rc(n, 0);
        std::vector<std::string> msg(n);
        std::vector<std::thread> th;
        const unsigned seq0 = m->seq;
        for (int d = 0; d < n; ++d)
            th.emplace_back([&, d] {
                rc[d] = issue_steps(m, d, per_round, seq0);
                if (rc[d]) msg[d] = fluidb200_last_error();
            });
        for (auto& t : th) t.join();
        m->seq = seq0 + (unsigned)(per_round * exchanges_per_step);
        if (n == 1) m->seq += (unsigned)(per_round * ((m->K + m->kmax - 1) / m->kmax));   // (counts sweeps launches: p_next aliasing)
        for (int d = 0; d < n; ++d)
            if (rc[d]) return mfail(rc[d], "multi_solve, slab %d: %s", d, msg[d].c_str());
        if (rounds > 1) M_CHECK(fluidb200_multi_synchronize(m));
    }
    return 0;
}

extern "C" int64_t fluidb200_multi_launch_count(const fluidb200_multi* m) {
    int64_t n = 0;
    if (m)
        for (const Slab& s : m->slabs) n += s.kernels;
    return n;
}
extern "C" int fluidb200_multi_num_slabs(const fluidb200_multi* m) { return m ? (int)m->slabs.size() : 0; }
