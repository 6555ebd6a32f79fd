// xlb_capi.cu -- extern "C" surface of libxlb_b200.so (declared in include/xlb_b200.h).
#include <atomic>
#include <cstdio>
#include <cstring>
#include <mutex>

#include "xlb_host.cuh"
#include "xlb_mesh.cuh"
#include "xlb_ops.cuh"

namespace xlb {

static thread_local char g_err[512] = "";
static std::atomic<long long> g_launches{0};

int set_error(int code, const char* msg) {
    std::snprintf(g_err, sizeof(g_err), "%s", msg);
    return code;
}

int count_launch(cudaError_t err) {
    if (err != cudaSuccess) {
        std::snprintf(g_err, sizeof(g_err), "CUDA error %d: %s", (int)err, cudaGetErrorString(err));
        return (int)err;
    }
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return 0;
}

static int check_domain(const XlbDomain* d, const char* who) {
    if (!d) return set_error(XLB_EINVAL, "null XlbDomain");
    const bool lat_ok = (d->d == 2 && d->q == 9) || (d->d == 3 && (d->q == 19 || d->q == 27));
    if (!lat_ok) return set_error(XLB_EUNSUPPORTED, "unsupported velocity set (need D2Q9, D3Q19 or D3Q27)");
    if (d->nx < 1 || d->ny < 1 || d->nz < 1) return set_error(XLB_EINVAL, "domain extents must be >= 1");
    if (d->d == 2 && d->nx != 1) return set_error(XLB_EINVAL, "2-D fields are passed as nx=1, ny=NX, nz=NY");
    (void)who;
    return 0;
}

static OpParams make_op(const XlbDomain* d) {
    OpParams p;
    std::memset(&p, 0, sizeof(p));
    p.nx = d->nx;
    p.ny = d->ny;
    p.nz = d->nz;
    p.x_periodic = d->x_periodic;
    p.ncell = (long long)d->nx * d->ny * d->nz;
    p.qs = d->q_stride ? d->q_stride : p.ncell;
    p.mqs = d->mask_q_stride ? d->mask_q_stride : p.ncell;
    return p;
}

static unsigned blocks_for(long long n, int b) { return (unsigned)((n + b - 1) / b); }

// dispatch helper: call F<T,Q>(...) for the runtime (compute dtype, q)
#define XLB_DISPATCH_TQ(dom, CALL)                                                       \
    do {                                                                                 \
        const int q_ = (dom)->q;                                                         \
        if ((dom)->compute_dtype == XLB_F64) {                                           \
            using T = double;                                                            \
            if (q_ == 9) { constexpr int Q = 9; CALL; }                                  \
            else if (q_ == 19) { constexpr int Q = 19; CALL; }                           \
            else { constexpr int Q = 27; CALL; }                                         \
        } else {                                                                         \
            using T = float;                                                             \
            if (q_ == 9) { constexpr int Q = 9; CALL; }                                  \
            else if (q_ == 19) { constexpr int Q = 19; CALL; }                           \
            else { constexpr int Q = 27; CALL; }                                         \
        }                                                                                \
    } while (0)

}  // namespace xlb

using namespace xlb;

// Tile-scheduler words of the persistent bulk kernel (xlb_step_bulk.cuh): XLB_SCHED_SLOTS pairs of zero-initialised counters;
// every launch gets the next slot (round robin) and its last block puts the pair back to zero, so launches that overlap on
// the GPU never share a pair unless more than XLB_SCHED_SLOTS launches are in flight at once.  (A launch captured in a CUDA
// graph keeps its slot on every replay.)  The device address is resolved once per device with cudaGetSymbolAddress.
#define XLB_SCHED_SLOTS 1024
static __device__ __align__(256) unsigned g_step_sched[2 * XLB_SCHED_SLOTS];
static std::atomic<unsigned> g_sched_next{0};

static int step_sched_slot(unsigned** out) {
    constexpr int MAX_DEV = 64;
    static std::mutex mu;
    static unsigned* cached[MAX_DEV] = {nullptr};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return count_launch(e);
    if (dev < 0 || dev >= MAX_DEV) {  // beyond the cache: static stride
        *out = nullptr;
        return 0;
    }
    {
        std::lock_guard<std::mutex> lock(mu);
        if (cached[dev] == nullptr) {
            void* ptr = nullptr;
            e = cudaGetSymbolAddress(&ptr, g_step_sched);
            if (e != cudaSuccess || ptr == nullptr) {
                std::snprintf(g_err, sizeof(g_err), "xlb_step: cudaGetSymbolAddress(g_step_sched) failed: %s", cudaGetErrorString(e));
                cudaGetLastError();
                return e != cudaSuccess ? (int)e : XLB_EINVAL;
            }
            cached[dev] = static_cast<unsigned*>(ptr);
        }
        *out = cached[dev] + 2 * (g_sched_next.fetch_add(1, std::memory_order_relaxed) % XLB_SCHED_SLOTS);
    }
    return 0;
}

extern "C" {

const char* xlb_last_error(void) { return g_err; }
int xlb_abi_version(void) { return XLB_B200_ABI_VERSION; }
int xlb_bc_entry_size(void) { return (int)sizeof(XlbBcEntry); }
int xlb_domain_size(void) { return (int)sizeof(XlbDomain); }
int xlb_collision_params_size(void) { return (int)sizeof(XlbCollisionParams); }
long long xlb_launch_count(void) { return g_launches.load(); }

int xlb_step(const XlbDomain* dom, int collision, const void* f0, void* f1, const uint8_t* bc_mask, const uint8_t* missing_mask,
             const XlbBcEntry* bc_table, int n_bc, double omega, int timestep, int x_begin, int x_end, int variant, void* stream) {
    return xlb_step_ex(dom, collision, nullptr, f0, f1, bc_mask, missing_mask, bc_table, n_bc, omega, timestep, x_begin, x_end, variant, stream);
}

int xlb_step_ex(const XlbDomain* dom, int collision, const XlbCollisionParams* cp, const void* f0, void* f1, const uint8_t* bc_mask,
                const uint8_t* missing_mask, const XlbBcEntry* bc_table, int n_bc, double omega, int timestep, int x_begin, int x_end,
                int variant, void* stream) {
    if (x_begin > x_end) return set_error(XLB_EINVAL, "xlb_step: bad x range");
    XlbStepRange r;
    r.x_begin = x_begin;
    r.x_stride = 1;
    r.n_planes = x_end - x_begin;
    r.y_begin = 0;
    r.y_stride = 1;
    r.n_rows = 0;
    return xlb_step_range(dom, collision, cp, &r, nullptr, f0, f1, bc_mask, missing_mask, bc_table, n_bc, omega, timestep, variant, stream);
}

int xlb_step_range(const XlbDomain* dom, int collision, const XlbCollisionParams* cp, const XlbStepRange* range, const XlbHaloPush* halo,
                   const void* f0, void* f1, const uint8_t* bc_mask, const uint8_t* missing_mask, const XlbBcEntry* bc_table, int n_bc,
                   double omega, int timestep, int variant, void* stream) {
    // ``timestep`` (nse_stepper.py:229; the JAX path never reads it) only chooses the ORDER in which the planes are walked: odd
    // steps go from the last plane to the first.  The planes a step wrote last are the ones still in the 126 MB L2; the next
    // step reads them first instead of evicting them before it gets there.  Results do not depend on the order.
    if (int rc = check_domain(dom, "xlb_step")) return rc;
    if (!f0 || !f1) return set_error(XLB_EINVAL, "xlb_step: null population pointer");
    if (f0 == f1) return set_error(XLB_EINVAL, "xlb_step: f0 and f1 must be distinct buffers (pull scheme, double buffered)");
    XlbStepRange r;
    if (range) {
        r = *range;
    } else {
        r.x_begin = 0, r.x_stride = 1, r.n_planes = dom->nx;
        r.y_begin = 0, r.y_stride = 1, r.n_rows = 0;
    }
    if (r.n_rows == 0) r.y_begin = 0, r.y_stride = 1, r.n_rows = dom->ny;
    if (r.n_planes < 0 || r.n_rows < 0 || r.x_stride < 1 || r.y_stride < 1) return set_error(XLB_EINVAL, "xlb_step: bad plane / row range");
    if (r.n_planes == 0 || r.n_rows == 0) return 0;
    const int x_last = r.x_begin + (r.n_planes - 1) * r.x_stride, y_last = r.y_begin + (r.n_rows - 1) * r.y_stride;
    if (r.x_begin < 0 || x_last >= dom->nx) return set_error(XLB_EINVAL, "xlb_step: bad x range");
    if (r.y_begin < 0 || y_last >= dom->ny) return set_error(XLB_EINVAL, "xlb_step: bad y range");
    if (!dom->x_periodic && (r.x_begin < 1 || x_last > dom->nx - 2) && dom->d == 3)
        return set_error(XLB_EINVAL, "xlb_step: non-periodic x needs ghost planes at x_begin-1 and x_end");
    StepParams p;
    std::memset(&p, 0, sizeof(p));
    p.f0 = f0;
    p.f1 = f1;
    const bool has_bc = n_bc > 0 && bc_mask && bc_table;
    if (n_bc > 0 && !has_bc) return set_error(XLB_EINVAL, "xlb_step: n_bc > 0 needs bc_mask and bc_table");
    p.bc_mask = has_bc ? bc_mask : nullptr;
    p.missing = has_bc ? missing_mask : nullptr;
    p.bct = has_bc ? bc_table : nullptr;
    p.nx = dom->nx;
    p.ny = dom->ny;
    p.nz = dom->nz;
    p.x_begin = r.x_begin;
    p.reverse = timestep & 1;
    p.x_end = x_last + 1;
    p.x_stride = r.x_stride;
    p.n_planes = r.n_planes;
    p.y_begin = r.y_begin;
    p.y_stride = r.y_stride;
    p.n_rows = r.n_rows;
    p.x_periodic = dom->x_periodic;
    const long long ncell = (long long)dom->nx * dom->ny * dom->nz;
    p.qs = dom->q_stride ? dom->q_stride : ncell;
    p.mqs = dom->mask_q_stride ? dom->mask_q_stride : ncell;
    p.omega = omega;
    p.bc_kind_mask = cp ? cp->bc_kind_mask : 0;
    p.halfway_fraction = cp ? (float)cp->halfway_fraction : 0.f;
    if (halo) {
        // slab axis: x planes for the 3-D lattices, rows (kernel y) for D2Q9
        const int extent = dom->d == 3 ? dom->nx : dom->ny;
        const long long unit = dom->d == 3 ? (long long)dom->ny * dom->nz : (long long)dom->nz;
        for (int k = 0; k < 2; ++k) {
            if (!halo->peer_f1[k]) continue;
            if (halo->src[k] < 0 || halo->src[k] >= extent || halo->dst[k] < 0) return set_error(XLB_EINVAL, "xlb_step: bad XlbHaloPush plane");
            if (halo->peer_f1[k] == f1 || halo->peer_f1[k] == f0) return set_error(XLB_EINVAL, "xlb_step: XlbHaloPush.peer_f1 must be the NEIGHBOUR's buffer");
            p.peer[k] = halo->peer_f1[k];
            p.peer_src[k] = halo->src[k];
            p.peer_off[k] = (long long)halo->dst[k] * unit;
            p.peer_qs[k] = halo->peer_q_stride[k] ? halo->peer_q_stride[k] : p.qs;
            if (p.peer_qs[k] < (long long)(halo->dst[k] + 1) * unit) return set_error(XLB_EINVAL, "xlb_step: XlbHaloPush.peer_q_stride is smaller than the ghost plane's offset");
        }
        if (p.peer[0] && p.peer[1] && p.peer_src[0] == p.peer_src[1])
            return set_error(XLB_EINVAL, "xlb_step: the slab must own at least two planes (first and last owned plane coincide)");
    }
    if (cp && cp->missing_bits && has_bc) {
        if (reinterpret_cast<uintptr_t>(cp->missing_bits) % 16 != 0) return set_error(XLB_EALIGN, "xlb_step: missing_bits must be 16-byte aligned");
        p.mbits = cp->missing_bits;
    }
    if (cp && cp->bc_cells) {
        if (cp->n_bc_cells < 0 || cp->n_bc_cells >= (1ll << 37)) return set_error(XLB_EINVAL, "xlb_step: bad n_bc_cells");
        p.cells = reinterpret_cast<const long long*>(cp->bc_cells);
        p.n_cells = cp->n_bc_cells;
    }
    if (collision != XLB_BGK && collision != XLB_KBC && collision != XLB_SMAGORINSKY_LES_BGK)
        return set_error(XLB_EINVAL, "xlb_step: unknown collision model");
    if (collision == XLB_SMAGORINSKY_LES_BGK) {
        const double cs = cp ? cp->smagorinsky_coef : 0.17;  // default of SmagorinskyLESBGK.__init__ (smagorinsky_les_bgk.py:39)
        p.les_c = 36.0 * cs * cs;
    }
    if (cp && cp->has_force) {
        for (int a = 0; a < 3; ++a) p.force3[a] = 3.0 * cp->force[a];
        if (dom->d == 2) p.force3[0] = 0.0;
        collision |= XLB_COLL_FORCED;
    }
    if (int rc = step_sched_slot(&p.sched)) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int key = dom->q * 100 + dom->compute_dtype * 10 + dom->store_dtype;
    switch (key) {
#define XLB_CASE(Q, C, S, tag) \
    case Q * 100 + C * 10 + S: return xlb_step_q##Q##_##tag(collision, p, variant, st);
#define XLB_CASES(Q)                        \
    XLB_CASE(Q, XLB_F32, XLB_F32, f32f32)   \
    XLB_CASE(Q, XLB_F32, XLB_F16, f32f16)   \
    XLB_CASE(Q, XLB_F64, XLB_F64, f64f64)   \
    XLB_CASE(Q, XLB_F64, XLB_F32, f64f32)   \
    XLB_CASE(Q, XLB_F64, XLB_F16, f64f16)
        XLB_CASES(9)
        XLB_CASES(19)
        XLB_CASES(27)
        default: return set_error(XLB_EUNSUPPORTED, "xlb_step: unsupported precision policy (precision_policy.py:57-97)");
    }
}

int xlb_step_inplace(const XlbDomain* dom, int collision, const XlbCollisionParams* cp, void* f, const uint8_t* bc_mask,
                     const XlbBcEntry* bc_table, int n_bc, double omega, int parity, void* stream) {
    if (int rc = check_domain(dom, "xlb_step_inplace")) return rc;
    if (!f) return set_error(XLB_EINVAL, "xlb_step_inplace: null population pointer");
    if (parity != 0 && parity != 1) return set_error(XLB_EINVAL, "xlb_step_inplace: parity must be 0 (P step) or 1 (L step)");
    if (!dom->x_periodic && dom->d == 3) return set_error(XLB_EUNSUPPORTED, "xlb_step_inplace: single periodic block only (no ghost planes)");
    const bool has_bc = n_bc > 0 && bc_mask && bc_table;
    if (n_bc > 0 && !has_bc) return set_error(XLB_EINVAL, "xlb_step_inplace: n_bc > 0 needs bc_mask and bc_table");
    if (has_bc) {
        const int allowed = (1 << XLB_BC_NONE) | (1 << XLB_BC_FULLWAY) | (1 << XLB_BC_EQUILIBRIUM);
        if (!cp || cp->bc_kind_mask == 0) return set_error(XLB_EINVAL, "xlb_step_inplace: cp->bc_kind_mask must list the BC kinds in bc_table");
        if (cp->bc_kind_mask & ~allowed)
            return set_error(XLB_EUNSUPPORTED, "xlb_step_inplace: only FullwayBounceBack and Equilibrium boundary conditions work on one buffer "
                                               "(half-way, Zou-He, regularized, extrapolation and do-nothing read f_pre or a neighbour)");
    }
    if (cp && cp->has_force) return set_error(XLB_EUNSUPPORTED, "xlb_step_inplace: the body-force collisions keep the two-buffer kernels");
    if (collision != XLB_BGK && collision != XLB_KBC && collision != XLB_SMAGORINSKY_LES_BGK)
        return set_error(XLB_EINVAL, "xlb_step_inplace: unknown collision model");
    StepParams p;
    std::memset(&p, 0, sizeof(p));
    p.f0 = f;
    p.f1 = f;
    p.bc_mask = has_bc ? bc_mask : nullptr;
    p.bct = has_bc ? bc_table : nullptr;
    p.nx = dom->nx, p.ny = dom->ny, p.nz = dom->nz;
    p.x_begin = 0, p.x_end = dom->nx, p.x_stride = 1, p.n_planes = dom->nx;
    p.reverse = parity;  // the L step walks the planes backwards: what the P step wrote last is read first (L2)
    p.y_begin = 0, p.y_stride = 1, p.n_rows = dom->ny;
    p.x_periodic = 1;
    const long long ncell = (long long)dom->nx * dom->ny * dom->nz;
    p.qs = dom->q_stride ? dom->q_stride : ncell;
    p.mqs = dom->mask_q_stride ? dom->mask_q_stride : ncell;
    p.omega = omega;
    p.bc_kind_mask = cp ? cp->bc_kind_mask : 0;
    if (collision == XLB_SMAGORINSKY_LES_BGK) {
        const double cs = cp ? cp->smagorinsky_coef : 0.17;
        p.les_c = 36.0 * cs * cs;
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int key = dom->q * 100 + dom->compute_dtype * 10 + dom->store_dtype;
    switch (key) {
#define XLB_AA_CASE(Q, C, S, tag) \
    case Q * 100 + C * 10 + S: return xlb_step_aa_q##Q##_##tag(collision, p, parity, st);
#define XLB_AA_CASES(Q)                        \
    XLB_AA_CASE(Q, XLB_F32, XLB_F32, f32f32)   \
    XLB_AA_CASE(Q, XLB_F32, XLB_F16, f32f16)   \
    XLB_AA_CASE(Q, XLB_F64, XLB_F64, f64f64)   \
    XLB_AA_CASE(Q, XLB_F64, XLB_F32, f64f32)   \
    XLB_AA_CASE(Q, XLB_F64, XLB_F16, f64f16)
        XLB_AA_CASES(9)
        XLB_AA_CASES(19)
        XLB_AA_CASES(27)
        default: return set_error(XLB_EUNSUPPORTED, "xlb_step_inplace: unsupported precision policy (precision_policy.py:57-97)");
    }
}

// ----------------------------------------------------------------------------- halo push ordering, peer mapping, batch copies
namespace xlb {
__global__ void halo_signal_kernel(unsigned long long* sent, unsigned long long* left_flag, unsigned long long* right_flag) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    const unsigned long long v = *sent + 1ull;
    *sent = v;
    // every store of the edge launch that preceded this kernel on the stream (the pushes into the neighbours' ghost planes
    // included) happens-before this fence; the release stores then publish "push number v has landed"
    __threadfence_system();
    if (left_flag) asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(left_flag), "l"(v) : "memory");
    if (right_flag) asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(right_flag), "l"(v) : "memory");
}
__global__ void halo_wait_kernel(unsigned long long* waited, const unsigned long long* from_left, const unsigned long long* from_right) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    const unsigned long long need = waited[0];
    const unsigned long long* flags[2] = {from_left, from_right};
    unsigned long long t0, t1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    for (int k = 0; k < 2; ++k) {
        if (!flags[k]) continue;
        unsigned long long v;
        do {
            asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(flags[k]) : "memory");
            if (v < need) {
                __nanosleep(64);
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
                if (t1 - t0 > 30000000000ull) {  // 30 s without the neighbour's signal: it is gone.  Record it, never hang the GPU
                    waited[1] = need;
                    break;
                }
            }
        } while (v < need);
    }
    waited[0] = need + 1ull;
}
}  // namespace xlb

int xlb_halo_signal(uint64_t* sent, uint64_t* left_neighbours_flag, uint64_t* right_neighbours_flag, void* stream) {
    if (!sent) return set_error(XLB_EINVAL, "xlb_halo_signal: null counter");
    halo_signal_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(reinterpret_cast<unsigned long long*>(sent),
                                                                          reinterpret_cast<unsigned long long*>(left_neighbours_flag),
                                                                          reinterpret_cast<unsigned long long*>(right_neighbours_flag));
    return count_launch(cudaGetLastError());
}

int xlb_halo_wait(uint64_t* waited, const uint64_t* flag_from_left, const uint64_t* flag_from_right, void* stream) {
    if (!waited) return set_error(XLB_EINVAL, "xlb_halo_wait: null counter");
    halo_wait_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(reinterpret_cast<unsigned long long*>(waited),
                                                                        reinterpret_cast<const unsigned long long*>(flag_from_left),
                                                                        reinterpret_cast<const unsigned long long*>(flag_from_right));
    return count_launch(cudaGetLastError());
}

static int cuda_rc(cudaError_t e, const char* what) {
    if (e == cudaSuccess) return 0;
    std::snprintf(g_err, sizeof(g_err), "%s: CUDA error %d: %s", what, (int)e, cudaGetErrorString(e));
    cudaGetLastError();  // clear the sticky-free error state
    return (int)e;
}

int xlb_ipc_export(const void* dptr, void* handle, int64_t* offset) {
    if (!dptr || !handle || !offset) return set_error(XLB_EINVAL, "xlb_ipc_export: null pointer");
    // base of the allocation that contains dptr: cuMemGetAddressRange through the runtime's driver entry point (no libcuda link)
    typedef int (*get_range_t)(unsigned long long*, size_t*, unsigned long long);
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qr;
    if (int rc = cuda_rc(cudaGetDriverEntryPoint("cuMemGetAddressRange", &fn, cudaEnableDefault, &qr), "xlb_ipc_export")) return rc;
    if (!fn || qr != cudaDriverEntryPointSuccess) return set_error(XLB_EUNSUPPORTED, "xlb_ipc_export: cuMemGetAddressRange is not available");
    unsigned long long base = 0;
    size_t size = 0;
    const int drc = reinterpret_cast<get_range_t>(fn)(&base, &size, (unsigned long long)reinterpret_cast<uintptr_t>(dptr));
    if (drc != 0) return set_error(XLB_EINVAL, "xlb_ipc_export: not a device allocation");
    cudaIpcMemHandle_t h;
    if (int rc = cuda_rc(cudaIpcGetMemHandle(&h, reinterpret_cast<void*>(base)), "xlb_ipc_export")) return rc;
    static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
    std::memcpy(handle, &h, sizeof(h));
    *offset = (int64_t)(reinterpret_cast<uintptr_t>(dptr) - base);
    return 0;
}

int xlb_ipc_import(const void* handle, int64_t offset, void** dptr, void** base) {
    if (!handle || !dptr || !base || offset < 0) return set_error(XLB_EINVAL, "xlb_ipc_import: bad argument");
    cudaIpcMemHandle_t h;
    std::memcpy(&h, handle, sizeof(h));
    void* b = nullptr;
    if (int rc = cuda_rc(cudaIpcOpenMemHandle(&b, h, cudaIpcMemLazyEnablePeerAccess), "xlb_ipc_import")) return rc;
    *base = b;
    *dptr = static_cast<char*>(b) + offset;
    return 0;
}

int xlb_ipc_close(void* base) {
    if (!base) return 0;
    return cuda_rc(cudaIpcCloseMemHandle(base), "xlb_ipc_close");
}

int xlb_copy_runs(void* dst, int64_t dst_stride_bytes, const void* src, int64_t src_stride_bytes, int64_t run_bytes, int n_runs, void* stream) {
    if (!dst || !src || run_bytes < 0 || n_runs < 0) return set_error(XLB_EINVAL, "xlb_copy_runs: bad argument");
    if (run_bytes == 0 || n_runs == 0) return 0;
    if (n_runs > 1 && (dst_stride_bytes < run_bytes || src_stride_bytes < run_bytes)) return set_error(XLB_EINVAL, "xlb_copy_runs: runs overlap");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    int dev = 0, max_pitch = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&max_pitch, cudaDevAttrMaxPitch, dev);
    if (n_runs == 1) return cuda_rc(cudaMemcpyAsync(dst, src, (size_t)run_bytes, cudaMemcpyDefault, st), "xlb_copy_runs");
    if (dst_stride_bytes <= max_pitch && src_stride_bytes <= max_pitch)  // one 2-D copy: row = run, pitch = stride
        return cuda_rc(cudaMemcpy2DAsync(dst, (size_t)dst_stride_bytes, src, (size_t)src_stride_bytes, (size_t)run_bytes, (size_t)n_runs, cudaMemcpyDefault, st),
                       "xlb_copy_runs");
    for (int r = 0; r < n_runs; ++r)  // strides beyond the 2-D copy's pitch limit (e.g. 1024^3 populations): run by run
        if (int rc = cuda_rc(cudaMemcpyAsync(static_cast<char*>(dst) + r * dst_stride_bytes, static_cast<const char*>(src) + r * src_stride_bytes,
                                             (size_t)run_bytes, cudaMemcpyDefault, st), "xlb_copy_runs"))
            return rc;
    return 0;
}

int xlb_stream(const XlbDomain* dom, int dtype, const void* f_in, void* f_out, void* stream) {
    if (int rc = check_domain(dom, "xlb_stream")) return rc;
    if (!f_in || !f_out || f_in == f_out) return set_error(XLB_EINVAL, "xlb_stream: need two distinct buffers");
    OpParams p = make_op(dom);
    p.a = f_in;
    p.o0 = f_out;
    p.dt_in = p.dt_out = dtype;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const unsigned nb = blocks_for(p.ncell, 256);
    if (dom->q == 9) stream_kernel<9><<<nb, 256, 0, st>>>(p);
    else if (dom->q == 19) stream_kernel<19><<<nb, 256, 0, st>>>(p);
    else stream_kernel<27><<<nb, 256, 0, st>>>(p);
    return count_launch(cudaGetLastError());
}

int xlb_macroscopic(const XlbDomain* dom, int f_dtype, const void* f, int out_dtype, void* rho, void* u, void* stream) {
    if (int rc = check_domain(dom, "xlb_macroscopic")) return rc;
    if (!f) return set_error(XLB_EINVAL, "xlb_macroscopic: null f");
    OpParams p = make_op(dom);
    p.a = f;
    p.o0 = rho;
    p.o1 = u;
    p.dt_in = f_dtype;
    p.dt_out = out_dtype;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const unsigned nb = blocks_for(p.ncell, 256);
    XLB_DISPATCH_TQ(dom, (macroscopic_kernel<T, Q><<<nb, 256, 0, st>>>(p)));
    return count_launch(cudaGetLastError());
}

int xlb_second_moment(const XlbDomain* dom, int f_dtype, const void* f, int out_dtype, void* pi, void* stream) {
    if (int rc = check_domain(dom, "xlb_second_moment")) return rc;
    if (!f || !pi) return set_error(XLB_EINVAL, "xlb_second_moment: null pointer");
    OpParams p = make_op(dom);
    p.a = f;
    p.o0 = pi;
    p.dt_in = f_dtype;
    p.dt_out = out_dtype;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const unsigned nb = blocks_for(p.ncell, 256);
    XLB_DISPATCH_TQ(dom, (second_moment_kernel<T, Q><<<nb, 256, 0, st>>>(p)));
    return count_launch(cudaGetLastError());
}

int xlb_equilibrium(const XlbDomain* dom, int in_dtype, const void* rho, const void* u, int f_dtype, void* f, void* stream) {
    if (int rc = check_domain(dom, "xlb_equilibrium")) return rc;
    if (!rho || !u || !f) return set_error(XLB_EINVAL, "xlb_equilibrium: null pointer");
    OpParams p = make_op(dom);
    p.a = rho;
    p.b = u;
    p.o0 = f;
    p.dt_in = in_dtype;
    p.dt_out = f_dtype;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const unsigned nb = blocks_for(p.ncell, 256);
    XLB_DISPATCH_TQ(dom, (equilibrium_kernel<T, Q><<<nb, 256, 0, st>>>(p)));
    return count_launch(cudaGetLastError());
}

int xlb_collide(const XlbDomain* dom, int collision, int dtype, const void* f, const void* feq, void* fout, double omega, void* stream) {
    return xlb_collide_ex(dom, collision, nullptr, dtype, f, feq, fout, omega, stream);
}

int xlb_collide_ex(const XlbDomain* dom, int collision, const XlbCollisionParams* cp, int dtype, const void* f, const void* feq, void* fout,
                   double omega, void* stream) {
    if (int rc = check_domain(dom, "xlb_collide")) return rc;
    if (!f || !feq || !fout) return set_error(XLB_EINVAL, "xlb_collide: null pointer");
    if (collision == XLB_KBC && dom->q == 19)
        return set_error(XLB_EUNSUPPORTED, "KBC: velocity set not supported (needs D2Q9 or D3Q27; kbc.py:65-66)");
    if (collision != XLB_BGK && collision != XLB_KBC && collision != XLB_SMAGORINSKY_LES_BGK)
        return set_error(XLB_EINVAL, "xlb_collide: unknown collision model");
    OpParams p = make_op(dom);
    p.a = f;
    p.b = feq;
    p.o0 = fout;
    p.dt_in = p.dt_out = dtype;
    p.omega = omega;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const unsigned nb = blocks_for(p.ncell, 128);
    if (collision == XLB_BGK) {
        XLB_DISPATCH_TQ(dom, (collide_kernel<T, Q, XLB_BGK><<<nb, 128, 0, st>>>(p)));
    } else if (collision == XLB_SMAGORINSKY_LES_BGK) {
        p.les_cs = cp ? cp->smagorinsky_coef : 0.17;
        XLB_DISPATCH_TQ(dom, (collide_kernel<T, Q, XLB_SMAGORINSKY_LES_BGK><<<nb, 128, 0, st>>>(p)));
    } else if (dom->q == 9) {
        if (dom->compute_dtype == XLB_F64) collide_kernel<double, 9, XLB_KBC><<<nb, 128, 0, st>>>(p);
        else collide_kernel<float, 9, XLB_KBC><<<nb, 128, 0, st>>>(p);
    } else {
        if (dom->compute_dtype == XLB_F64) collide_kernel<double, 27, XLB_KBC><<<nb, 128, 0, st>>>(p);
        else collide_kernel<float, 27, XLB_KBC><<<nb, 128, 0, st>>>(p);
    }
    return count_launch(cudaGetLastError());
}

int xlb_apply_bc(const XlbDomain* dom, int dtype, const void* f_pre, const void* f_post, void* f_out, const uint8_t* bc_mask,
                 const uint8_t* missing_mask, const XlbBcEntry* bc_table, int id, int assemble, void* stream) {
    if (int rc = check_domain(dom, "xlb_apply_bc")) return rc;
    if (!f_pre || !f_post || !f_out || !bc_mask || !missing_mask || !bc_table) return set_error(XLB_EINVAL, "xlb_apply_bc: null pointer");
    if (id < 1 || id > 255) return set_error(XLB_EINVAL, "xlb_apply_bc: id must be in 1..255");
    OpParams p = make_op(dom);
    p.a = f_pre;
    p.b = f_post;
    p.o0 = f_out;
    p.bc_mask = bc_mask;
    p.missing = missing_mask;
    p.bct = bc_table;
    p.dt_in = p.dt_out = dtype;
    p.id = id;
    p.flag = assemble;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const unsigned nb = blocks_for(p.ncell, 128);
    XLB_DISPATCH_TQ(dom, (apply_bc_kernel<T, Q><<<nb, 128, 0, st>>>(p)));
    return count_launch(cudaGetLastError());
}

int xlb_momentum_transfer(const XlbDomain* dom, int dtype, const void* f_post_collision, const uint8_t* bc_mask, const uint8_t* missing_mask,
                          const XlbBcEntry* bc_table, int id, int x_begin, int x_end, double* force, void* stream) {
    if (int rc = check_domain(dom, "xlb_momentum_transfer")) return rc;
    if (!f_post_collision || !bc_mask || !missing_mask || !bc_table || !force) return set_error(XLB_EINVAL, "xlb_momentum_transfer: null pointer");
    if (id < 1 || id > 255) return set_error(XLB_EINVAL, "xlb_momentum_transfer: id must be in 1..255");
    if (x_begin < 0 || x_end > dom->nx || x_begin > x_end) return set_error(XLB_EINVAL, "xlb_momentum_transfer: bad x range");
    if (!dom->x_periodic && (x_begin < 1 || x_end > dom->nx - 1) && dom->d == 3)
        return set_error(XLB_EINVAL, "xlb_momentum_transfer: non-periodic x needs ghost planes at x_begin-1 and x_end");
    OpParams p = make_op(dom);
    p.a = f_post_collision;
    p.o0 = force;
    p.bc_mask = bc_mask;
    p.missing = missing_mask;
    p.bct = bc_table;
    p.dt_in = p.dt_out = dtype;
    p.id = id;
    p.flag = x_begin;
    p.flag2 = x_end;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const unsigned nb = blocks_for(p.ncell, 128);
    XLB_DISPATCH_TQ(dom, (momentum_transfer_kernel<T, Q><<<nb, 128, 0, st>>>(p)));
    return count_launch(cudaGetLastError());
}

static MaskParams make_mask(const XlbDomain* d, const int32_t origin[3], const int32_t gshape[3]) {
    MaskParams p;
    std::memset(&p, 0, sizeof(p));
    p.nx = d->nx;
    p.ny = d->ny;
    p.nz = d->nz;
    for (int a = 0; a < 3; ++a) {
        p.origin[a] = origin ? origin[a] : 0;
        p.gshape[a] = gshape ? gshape[a] : (a == 0 ? d->nx : (a == 1 ? d->ny : d->nz));
    }
    const long long ncell = (long long)d->nx * d->ny * d->nz;
    p.mqs = d->mask_q_stride ? d->mask_q_stride : ncell;
    return p;
}

int xlb_mask_tag(const XlbDomain* dom, const int32_t* indices, int64_t n, const int32_t origin[3], const int32_t global_shape[3], int id,
                   int mark_solid, int pad, uint8_t* bc_mask, uint8_t* solid, void* stream) {
    if (int rc = check_domain(dom, "xlb_mask_tag")) return rc;
    if (n == 0) return 0;
    if (!indices) return set_error(XLB_EINVAL, "xlb_mask_tag: null indices");
    if (id > 255) return set_error(XLB_EINVAL, "xlb_mask_tag: id must fit a uint8");
    MaskParams p = make_mask(dom, origin, global_shape);
    p.idx = indices;
    p.n = n;
    p.id = id;
    p.mark_solid = mark_solid;
    p.pad = pad;
    p.bc_mask = bc_mask;
    p.solid = solid;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const long long total = pad ? n * dom->q : n;
    const unsigned nb = blocks_for(total, 256);
    if (dom->q == 9) mask_tag_kernel<9><<<nb, 256, 0, st>>>(p);
    else if (dom->q == 19) mask_tag_kernel<19><<<nb, 256, 0, st>>>(p);
    else mask_tag_kernel<27><<<nb, 256, 0, st>>>(p);
    return count_launch(cudaGetLastError());
}

int xlb_mask_missing(const XlbDomain* dom, const uint8_t* solid, const int32_t origin[3], const int32_t global_shape[3],
                     uint8_t* missing_mask, void* stream) {
    if (int rc = check_domain(dom, "xlb_mask_missing")) return rc;
    if (!missing_mask) return set_error(XLB_EINVAL, "xlb_mask_missing: null missing_mask");
    MaskParams p = make_mask(dom, origin, global_shape);
    p.solid_in = solid;
    p.missing = missing_mask;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const long long ncell = (long long)dom->nx * dom->ny * dom->nz;
    const unsigned nb = blocks_for(ncell, 256);
    if (dom->q == 9) mask_missing_kernel<9><<<nb, 256, 0, st>>>(p);
    else if (dom->q == 19) mask_missing_kernel<19><<<nb, 256, 0, st>>>(p);
    else mask_missing_kernel<27><<<nb, 256, 0, st>>>(p);
    return count_launch(cudaGetLastError());
}

int xlb_pack_missing(const XlbDomain* dom, const uint8_t* missing_mask, uint32_t* missing_bits, void* stream) {
    if (int rc = check_domain(dom, "xlb_pack_missing")) return rc;
    if (!missing_mask || !missing_bits) return set_error(XLB_EINVAL, "xlb_pack_missing: null pointer");
    const long long ncell = (long long)dom->nx * dom->ny * dom->nz;
    const long long mqs = dom->mask_q_stride ? dom->mask_q_stride : ncell;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const unsigned nb = blocks_for(ncell, 256);
    if (dom->q == 9) pack_missing_kernel<9><<<nb, 256, 0, st>>>(missing_mask, mqs, ncell, missing_bits);
    else if (dom->q == 19) pack_missing_kernel<19><<<nb, 256, 0, st>>>(missing_mask, mqs, ncell, missing_bits);
    else pack_missing_kernel<27><<<nb, 256, 0, st>>>(missing_mask, mqs, ncell, missing_bits);
    return count_launch(cudaGetLastError());
}

int xlb_list_bc_cells(const XlbDomain* dom, const uint8_t* bc_mask, const XlbBcEntry* bc_table, int kind_mask, int x_begin, int x_end,
                      int64_t* cells, int64_t capacity, int64_t* count, void* stream) {
    if (int rc = check_domain(dom, "xlb_list_bc_cells")) return rc;
    if (!bc_mask || !bc_table || !count) return set_error(XLB_EINVAL, "xlb_list_bc_cells: null pointer");
    if (x_begin < 0 || x_end > dom->nx || x_begin > x_end) return set_error(XLB_EINVAL, "xlb_list_bc_cells: bad x range");
    if (capacity < 0 || (cells == nullptr && capacity != 0)) return set_error(XLB_EINVAL, "xlb_list_bc_cells: bad capacity");
    if (reinterpret_cast<uintptr_t>(bc_mask) % 16 != 0) return set_error(XLB_EALIGN, "xlb_list_bc_cells: bc_mask must be 16-byte aligned");
    if (x_begin == x_end) return 0;
    ListParams p;
    std::memset(&p, 0, sizeof(p));
    const long long plane = (long long)dom->ny * dom->nz;
    p.bc_mask = bc_mask;
    p.bct = bc_table;
    p.first = x_begin * plane;
    p.last = x_end * plane;
    p.ncell = dom->nx * plane;
    p.kind_mask = kind_mask ? kind_mask : XLB_BC_LISTED_KINDS;
    p.cells = reinterpret_cast<long long*>(cells);
    p.capacity = capacity;
    p.count = reinterpret_cast<unsigned long long*>(count);
    const long long span = p.last - (p.first & ~15ll);
    const long long nthreads = (span + 15) / 16;
    list_bc_cells_kernel<<<blocks_for(nthreads, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(p);
    return count_launch(cudaGetLastError());
}

int xlb_mask_mesh_ray(const XlbDomain* dom, const float* triangles, int64_t n_triangles, const int32_t origin[3],
                      const int32_t global_shape[3], int id, uint8_t* bc_mask, uint8_t* missing_mask, uint64_t* distance_keys, void* stream) {
    if (int rc = check_domain(dom, "xlb_mask_mesh_ray")) return rc;
    if (dom->d != 3) return set_error(XLB_EUNSUPPORTED, "xlb_mask_mesh_ray: mesh maskers are 3-D only (mesh_boundary_masker.py:36-39)");
    if (n_triangles < 0 || id < 1 || id > 255) return set_error(XLB_EINVAL, "xlb_mask_mesh_ray: bad triangle count or id");
    if (!bc_mask || !missing_mask || (n_triangles > 0 && !triangles)) return set_error(XLB_EINVAL, "xlb_mask_mesh_ray: null pointer");
    MeshParams p;
    std::memset(&p, 0, sizeof(p));
    p.tri = triangles;
    p.ntri = n_triangles;
    p.nx = dom->nx;
    p.ny = dom->ny;
    p.nz = dom->nz;
    for (int a = 0; a < 3; ++a) {
        p.origin[a] = origin ? origin[a] : 0;
        p.gshape[a] = global_shape ? global_shape[a] : (a == 0 ? dom->nx : (a == 1 ? dom->ny : dom->nz));
    }
    p.id = id;
    p.bc_mask = bc_mask;
    p.missing = missing_mask;
    const long long ncell = (long long)dom->nx * dom->ny * dom->nz;
    p.mqs = dom->mask_q_stride ? dom->mask_q_stride : ncell;
    p.dist_key = reinterpret_cast<unsigned long long*>(distance_keys);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (n_triangles > 0) {
        const unsigned nb = blocks_for(n_triangles * 32, 128);
        if (dom->q == 19) mesh_ray_kernel<19><<<nb, 128, 0, st>>>(p);
        else mesh_ray_kernel<27><<<nb, 128, 0, st>>>(p);
        if (int rc = count_launch(cudaGetLastError())) return rc;
    }
    const unsigned nb2 = blocks_for(ncell, 256);
    if (dom->q == 19) mesh_resolve_oob_kernel<19><<<nb2, 256, 0, st>>>(p);
    else mesh_resolve_oob_kernel<27><<<nb2, 256, 0, st>>>(p);
    return count_launch(cudaGetLastError());
}


// ----------------------------------------------------------------------------- allocation-free indices masker
// The per-cell ``solid`` scratch of the two masker passes, inside the library (see include/xlb_b200.h).  In HOST code the
// name g_mask_solid_scratch is only a stand-in for the symbol: the device address is asked for with cudaGetSymbolAddress,
// once per device with that device current, and that pointer is what the kernels get.
}  // extern "C"
static __device__ __align__(256) uint8_t g_mask_solid_scratch[XLB_MASK_STATIC_MAX_CELLS];

static int mask_scratch_address(uint8_t** out) {
    constexpr int MAX_DEV = 64;
    static std::mutex mu;
    static uint8_t* cached[MAX_DEV] = {nullptr};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return count_launch(e);
    if (dev < 0 || dev >= MAX_DEV) return set_error(XLB_EUNSUPPORTED, "xlb_mask_indices_static: device ordinal beyond 63");
    std::lock_guard<std::mutex> lock(mu);
    if (cached[dev] == nullptr) {
        void* ptr = nullptr;
        e = cudaGetSymbolAddress(&ptr, g_mask_solid_scratch);  // the current device's instance of the array
        if (e != cudaSuccess || ptr == nullptr) {
            std::snprintf(g_err, sizeof(g_err), "xlb_mask_indices_static: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e));
            cudaGetLastError();
            return e != cudaSuccess ? (int)e : XLB_EINVAL;
        }
        cached[dev] = static_cast<uint8_t*>(ptr);
    }
    *out = cached[dev];
    return 0;
}
extern "C" {

int xlb_mask_indices_static(const XlbDomain* dom, const int32_t* indices, int64_t n, const int32_t origin[3], const int32_t global_shape[3],
                            int id, int mark_solid, int pad, int begin, int finish, uint8_t* bc_mask, uint8_t* missing_mask, void* stream) {
    if (int rc = check_domain(dom, "xlb_mask_indices_static")) return rc;
    const long long ncell = (long long)dom->nx * dom->ny * dom->nz;
    if (ncell > (long long)XLB_MASK_STATIC_MAX_CELLS) {
        std::snprintf(g_err, sizeof(g_err), "xlb_mask_indices_static: block of %lld cells exceeds the static scratch of %d cells (XLB_MASK_STATIC_MAX_CELLS); "
                      "use xlb_mask_tag / xlb_mask_missing with a caller-allocated solid array", ncell, (int)XLB_MASK_STATIC_MAX_CELLS);
        return XLB_EINVAL;
    }
    if (n < 0 || (n > 0 && !indices)) return set_error(XLB_EINVAL, "xlb_mask_indices_static: null indices");
    if (n > 0 && !bc_mask && id >= 0) return set_error(XLB_EINVAL, "xlb_mask_indices_static: null bc_mask");
    if (id > 255) return set_error(XLB_EINVAL, "xlb_mask_indices_static: id must fit a uint8");
    if (finish && !missing_mask) return set_error(XLB_EINVAL, "xlb_mask_indices_static: finish needs missing_mask");
    uint8_t* scratch = nullptr;
    if (int rc = mask_scratch_address(&scratch)) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (begin) {  // the array keeps the last call's contents: everything the missing pass reads must start at zero
        cudaError_t e = cudaMemsetAsync(scratch, 0, (size_t)ncell, st);
        if (e != cudaSuccess) return count_launch(e);
    }
    if (n > 0) {
        MaskParams p = make_mask(dom, origin, global_shape);
        p.idx = indices;
        p.n = n;
        p.id = id;
        p.mark_solid = mark_solid;
        p.pad = pad;
        p.bc_mask = bc_mask;
        p.solid = scratch;
        const long long total = pad ? n * dom->q : n;
        const unsigned nb = blocks_for(total, 256);
        if (dom->q == 9) mask_tag_kernel<9><<<nb, 256, 0, st>>>(p);
        else if (dom->q == 19) mask_tag_kernel<19><<<nb, 256, 0, st>>>(p);
        else mask_tag_kernel<27><<<nb, 256, 0, st>>>(p);
        if (int rc = count_launch(cudaGetLastError())) return rc;
    }
    if (finish) {
        MaskParams p = make_mask(dom, origin, global_shape);
        p.solid_in = scratch;
        p.missing = missing_mask;
        const unsigned nb = blocks_for(ncell, 256);
        if (dom->q == 9) mask_missing_kernel<9><<<nb, 256, 0, st>>>(p);
        else if (dom->q == 19) mask_missing_kernel<19><<<nb, 256, 0, st>>>(p);
        else mask_missing_kernel<27><<<nb, 256, 0, st>>>(p);
        if (int rc = count_launch(cudaGetLastError())) return rc;
    }
    return 0;
}

}  // extern "C"
