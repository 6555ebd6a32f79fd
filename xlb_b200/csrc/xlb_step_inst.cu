// xlb_step_inst.cu -- one translation unit per (lattice, precision policy); compiled many times by
// build.py with -DXLB_Q=<9|19|27> -DXLB_TC=<compute type> -DXLB_TS=<store type> -DXLB_TAG=<suffix>.
// Exposes   int xlb_step_<tag>(const XlbDomain*, int collision, StepParams, int variant, cudaStream_t)
#include <mutex>

#include "xlb_host.cuh"
#include "xlb_step.cuh"
#include "xlb_step_bulk.cuh"
#include "xlb_step_h2.cuh"
#include "xlb_step_aa.cuh"

#ifndef XLB_Q
#error "compile with -DXLB_Q -DXLB_TC -DXLB_TS -DXLB_TAG"
#endif

// tuning knobs (overridable from the command line for sweeps; see tools/sweep.py)
// 1: FP32-compute / FP16-store instances run step_vector_h2_kernel (populations packed as half2 from load to store)
#ifndef XLB_TUNE_H2
#define XLB_TUNE_H2 1
#endif
#ifndef XLB_TUNE_BLOCK
#define XLB_TUNE_BLOCK 128
#else
#define XLB_TUNE_BLOCK_SET 1
#endif
// default (V, MINB) per instance from the B200 sweeps recorded in profiles/ (tools/sweep.py)
// FP16 storage (packed-half2 kernel): 8 cells per thread = one 128-bit load per population at 12 warps per SM
// (profiles/sweep9_f16_h2.log, sweep10_f16_h2_v8.log); rows that are only a multiple of 4 cells wide run V = 4.
template <typename T, typename S, int Q>
constexpr int default_vec() { return sizeof(S) == 8 ? 2 : (Q == 27 ? 2 : ((sizeof(S) == 2 && sizeof(T) == 4 && XLB_TUNE_H2) ? 8 : 4)); }
template <typename S, int Q, int V>
constexpr int default_minb() { return sizeof(S) == 8 ? 4 : (Q == 27 ? 4 : (sizeof(S) == 2 ? 5 : 4)); }  // FP16: 5 blocks of 64 (V = 8) or of 128 (V = 4)

namespace xlb {

template <typename T, typename S, int Q, int COLL, int V, bool HAS_BC, bool HALF = false, bool PUSH = false>
static int launch_vector(const StepParams& p0, cudaStream_t st) {
    constexpr bool USE_H2 = XLB_TUNE_H2 && std::is_same<S, __half>::value && std::is_same<T, float>::value && V >= 2;
#ifdef XLB_TUNE_BLOCK_SET
    constexpr int BLOCK = XLB_TUNE_BLOCK;
#else
    constexpr int BLOCK = (USE_H2 && V == 8) ? 64 : XLB_TUNE_BLOCK;  // finer block granularity at 168-204 registers per thread
#endif
    StepParams p = p0;
    p.ntz = p.nz / V;
    p.ntz_shift = -1;
    for (int b = 0; b < 31; ++b)
        if ((1 << b) == p.ntz) p.ntz_shift = b;
    const long long nvec = (long long)p.n_rows * p.ntz;
    dim3 grid((unsigned)((nvec + BLOCK - 1) / BLOCK), (unsigned)p.n_planes, 1);
#ifdef XLB_TUNE_MINB
    constexpr int MINB = XLB_TUNE_MINB;
#else
    constexpr int MINB = (V == 1) ? 4 : default_minb<S, Q, V>();
#endif
    if constexpr (USE_H2) {
        step_vector_h2_kernel<Q, COLL, V, BLOCK, MINB, HAS_BC, HALF, PUSH><<<grid, BLOCK, 0, st>>>(p);
    } else {
        step_vector_kernel<T, S, Q, COLL, V, BLOCK, MINB, HAS_BC, HALF, PUSH><<<grid, BLOCK, 0, st>>>(p);
    }
    return count_launch(cudaGetLastError());
}

#ifndef XLB_TUNE_STAGES
#define XLB_TUNE_STAGES 2
#endif
#ifndef XLB_TUNE_BMINB
#define XLB_TUNE_BMINB 3
#endif
#ifndef XLB_TUNE_C
#define XLB_TUNE_C 2
#endif

// persistent bulk-copy-staged kernel (xlb_step_bulk.cuh); returns -1000 when the shape does not qualify
// FP16 storage: packed-half2 consumer of the staged rows (H2), tile of C = 4 cells per thread (1 KB rows) where nz allows
#ifndef XLB_TUNE_HC
#define XLB_TUNE_HC 4
#endif
#ifndef XLB_TUNE_HSTAGES
#define XLB_TUNE_HSTAGES 3
#endif
#ifndef XLB_TUNE_HBMINB
#define XLB_TUNE_HBMINB 3
#endif

#ifndef XLB_TUNE_MULTIROW
#define XLB_TUNE_MULTIROW 1
#endif

template <typename T, typename S, int Q, int COLL, int C, int NT, bool HAS_BC, bool HALF = false, bool H2 = false, bool MULTI = false>
static int launch_bulk(const StepParams& p, cudaStream_t st) {
    constexpr int TZ = NT * C;
    constexpr int STAGES = H2 ? XLB_TUNE_HSTAGES : XLB_TUNE_STAGES;
    using L = BulkLayout<S, Q, TZ, HAS_BC>;
    constexpr int BM0 = H2 ? XLB_TUNE_HBMINB : XLB_TUNE_BMINB;
    constexpr int BMINB = (BM0 * 128 / NT) > 0 ? (BM0 * 128 / NT) : 1;  // same register budget per thread for every NT
    auto kern = step_bulk_kernel<T, S, Q, COLL, C, NT, STAGES, BMINB, HAS_BC, HALF, H2, MULTI>;
    // the shared-memory opt-in and the occupancy are per device: cached per device ordinal (one slot each, filled under a lock)
    constexpr int MAX_DEV = 64;
    static std::mutex mu;
    static int cached_blocks[MAX_DEV] = {0}, cached_sms[MAX_DEV] = {0};
    const size_t smem = L::smem_bytes(STAGES);
    int dev = 0;
    cudaGetDevice(&dev);
    int blocks_per_sm = 0, n_sm = 0;
    {
        std::lock_guard<std::mutex> lock(mu);
        const int slot = dev < MAX_DEV ? dev : MAX_DEV - 1;
        if (cached_blocks[slot] == 0 || dev >= MAX_DEV) {
            cudaDeviceGetAttribute(&cached_sms[slot], cudaDevAttrMultiProcessorCount, dev);
            cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            if (e != cudaSuccess) return count_launch(e);
            e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&cached_blocks[slot], kern, NT, smem);
            if (e != cudaSuccess) return count_launch(e);
            if (cached_blocks[slot] < 1) cached_blocks[slot] = 1;
        }
        blocks_per_sm = cached_blocks[slot];
        n_sm = cached_sms[slot];
    }
    constexpr bool multi = MULTI;  // multi-row tiles (xlb_step_bulk.cuh): C rows of NT cells per tile, p.nz == NT
    if (multi && (p.nz != NT || p.n_rows % C != 0 || p.y_stride != 1)) return -1000;
    const long long total = multi ? (long long)p.n_planes * (p.n_rows / C) : (long long)p.n_planes * p.n_rows * (p.nz / TZ);
    if (total >= (1ll << 30)) return -1000;  // tile indices are 32-bit in the kernel
    long long grid = (long long)blocks_per_sm * n_sm;
    if (grid > total) grid = total;
    kern<<<(unsigned)grid, NT, smem, st>>>(p);
    return count_launch(cudaGetLastError());
}

template <typename T, typename S, int Q, int COLL, bool HAS_BC, bool HALF = false>
static int try_bulk(const StepParams& p, cudaStream_t st) {
    constexpr int C = XLB_TUNE_C;
    const uintptr_t a0 = reinterpret_cast<uintptr_t>(p.f0), am = reinterpret_cast<uintptr_t>(p.bc_mask);
    bool aligned = (a0 % 16 == 0) && ((p.qs * sizeof(S)) % 16 == 0) && ((p.nz * sizeof(S)) % 16 == 0) && (!HAS_BC || am % 16 == 0);
    // the half-way path of the packed-half2 consumer reads the missing bytes of a cell pair with one 16-bit load
    if (HAS_BC && sizeof(S) == 2) aligned = aligned && reinterpret_cast<uintptr_t>(p.missing) % 2 == 0 && p.mqs % 2 == 0;
    if (HAS_BC) aligned = aligned && reinterpret_cast<uintptr_t>(p.mbits) % 8 == 0;
    if (p.peer[0] || p.peer[1]) return -1000;  // halo push: the vector kernel's PUSH instance
    if (!aligned) return -1000;
    if constexpr (XLB_TUNE_H2 && std::is_same<S, __half>::value && std::is_same<T, float>::value) {
        if (p.nz % (128 * XLB_TUNE_HC) == 0) return launch_bulk<T, S, Q, COLL, XLB_TUNE_HC, 128, HAS_BC, HALF, true>(p, st);
        if (XLB_TUNE_HC != 2 && p.nz % 256 == 0) return launch_bulk<T, S, Q, COLL, 2, 128, HAS_BC, HALF, true>(p, st);
        return -1000;
    }
#ifdef XLB_TUNE_NT
    if (p.nz % (XLB_TUNE_NT * C) == 0) return launch_bulk<T, S, Q, COLL, C, XLB_TUNE_NT, HAS_BC, HALF>(p, st);
#else
    if (p.nz % (128 * C) == 0) return launch_bulk<T, S, Q, COLL, C, 128, HAS_BC, HALF>(p, st);
#if XLB_TUNE_MULTIROW
    // rows of exactly 128 cells (BASELINE configs[2] is 128 wide): tiles of C whole rows on the 128-thread instance
    if constexpr (C > 1) {
        if (p.nz == 128 && p.n_rows % C == 0 && p.y_stride == 1) return launch_bulk<T, S, Q, COLL, C, 128, HAS_BC, HALF, false, true>(p, st);
    }
#endif
    if constexpr (Q == 27 && COLL == XLB_KBC) {  // narrow grids
        if (p.nz % (64 * C) == 0) return launch_bulk<T, S, Q, COLL, C, 64, HAS_BC, HALF>(p, st);
    }
#endif
    return -1000;
}

template <typename T, typename S, int Q, int COLL>
static int launch_generic(const StepParams& p, cudaStream_t st) {
    const long long row = (long long)p.n_rows * p.nz;
    dim3 grid((unsigned)((row + 127) / 128), (unsigned)p.n_planes, 1);
    step_generic_kernel<T, S, Q, COLL><<<grid, 128, 0, st>>>(p);
    return count_launch(cudaGetLastError());
}

template <typename T, typename S, int Q, int COLL>
static int step_main(const StepParams& p, int variant, cudaStream_t st);

// the fused kernel, then (when the caller passed XlbCollisionParams.bc_cells) the dense launch over the BC cell list
template <typename T, typename S, int Q, int COLL>
static int step_coll(const StepParams& p0, int variant, cudaStream_t st) {
    StepParams p = p0;
    const bool has_bc = p.bc_mask != nullptr && p.bct != nullptr;
    if (!has_bc || variant == XLB_VARIANT_GENERIC || p.n_cells <= 0) p.cells = nullptr;
    if (int rc = step_main<T, S, Q, COLL>(p, variant, st)) return rc;
    if (p.cells == nullptr) return 0;
    step_cells_kernel<T, S, Q, COLL><<<(unsigned)((p.n_cells + 127) / 128), 128, 0, st>>>(p);
    return count_launch(cudaGetLastError());
}

template <typename T, typename S, int Q, int COLL>
static int step_main(const StepParams& p, int variant, cudaStream_t st) {
#ifdef XLB_TUNE_V
    constexpr int VMAX = XLB_TUNE_V;
#else
    constexpr int VMAX = default_vec<T, S, Q>();
#endif
    const bool has_bc = p.bc_mask != nullptr && p.bct != nullptr;
    // half-way obstacles get the instance with the vectorised half-way path (unknown kind mask: assume present)
    const bool half = has_bc && (p.bc_kind_mask == 0 || (p.bc_kind_mask & (1 << XLB_BC_HALFWAY)));
    auto bulk = [&]() -> int {
        if constexpr (COLL == XLB_BGK || COLL == XLB_KBC) {
            if (!has_bc) return try_bulk<T, S, Q, COLL, false>(p, st);
            return half ? try_bulk<T, S, Q, COLL, true, true>(p, st) : try_bulk<T, S, Q, COLL, true, false>(p, st);
        } else {
            return -1000;
        }
    };
    if (variant == XLB_VARIANT_GENERIC) return launch_generic<T, S, Q, COLL>(p, st);
    const bool push = p.peer[0] != nullptr || p.peer[1] != nullptr;
    if (push) {
        // edge-plane launch of a slab with the halo push (XlbHaloPush): the vector kernel's PUSH instance where the shape allows
        // the full vector width, else the generic kernel (every variant performs the same arithmetic, bit for bit)
        const uintptr_t b0 = reinterpret_cast<uintptr_t>(p.f0), b1 = reinterpret_cast<uintptr_t>(p.f1);
        bool ok = (p.nz % VMAX == 0) && (p.qs % VMAX == 0) && (b0 % (VMAX * sizeof(S)) == 0) && (b1 % (VMAX * sizeof(S)) == 0);
        if (has_bc) ok = ok && reinterpret_cast<uintptr_t>(p.bc_mask) % VMAX == 0 && reinterpret_cast<uintptr_t>(p.missing) % VMAX == 0 && p.mqs % VMAX == 0 &&
                         reinterpret_cast<uintptr_t>(p.mbits) % (4 * VMAX) == 0;
        for (int k = 0; k < 2; ++k)
            if (p.peer[k]) ok = ok && reinterpret_cast<uintptr_t>(p.peer[k]) % (VMAX * sizeof(S)) == 0 && p.peer_off[k] % VMAX == 0 && p.peer_qs[k] % VMAX == 0;
        if (!ok || variant == XLB_VARIANT_SCALAR) return launch_generic<T, S, Q, COLL>(p, st);
        if (!has_bc) return launch_vector<T, S, Q, COLL, VMAX, false, false, true>(p, st);
        return half ? launch_vector<T, S, Q, COLL, VMAX, true, true, true>(p, st) : launch_vector<T, S, Q, COLL, VMAX, true, false, true>(p, st);
    }
    if (variant == XLB_VARIANT_BULK) {
        if constexpr (COLL == XLB_BGK || COLL == XLB_KBC) {
            const int rc = bulk();
            if (rc != -1000) return rc;
            return set_error(XLB_EALIGN, "xlb_step: the bulk variant needs nz to be a multiple of the 256-cell tile and 16-byte aligned rows");
        } else {
            return set_error(XLB_EUNSUPPORTED, "xlb_step: the bulk variant is built for plain BGK / KBC only");
        }
    }
    const uintptr_t a0 = reinterpret_cast<uintptr_t>(p.f0), a1 = reinterpret_cast<uintptr_t>(p.f1);
    // the vectorised paths move V bc_mask / missing_mask bytes and V populations per access: every base pointer and stride
    // must be aligned to the vector
    const uintptr_t am = reinterpret_cast<uintptr_t>(p.missing);
    auto masks_ok = [&](int v) {
        return !has_bc || (reinterpret_cast<uintptr_t>(p.bc_mask) % v == 0 && am % v == 0 && p.mqs % v == 0 &&
                           reinterpret_cast<uintptr_t>(p.mbits) % (4 * v) == 0);
    };
    const bool vec_ok = (p.nz % VMAX == 0) && (p.qs % VMAX == 0) && (a0 % (VMAX * sizeof(S)) == 0) && (a1 % (VMAX * sizeof(S)) == 0) &&
                        masks_ok(VMAX);
    if (variant == XLB_VARIANT_VECTOR && !vec_ok && !(VMAX == 8 && p.nz % 4 == 0)) return set_error(XLB_EALIGN, "xlb_step: vector variant needs nz, q_stride and base pointers aligned to the vector width");
    // AUTO: the bulk-staged pipeline where the B200 measurements show it ahead of the vector kernel: D3Q27 with 4-byte storage,
    // KBC (compute-heavy) and BGK (384x256x512: 30.4k vs 28.2k MLUPS, profiles/r2u_bench_more.jsonl).  D3Q19 stays on the
    // vector kernel (512^3: 42.4k vs 40.9k; 128^3 equal).
    if constexpr (Q == 27 && (COLL == XLB_KBC || COLL == XLB_BGK) && sizeof(S) == 4) {
        // (round 1 fell back to the vector kernel on blocks where a half-way obstacle covers more than ~1 % of the cells; with the
        // dynamic tile scheduler and the packed missing bits the pipeline is ahead there too: 22.7k vs 21.1k MLUPS on a
        // 192x1536x1536 block holding half of an r = 192 sphere)
        if (variant == XLB_VARIANT_AUTO) {
            const int rc = bulk();
            if (rc != -1000) return rc;
        }
    }
    if (vec_ok && variant != XLB_VARIANT_SCALAR) {
        if (!has_bc) return launch_vector<T, S, Q, COLL, VMAX, false>(p, st);
        return half ? launch_vector<T, S, Q, COLL, VMAX, true, true>(p, st) : launch_vector<T, S, Q, COLL, VMAX, true, false>(p, st);
    }
    if constexpr (VMAX == 8) {  // FP16 rows that are a multiple of 4 but not of 8 cells
        constexpr int V4 = 4;
        const bool v4_ok = (p.nz % V4 == 0) && (p.qs % V4 == 0) && (a0 % (V4 * sizeof(S)) == 0) && (a1 % (V4 * sizeof(S)) == 0) && masks_ok(V4);
        if (v4_ok && variant != XLB_VARIANT_SCALAR) {
            if (!has_bc) return launch_vector<T, S, Q, COLL, V4, false>(p, st);
            return half ? launch_vector<T, S, Q, COLL, V4, true, true>(p, st) : launch_vector<T, S, Q, COLL, V4, true, false>(p, st);
        }
    }
    return has_bc ? launch_vector<T, S, Q, COLL, 1, true>(p, st) : launch_vector<T, S, Q, COLL, 1, false>(p, st);
}

// ---- in-place (AA) step, xlb_step_aa.cuh: P step (local_step == 0) or L step
template <typename T, typename S, int Q, int COLL, int V, bool HAS_BC, bool LOCAL>
static int launch_aa(const StepParams& p0, cudaStream_t st) {
    constexpr int BLOCK = 128;
#ifdef XLB_TUNE_AA_MINB
    constexpr int MINB = XLB_TUNE_AA_MINB;
#else
    constexpr int MINB = (V == 1) ? 4 : default_minb<S, Q, V>();
#endif
    StepParams p = p0;
    p.ntz = p.nz / V;
    p.ntz_shift = -1;
    for (int b = 0; b < 31; ++b)
        if ((1 << b) == p.ntz) p.ntz_shift = b;
    const long long nvec = (long long)p.n_rows * p.ntz;
    dim3 grid((unsigned)((nvec + BLOCK - 1) / BLOCK), (unsigned)p.n_planes, 1);
    step_aa_kernel<T, S, Q, COLL, V, BLOCK, MINB, HAS_BC, LOCAL><<<grid, BLOCK, 0, st>>>(p);
    return count_launch(cudaGetLastError());
}

template <typename T, typename S, int Q, int COLL>
static int step_aa(const StepParams& p, int local_step, cudaStream_t st) {
    // FP16 storage: 4 cells per thread here (the 8-cell packed-half2 layout belongs to the two-buffer kernel)
    constexpr int VMAX = (sizeof(S) == 2) ? 4 : default_vec<T, S, Q>();
    const bool has_bc = p.bc_mask != nullptr && p.bct != nullptr;
    const uintptr_t a = reinterpret_cast<uintptr_t>(p.f1);
    const bool vec_ok = (p.nz % VMAX == 0) && (p.qs % VMAX == 0) && (a % (VMAX * sizeof(S)) == 0) &&
                        (!has_bc || reinterpret_cast<uintptr_t>(p.bc_mask) % VMAX == 0);
#define XLB_AA_GO(V)                                                                                                        \
    do {                                                                                                                    \
        if (has_bc) return local_step ? launch_aa<T, S, Q, COLL, V, true, true>(p, st) : launch_aa<T, S, Q, COLL, V, true, false>(p, st);  \
        return local_step ? launch_aa<T, S, Q, COLL, V, false, true>(p, st) : launch_aa<T, S, Q, COLL, V, false, false>(p, st);            \
    } while (0)
    if (vec_ok) XLB_AA_GO(VMAX);
    XLB_AA_GO(1);
#undef XLB_AA_GO
}

}  // namespace xlb

#define XLB_CAT2(a, b) a##b
#define XLB_CAT(a, b) XLB_CAT2(a, b)

int XLB_CAT(xlb_step_, XLB_TAG)(int collision, const xlb::StepParams& p, int variant, cudaStream_t st) {
    using namespace xlb;
#ifdef XLB_STUB  // tuning builds compile only the instance under test
    (void)collision; (void)p; (void)variant; (void)st;
    return set_error(XLB_EUNSUPPORTED, "xlb_step: this (lattice, policy) instance was stubbed out of a tuning build");
#else
    // ``collision`` = model | XLB_COLL_FORCED (set by xlb_step_ex when a force vector is given)
    switch (collision) {
        case XLB_BGK: return step_coll<XLB_TC, XLB_TS, XLB_Q, XLB_BGK>(p, variant, st);
        case XLB_SMAGORINSKY_LES_BGK: return step_coll<XLB_TC, XLB_TS, XLB_Q, XLB_SMAGORINSKY_LES_BGK>(p, variant, st);
        case XLB_BGK | XLB_COLL_FORCED: return step_coll<XLB_TC, XLB_TS, XLB_Q, XLB_BGK | XLB_COLL_FORCED>(p, variant, st);
        case XLB_SMAGORINSKY_LES_BGK | XLB_COLL_FORCED: return step_coll<XLB_TC, XLB_TS, XLB_Q, XLB_SMAGORINSKY_LES_BGK | XLB_COLL_FORCED>(p, variant, st);
#if XLB_Q != 19
        case XLB_KBC: return step_coll<XLB_TC, XLB_TS, XLB_Q, XLB_KBC>(p, variant, st);
        case XLB_KBC | XLB_COLL_FORCED: return step_coll<XLB_TC, XLB_TS, XLB_Q, XLB_KBC | XLB_COLL_FORCED>(p, variant, st);
#endif
        default: break;
    }
    return set_error(XLB_EUNSUPPORTED, "xlb_step: collision model not supported for this velocity set (KBC needs D2Q9 or D3Q27; kbc.py:65-66)");
#endif
}

int XLB_CAT(xlb_step_aa_, XLB_TAG)(int collision, const xlb::StepParams& p, int local_step, cudaStream_t st) {
    using namespace xlb;
#ifdef XLB_STUB
    (void)collision; (void)p; (void)local_step; (void)st;
    return set_error(XLB_EUNSUPPORTED, "xlb_step_inplace: this (lattice, policy) instance was stubbed out of a tuning build");
#else
    switch (collision) {  // the body-force variants read their own output between two passes; they keep the two-buffer kernels
        case XLB_BGK: return step_aa<XLB_TC, XLB_TS, XLB_Q, XLB_BGK>(p, local_step, st);
        case XLB_SMAGORINSKY_LES_BGK: return step_aa<XLB_TC, XLB_TS, XLB_Q, XLB_SMAGORINSKY_LES_BGK>(p, local_step, st);
#if XLB_Q != 19
        case XLB_KBC: return step_aa<XLB_TC, XLB_TS, XLB_Q, XLB_KBC>(p, local_step, st);
#endif
        default: break;
    }
    return set_error(XLB_EUNSUPPORTED, "xlb_step_inplace: BGK, KBC (D2Q9 / D3Q27) and SmagorinskyLESBGK without a body force");
#endif
}
