// xlb_common.cuh -- lattice traits, typed vector access and shared parameter blocks.
//
// Lattice direction order is the reference's (xlb/velocity_set/d3q19.py:17-30, d3q27.py:17-32,
// d2q9.py:16-24): itertools.product([0,-1,1], repeat=3), x slowest; D3Q19 keeps |c|_1 <= 2.
// All tables are constexpr functions so that fully unrolled loops fold them into immediates.
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <type_traits>
#include <utility>

#include "../../include/xlb_b200.h"

#define XLB_HD __host__ __device__ __forceinline__
#define XLB_D __device__ __forceinline__

namespace xlb {

// ----------------------------------------------------------------------------- lattices
// Kernel axes are always 3-D (x slowest, z fastest).  D2Q9 fields [q][NX][NY] run as
// nx=1, ny=NX, nz=NY, i.e. user axis a maps to kernel axis a+1.
template <int Q>
struct Lat;

XLB_HD constexpr int digit3(int v) { return v == 0 ? 0 : (v == 1 ? -1 : 1); }

template <>
struct Lat<27> {
    static constexpr int Q = 27, D = 3;
    XLB_HD static constexpr int c(int l, int a) { return a == 0 ? digit3(l / 9) : (a == 1 ? digit3((l / 3) % 3) : digit3(l % 3)); }
    XLB_HD static constexpr double w(int l) {
        int n = (c(l, 0) != 0) + (c(l, 1) != 0) + (c(l, 2) != 0);
        return n == 0 ? 8.0 / 27.0 : (n == 1 ? 2.0 / 27.0 : (n == 2 ? 1.0 / 54.0 : 1.0 / 216.0));
    }
};

template <>
struct Lat<19> {
    static constexpr int Q = 19, D = 3;
    // index of the l-th member of the D3Q27 enumeration with |c|_1 <= 2
    XLB_HD static constexpr int l27(int l) {
        int k = 0;
        for (int m = 0; m < 27; ++m) {
            int n = (Lat<27>::c(m, 0) != 0) + (Lat<27>::c(m, 1) != 0) + (Lat<27>::c(m, 2) != 0);
            if (n <= 2) {
                if (k == l) return m;
                ++k;
            }
        }
        return 0;
    }
    XLB_HD static constexpr int c(int l, int a) { return Lat<27>::c(l27(l), a); }
    XLB_HD static constexpr double w(int l) {
        int n = (c(l, 0) != 0) + (c(l, 1) != 0) + (c(l, 2) != 0);
        return n == 0 ? 1.0 / 3.0 : (n == 1 ? 1.0 / 18.0 : 1.0 / 36.0);
    }
};

template <>
struct Lat<9> {
    static constexpr int Q = 9, D = 2;
    XLB_HD static constexpr int cx2(int l) {
        constexpr int t[9] = {0, 0, 0, 1, -1, 1, -1, 1, -1};
        return t[l];
    }
    XLB_HD static constexpr int cy2(int l) {
        constexpr int t[9] = {0, 1, -1, 0, 1, -1, 0, 1, -1};
        return t[l];
    }
    XLB_HD static constexpr int c(int l, int a) { return a == 0 ? 0 : (a == 1 ? cx2(l) : cy2(l)); }
    XLB_HD static constexpr double w(int l) {
        int n = (cx2(l) != 0) + (cy2(l) != 0);
        return n == 0 ? 4.0 / 9.0 : (n == 1 ? 1.0 / 9.0 : 1.0 / 36.0);
    }
};

template <int Q>
XLB_HD constexpr int opp(int l) {  // velocity_set.py:182-196
    for (int m = 0; m < Q; ++m)
        if (Lat<Q>::c(m, 0) == -Lat<Q>::c(l, 0) && Lat<Q>::c(m, 1) == -Lat<Q>::c(l, 1) && Lat<Q>::c(m, 2) == -Lat<Q>::c(l, 2)) return m;
    return 0;
}
template <int Q>
XLB_HD constexpr bool is_main(int l) {  // velocity_set.py:198-212
    return (Lat<Q>::c(l, 0) != 0) + (Lat<Q>::c(l, 1) != 0) + (Lat<Q>::c(l, 2) != 0) == 1;
}
// first kernel axis that carries a lattice dimension (1 for D2Q9, 0 otherwise)
template <int Q>
XLB_HD constexpr int axis0() { return 3 - Lat<Q>::D; }

// Compile-time loop: f(std::integral_constant<int, I>) for I in [0, N).  Every lattice-table lookup
// inside the body is then a true constant expression (folded by the front end, not by unrolling).
template <typename F, int... Is>
XLB_D void static_for_impl(F&& f, std::integer_sequence<int, Is...>) {
    (f(std::integral_constant<int, Is>{}), ...);
}
template <int N, typename F>
XLB_D void static_for(F&& f) {
    static_for_impl(static_cast<F&&>(f), std::make_integer_sequence<int, N>{});
}
#define XLB_IDX(name, tag) constexpr int name = decltype(tag)::value

// ----------------------------------------------------------------------------- element access
template <typename S>
struct DtypeOf;
template <>
struct DtypeOf<__half> { static constexpr int value = XLB_F16; };
template <>
struct DtypeOf<float> { static constexpr int value = XLB_F32; };
template <>
struct DtypeOf<double> { static constexpr int value = XLB_F64; };

template <typename T, typename S>
XLB_D T to_compute(S v) { return static_cast<T>(v); }
template <>
XLB_D float to_compute<float, __half>(__half v) { return __half2float(v); }
template <>
XLB_D double to_compute<double, __half>(__half v) { return static_cast<double>(__half2float(v)); }

template <typename S, typename T>
XLB_D S to_store(T v) { return static_cast<S>(v); }  // cvt.rn (round-to-nearest-even), precision_policy.py:105-110
template <>
XLB_D __half to_store<__half, float>(float v) { return __float2half_rn(v); }
template <>
XLB_D __half to_store<__half, double>(double v) { return __double2half(v); }

// V consecutive elements, naturally aligned so that one LDG/STG moves all of them (<= 16 B)
template <typename S, int V>
struct alignas(sizeof(S) * V) Vec {
    S v[V];
};

template <typename S, int V>
XLB_D Vec<S, V> ld_vec(const S* p) { return *reinterpret_cast<const Vec<S, V>*>(p); }
template <typename S, int V>
XLB_D void st_vec(S* p, const Vec<S, V>& x) { *reinterpret_cast<Vec<S, V>*>(p) = x; }

// ---- population loads / stores with an L2 eviction policy (XLB_TUNE_L2HINT: bit 0 stores evict_last, bit 1 loads evict_first).
// What a step writes is what the next step reads; what it reads is dead.  On lattices of the order of the 126 MB L2 the written
// lines should outlive the read ones.  16-byte accesses only (the widths the fused kernels use); other widths take the plain path.
#ifndef XLB_TUNE_L2HINT
#define XLB_TUNE_L2HINT 0
#endif
XLB_D uint64_t l2_policy_evict_last() {
    uint64_t pol;
    asm("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
XLB_D uint64_t l2_policy_evict_first() {
    uint64_t pol;
    asm("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
template <typename S, int V>
XLB_D Vec<S, V> ld_pop(const S* p, uint64_t pol) {
    if constexpr ((XLB_TUNE_L2HINT & 2) && sizeof(S) * V == 16) {
        uint32_t a, b, c, d;
        asm("ld.global.L2::cache_hint.v4.u32 {%0, %1, %2, %3}, [%4], %5;" : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "l"(p), "l"(pol));
        Vec<S, V> r;
        uint32_t w[4] = {a, b, c, d};
        memcpy(&r, w, 16);
        return r;
    } else {
        return ld_vec<S, V>(p);
    }
}
template <typename S, int V>
XLB_D void st_pop(S* p, const Vec<S, V>& x, uint64_t pol) {
    if constexpr ((XLB_TUNE_L2HINT & 1) && sizeof(S) * V == 16) {
        uint32_t w[4];
        memcpy(w, &x, 16);
        asm volatile("st.global.L2::cache_hint.v4.u32 [%0], {%1, %2, %3, %4}, %5;" ::"l"(p), "r"(w[0]), "r"(w[1]), "r"(w[2]), "r"(w[3]), "l"(pol) : "memory");
    } else {
        st_vec<S, V>(p, x);
    }
}

XLB_D float shfl_up1(float v) { return __shfl_up_sync(0xffffffffu, v, 1); }
XLB_D double shfl_up1(double v) { return __shfl_up_sync(0xffffffffu, v, 1); }
XLB_D __half shfl_up1(__half v) { return __shfl_up_sync(0xffffffffu, v, 1); }
XLB_D float shfl_down1(float v) { return __shfl_down_sync(0xffffffffu, v, 1); }
XLB_D double shfl_down1(double v) { return __shfl_down_sync(0xffffffffu, v, 1); }
XLB_D __half shfl_down1(__half v) { return __shfl_down_sync(0xffffffffu, v, 1); }

// runtime-typed scalar access for the stand-alone operators and BC side data
template <typename T>
XLB_D T ld_any(const void* p, int dtype, int64_t i) {
    if (dtype == XLB_F32) return static_cast<T>(static_cast<const float*>(p)[i]);
    if (dtype == XLB_F64) return static_cast<T>(static_cast<const double*>(p)[i]);
    return static_cast<T>(__half2float(static_cast<const __half*>(p)[i]));
}
template <typename T>
XLB_D void st_any(void* p, int dtype, int64_t i, T v) {
    if (dtype == XLB_F32)
        static_cast<float*>(p)[i] = static_cast<float>(v);
    else if (dtype == XLB_F64)
        static_cast<double*>(p)[i] = static_cast<double>(v);
    else
        static_cast<__half*>(p)[i] = to_store<__half, T>(v);
}

// ----------------------------------------------------------------------------- kernel parameter block
struct StepParams {
    const void* f0;
    void* f1;
    const uint8_t* bc_mask;
    const uint8_t* missing;
    const XlbBcEntry* bct;
    int nx, ny, nz;
    int x_begin, x_end;      // planes updated: x_begin + k * x_stride, k < n_planes (x_end = one past the last of them)
    int x_stride, n_planes;
    int y_begin, y_stride, n_rows;  // rows updated in each plane: y_begin + k * y_stride, k < n_rows (whole planes: 0, 1, ny)
    int x_periodic;
    int reverse;  // 1: walk the planes from the last to the first (odd time steps; see xlb_step_range)
    int ntz;  // nz / V
    int ntz_shift;  // log2(ntz) when ntz is a power of two, else -1
    long long qs;   // population stride (elements)
    long long mqs;  // missing-mask direction stride
    double omega;
    float halfway_fraction;  // hint for the variant choice (see XlbCollisionParams)
    int bc_kind_mask;  // bit k set: a BC of kind k may be present (0 = unknown, assume any)
    double les_c;      // 36 * smagorinsky_coef^2 (smagorinsky_les_bgk.py:62)
    double force3[3];  // 3 * force_vector, kernel axes (exact_difference_force.py:72-75)
    const long long* cells;  // XlbCollisionParams.bc_cells: cells of listed BC kinds, updated by step_cells_kernel
    long long n_cells;
    // halo push (XlbHaloPush): while updating slab-axis coordinate peer_src[k], the populations that leave the slab through
    // that side (c = -1 for k = 0, +1 for k = 1) are ALSO stored into the neighbour's buffer peer[k] (peer-mapped memory,
    // same q stride) at element offset peer_off[k] + (offset of the cell inside its plane / row)
    void* peer[2];
    long long peer_off[2];
    long long peer_qs[2];  // population stride of that neighbour's buffer (its slab may be thicker or thinner than mine)
    int peer_src[2];
    // tile scheduler of the persistent bulk kernel: two zero-initialised words in the library's static scratch
    // ([0] next dynamic tile, [1] blocks finished), private to this launch; nullptr = static stride
    unsigned* sched;
    // XlbCollisionParams.missing_bits: missing_mask packed to one word per cell (bit l = missing[l]); nullptr = read the q bytes
    const uint32_t* mbits;
};

// slab axis of the multi-GPU decomposition in kernel axes: x for the 3-D lattices, y (= user x) for D2Q9, whose fields
// run as nx = 1
#ifndef XLB_TUNE_REVERSE
#define XLB_TUNE_REVERSE 1
#endif
// k-th plane of a launch in the order the grid walks them (StepParams.reverse)
XLB_D int plane_index(const StepParams& p, int k) {
#if XLB_TUNE_REVERSE
    return p.reverse ? p.n_planes - 1 - k : k;
#else
    return k;
#endif
}

template <int Q>
XLB_HD constexpr int slab_axis() { return Q == 9 ? 1 : 0; }

// where a thread's outgoing populations go on the neighbour (nullptr / c = 0: nowhere)
template <typename S>
struct PeerDst {
    S* base;       // neighbour's population 0 at this thread's cell(s); population l is at base + l * qs
    long long qs;  // the NEIGHBOUR's population stride
    int c;         // lattice component along the slab axis that crosses this side: -1 or +1
};
template <typename S, int Q>
XLB_D PeerDst<S> peer_dst(const StepParams& p, int x, int y, long long offset_in_plane) {
    PeerDst<S> d{nullptr, 0, 0};
    if (p.peer[0] == nullptr && p.peer[1] == nullptr) return d;
    const int coord = slab_axis<Q>() == 0 ? x : y;
    const long long inner = slab_axis<Q>() == 0 ? offset_in_plane : offset_in_plane - (long long)y * p.nz;
    if (coord == p.peer_src[0] && p.peer[0] != nullptr) {
        d.base = static_cast<S*>(p.peer[0]) + (p.peer_off[0] + inner);
        d.qs = p.peer_qs[0];
        d.c = -1;
    } else if (coord == p.peer_src[1] && p.peer[1] != nullptr) {
        d.base = static_cast<S*>(p.peer[1]) + (p.peer_off[1] + inner);
        d.qs = p.peer_qs[1];
        d.c = 1;
    }
    return d;
}
// compile-time component of direction (cx, cy) along the slab axis
template <int Q>
XLB_HD constexpr int slab_c(int cx, int cy) { return slab_axis<Q>() == 0 ? cx : cy; }

// collision template codes: low bits = model (XLB_BGK / XLB_KBC / XLB_SMAGORINSKY_LES_BGK), bit 2 = exact-difference force
#define XLB_COLL_FORCED 4
#define XLB_COLL_MODEL(COLL) ((COLL) & 3)

XLB_HD bool bc_is_complex(int kind) { return kind >= XLB_BC_HALFWAY; }
// kinds that go on the BC cell list (XlbCollisionParams.bc_cells): every complex kind except the half-way bounce-back,
// whose cells fill obstacle volumes and have their own vectorised in-register path
XLB_HD bool bc_is_listed(int kind) { return kind >= XLB_BC_DONOTHING; }
#define XLB_BC_LISTED_KINDS                                                                                                      \
    ((1 << XLB_BC_DONOTHING) | (1 << XLB_BC_ZOUHE_VELOCITY) | (1 << XLB_BC_ZOUHE_PRESSURE) | (1 << XLB_BC_EXTRAPOLATION) | \
     (1 << XLB_BC_REGULARIZED_VELOCITY) | (1 << XLB_BC_REGULARIZED_PRESSURE))

}  // namespace xlb
