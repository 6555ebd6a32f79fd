// xlb_step.cuh -- the fused per-timestep kernel (K1) and its generic per-cell engine.
//
// Replaces IncompressibleNavierStokesStepper.__call__ (xlb/operator/stepper/nse_stepper.py:
// 238-282 JAX pull path = semantics; :427-464 Warp kernel = GPU design reference).
//
// Layout: SoA populations f[l][x][y][z], z fastest.  One thread owns V consecutive z cells
// (V * sizeof(store type) <= 16 B), so every population is moved with one aligned vector
// LDG/STG per thread; the +-1 z shift of the pull is resolved with a warp shuffle of the
// single edge element (edge lanes of a row fall back to one scalar load).  Populations stay
// in registers from load to store; boundary handling is keyed on the packed bc_mask bytes:
//   * bc == 0 for all V cells        -> branch-free fast path
//   * FullwayBounceBack/EquilibriumBC -> in-register selects (no extra memory traffic)
//   * every other BC kind             -> cell_update_generic() below, per cell.
#pragma once
#include "xlb_math.cuh"

// 1: FP32 collision on the packed pipe (two cells per FFMA2 / FADD2 / FMUL2); 0: two independent scalar chains.
// Measured on B200 (profiles/*sweep*): packed is ahead for BGK and, once the KBC passes were trimmed, for KBC too.
#ifndef XLB_TUNE_PACKED
#define XLB_TUNE_PACKED 1
#endif
#define XLB_PACKED(COLL) (XLB_TUNE_PACKED)

namespace xlb {

XLB_D int wrap1(int v, int n) { return v < 0 ? v + n : (v >= n ? v - n : v); }
XLB_D int wrapn(int v, int n) {  // periodic wrap of an index that is at most a few cells outside [0, n)
    v = v < 0 ? v + n : (v >= n ? v - n : v);
    if ((unsigned)v >= (unsigned)n) {  // extents of 1 or 2 cells: general modulo (integer division is slow, keep it off the common path)
        v %= n;
        v = v < 0 ? v + n : v;
    }
    return v;
}

// ------------------------------------------------------------------ generic per-cell engine
// Handles every BC kind; all reads come from f0, the only write is the cell's own f1 column.
// (body; the fused kernels call it through the __noinline__ wrapper below so that their fast path keeps its registers,
// the dense kernels -- step_cells_kernel, step_generic_kernel -- inline it, which turns every p.* into a constant-bank
// operand instead of a generic load through the by-reference parameter block)
template <typename T, typename S, int Q, int COLL>
XLB_D void cell_update_generic_impl(const StepParams& p, int x, int y, int z) {
    const S* __restrict__ f0 = static_cast<const S*>(p.f0);
    S* __restrict__ f1 = static_cast<S*>(p.f1);
    const long long plane = (long long)p.ny * p.nz;
    const long long cell = x * plane + (long long)y * p.nz + z;
    // source offsets of the pull per axis, indexed by c (0 -> same, +1 -> minus side, -1 -> plus side): 6 wraps per cell
    // instead of 3 per direction
    const long long ox0 = x * plane, oxm = (p.x_periodic ? wrap1(x - 1, p.nx) : x - 1) * plane, oxp = (p.x_periodic ? wrap1(x + 1, p.nx) : x + 1) * plane;
    const int oy0 = y * p.nz, oym = wrap1(y - 1, p.ny) * p.nz, oyp = wrap1(y + 1, p.ny) * p.nz;
    const int oz0 = z, ozm = wrap1(z - 1, p.nz), ozp = wrap1(z + 1, p.nz);
    const int id = p.bc_mask ? p.bc_mask[cell] : 0;
    int kind = XLB_BC_NONE;
    const XlbBcEntry* e = nullptr;
    if (id != 0) {
        e = p.bct + id;
        kind = e->kind;
    }
    T f[Q][1];
    // pull-stream (stream.py:29-62)
    static_for<Q>([&](auto lc) {
        XLB_IDX(l, lc);
        using D = Dir<Q, l>;
        const long long ox = D::cx == 0 ? ox0 : (D::cx == 1 ? oxm : oxp);
        const int oy = D::cy == 0 ? oy0 : (D::cy == 1 ? oym : oyp);
        const int oz = D::cz == 0 ? oz0 : (D::cz == 1 ? ozm : ozp);
        f[l][0] = to_compute<T, S>(f0[l * p.qs + (ox + (oy + oz))]);
    });
    uint32_t miss = 0;
    if (kind >= XLB_BC_HALFWAY && kind != XLB_BC_DONOTHING) {
        if (p.mbits) {
            miss = p.mbits[cell];
        } else {
#pragma unroll
            for (int l = 0; l < Q; ++l) miss |= (p.missing[l * p.mqs + cell] ? 1u : 0u) << l;
        }
    }
    // ---- STREAMING-step boundary conditions (nse_stepper.py:250-257)
    if (kind == XLB_BC_EQUILIBRIUM) {
#pragma unroll
        for (int l = 0; l < Q; ++l) f[l][0] = static_cast<T>(e->table[l]);
    } else if (kind == XLB_BC_DONOTHING) {
#pragma unroll
        for (int l = 0; l < Q; ++l) f[l][0] = to_compute<T, S>(f0[l * p.qs + cell]);
    } else if (kind == XLB_BC_HALFWAY || kind == XLB_BC_EXTRAPOLATION) {
        // f_post[l] = f_pre[opp l] (+ 6 w_l c_l.u_wall) where missing[l]
        static_for<Q>([&](auto lc) {
            XLB_IDX(l, lc);
            if (miss & (1u << l)) {
                T v = to_compute<T, S>(f0[Dir<Q, l>::o * p.qs + cell]);
                if (kind == XLB_BC_HALFWAY) v = v + static_cast<T>(e->table[l]);
                f[l][0] = v;
            }
        });
    } else if (kind == XLB_BC_ZOUHE_VELOCITY || kind == XLB_BC_ZOUHE_PRESSURE || kind == XLB_BC_REGULARIZED_VELOCITY || kind == XLB_BC_REGULARIZED_PRESSURE) {
        zouhe_apply<T, Q>(f, miss, *e, kind == XLB_BC_ZOUHE_VELOCITY || kind == XLB_BC_REGULARIZED_VELOCITY, x, y, z, kind >= XLB_BC_REGULARIZED_VELOCITY);
    }
    // ---- macroscopic, equilibrium, collision (+ COLLISION-step FullwayBounceBackBC)
    T fps[Q];
    if (kind == XLB_BC_EXTRAPOLATION) {
#pragma unroll
        for (int l = 0; l < Q; ++l) fps[l] = f[l][0];
    }
    collide_column<T, Q, COLL, 1, 0>(f, coll_args<T>(p), kind == XLB_BC_FULLWAY);
    // ---- ExtrapolationOutflowBC.assemble_auxiliary_data (bc_extrapolation_outflow.py:106-133)
    if (kind == XLB_BC_EXTRAPOLATION) {
        const T cs = T(1.0) / sqrt(T(3.0));
        const int xn = x - e->normal[0], yn = y - e->normal[1], zn = z - e->normal[2];
        // per-axis source offsets of the pull at the interior neighbour x - n (same indexing by c as above)
        const long long nx0 = (p.x_periodic ? wrapn(xn, p.nx) : xn) * plane, nxm = (p.x_periodic ? wrapn(xn - 1, p.nx) : xn - 1) * plane,
                        nxp = (p.x_periodic ? wrapn(xn + 1, p.nx) : xn + 1) * plane;
        const int ny0 = wrapn(yn, p.ny) * p.nz, nym = wrapn(yn - 1, p.ny) * p.nz, nyp = wrapn(yn + 1, p.ny) * p.nz;
        const int nz0 = wrapn(zn, p.nz), nzm = wrapn(zn - 1, p.nz), nzp = wrapn(zn + 1, p.nz);
        static_for<Q>([&](auto lc) {
            XLB_IDX(l, lc);
            using D = Dir<Q, l>;
            if (miss & (1u << l)) {
                const long long ox = D::cx == 0 ? nx0 : (D::cx == 1 ? nxm : nxp);
                const int oy = D::cy == 0 ? ny0 : (D::cy == 1 ? nym : nyp);
                const int oz = D::cz == 0 ? nz0 : (D::cz == 1 ? nzm : nzp);
                const T nb = to_compute<T, S>(f0[l * p.qs + (ox + (oy + oz))]);
                f[D::o][0] = cs * nb + (T(1.0) - cs) * fps[l];
            }
        });
    }
    const PeerDst<S> pd = peer_dst<S, Q>(p, x, y, (long long)y * p.nz + z);
    static_for<Q>([&](auto lc) {
        XLB_IDX(l, lc);
        const S v = to_store<S, T>(f[l][0]);
        f1[l * p.qs + cell] = v;
        constexpr int ca = slab_c<Q>(Dir<Q, l>::cx, Dir<Q, l>::cy);
        if (ca != 0 && pd.c == ca) pd.base[l * pd.qs] = v;  // halo push: this population crosses into the neighbour's slab
    });
}

template <typename T, typename S, int Q, int COLL>
__device__ __noinline__ void cell_update_generic(const StepParams& p, int x, int y, int z) {
    cell_update_generic_impl<T, S, Q, COLL>(p, x, y, z);
}

// every cell through the generic engine (variant XLB_VARIANT_GENERIC; cross-check + odd shapes)
template <typename T, typename S, int Q, int COLL>
__global__ void __launch_bounds__(128) step_generic_kernel(const __grid_constant__ StepParams p) {
    const long long row = (long long)p.n_rows * p.nz;
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= row) return;
    const int x = p.x_begin + plane_index(p, blockIdx.y) * p.x_stride;
    cell_update_generic_impl<T, S, Q, COLL>(p, x, p.y_begin + (int)(i / p.nz) * p.y_stride, (int)(i % p.nz));
}

// the cells of the BC cell list (XlbCollisionParams.bc_cells) through the generic engine: consecutive threads take
// consecutive list entries, so a sorted list of a boundary plane is read and written coalesced, at full occupancy.
// Runs after the fused kernel on the same stream and overwrites what that kernel's fast path stored for these cells;
// the engine reads f0 only, so the order between the two launches does not change the result.
template <typename T, typename S, int Q, int COLL>
__global__ void __launch_bounds__(128) step_cells_kernel(const __grid_constant__ StepParams p) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= p.n_cells) return;
    const long long cell = p.cells[i];
    const long long plane = (long long)p.ny * p.nz;
    int x, r;
    if (cell < (1ll << 31) && plane < (1ll << 31)) {  // 32-bit division where the block allows it
        x = (int)((unsigned)cell / (unsigned)plane);
        r = (int)((unsigned)cell - (unsigned)x * (unsigned)plane);
    } else {
        x = (int)(cell / plane);
        r = (int)(cell - x * plane);
    }
    const int y = r / p.nz;
    // only cells of the planes / rows this launch updates (a strided range lists e.g. the two edge planes of a slab)
    const int dx = x - p.x_begin, dy = y - p.y_begin;
    if (dx < 0 || dx % p.x_stride != 0 || dx / p.x_stride >= p.n_planes) return;
    if (dy < 0 || dy % p.y_stride != 0 || dy / p.y_stride >= p.n_rows) return;
    cell_update_generic_impl<T, S, Q, COLL>(p, x, y, r - y * p.nz);
}

// ------------------------------------------------------------------ fused vector kernel (K1)
// predicated scalar load (no branch): returns *ptr where pred, else keep
XLB_D float ld_if(const float* ptr, bool pred, float keep) {
    asm("{ .reg .pred q; setp.ne.s32 q, %2, 0; @q ld.global.nc.f32 %0, [%1]; }" : "+f"(keep) : "l"(ptr), "r"((int)pred));
    return keep;
}
XLB_D double ld_if(const double* ptr, bool pred, double keep) {
    asm("{ .reg .pred q; setp.ne.s32 q, %2, 0; @q ld.global.nc.f64 %0, [%1]; }" : "+d"(keep) : "l"(ptr), "r"((int)pred));
    return keep;
}
XLB_D __half ld_if(const __half* ptr, bool pred, __half keep) {
    unsigned short k = __half_as_ushort(keep);
    asm("{ .reg .pred q; setp.ne.s32 q, %2, 0; @q ld.global.nc.u16 %0, [%1]; }" : "+h"(k) : "l"(ptr), "r"((int)pred));
    return __ushort_as_half(k);
}

// HALF: the BC list holds a HalfwayBounceBackBC; its cells (obstacle surface AND interior -- every solid cell is tagged)
// then take a vectorised in-register path instead of the per-cell generic engine.  It is a template flag because the
// fast path of this kernel is register-critical: compiling the extra path into the BC-free / cavity instance costs 5-8 %.
// PUSH: the halo-push instance (XlbHaloPush) -- used for the edge-plane launch of a slab only; the instance that updates the
// bulk of the lattice is compiled without it, because even three more live registers in the store loop cost it spills.
#ifndef XLB_TUNE_K1_PHASED
#define XLB_TUNE_K1_PHASED 0  // 1: ids -> all population loads -> classification -> shifts
#endif

template <typename T, typename S, int Q, int COLL, int V, int BLOCK, int MINB, bool HAS_BC, bool HALF = false, bool PUSH = false>
__global__ void __launch_bounds__(BLOCK, MINB) step_vector_kernel(const __grid_constant__ StepParams p) {
    const S* __restrict__ f0 = static_cast<const S*>(p.f0);
    S* __restrict__ f1 = static_cast<S*>(p.f1);
    const int nvec = p.n_rows * p.ntz;  // vectors per x plane
    int i = blockIdx.x * BLOCK + threadIdx.x;
    const bool active = i < nvec;
    if (!active) i = nvec - 1;  // keep the lane alive for the shuffles
    const int x = p.x_begin + plane_index(p, blockIdx.y) * p.x_stride;
    int y, tz;
    if (p.ntz_shift >= 0) {
        y = i >> p.ntz_shift;
        tz = i & (p.ntz - 1);
    } else {
        y = i / p.ntz;
        tz = i - y * p.ntz;
    }
    y = p.y_begin + y * p.y_stride;
    const int z0 = tz * V;
    const int lane = threadIdx.x & 31;
    const long long plane = (long long)p.ny * p.nz;
    const long long cell = x * plane + (long long)y * p.nz + z0;

    // ---- packed bc ids of my V cells
    uint8_t ids[V];
    bool any_bc = false, slow = false, any_half = false;
    bool k_full[V], k_eq[V], k_half[V];
#pragma unroll
    for (int j = 0; j < V; ++j) k_full[j] = k_eq[j] = k_half[j] = false;
    Vec<uint8_t, V> idv;
    if (HAS_BC) idv = ld_vec<uint8_t, V>(p.bc_mask + cell);
    auto classify = [&]() {
#pragma unroll
        for (int j = 0; j < V; ++j) {
            ids[j] = idv.v[j];
            any_bc |= ids[j] != 0;
        }
        if (any_bc) {
#pragma unroll
            for (int j = 0; j < V; ++j) {
                if (ids[j]) {
                    const int k = p.bct[ids[j]].kind;
                    k_full[j] = k == XLB_BC_FULLWAY;
                    k_eq[j] = k == XLB_BC_EQUILIBRIUM;
                    // listed kinds are left to step_cells_kernel when the caller passed the BC cell list
                    const bool here = p.cells == nullptr || !bc_is_listed(k);
                    if constexpr (HALF) {
                        k_half[j] = k == XLB_BC_HALFWAY;
                        any_half |= k_half[j];
                        slow |= bc_is_complex(k) && !k_half[j] && here;
                    } else {
                        slow |= bc_is_complex(k) && here;
                    }
                }
            }
        }
    };
#if XLB_TUNE_K1_PHASED == 0
    if (HAS_BC) classify();
#endif

    // ---- pull-stream: one aligned vector load per population
    const int xm = p.x_periodic ? wrap1(x - 1, p.nx) : x - 1;
    const int xp = p.x_periodic ? wrap1(x + 1, p.nx) : x + 1;
    const int ym = wrap1(y - 1, p.ny), yp = wrap1(y + 1, p.ny);
    // element offsets of the source rows, indexed by c (0 -> same, +1 -> pull from the minus side, -1 -> plus side)
    const long long ox0 = x * plane, oxm = xm * plane, oxp = xp * plane;
    const int oy0 = y * p.nz, oym = ym * p.nz, oyp = yp * p.nz;
    const bool edge_lo = (lane == 0) || (tz == 0);
    const bool edge_hi = (lane == 31) || (tz == p.ntz - 1);
    const int z_lo = (tz == 0) ? p.nz - 1 : z0 - 1;   // source of the element left of my vector
    const int z_hi = (tz == p.ntz - 1) ? 0 : z0 + V;  // source of the element right of my vector
    T f[Q][V];
    const S* pl = f0;
    uint64_t pol_ld = 0, pol_st = 0;
    if constexpr ((XLB_TUNE_L2HINT & 2) != 0) pol_ld = l2_policy_evict_first();
    if constexpr ((XLB_TUNE_L2HINT & 1) != 0) pol_st = l2_policy_evict_last();
#if XLB_TUNE_K1_PHASED == 1
    // phase 1: issue every population's vector load before anything waits on one.  With the classification of tagged cells in
    // front of the loads, and the exit of inactive lanes between the z-shifted populations and the rest, a thread made three
    // dependent round trips to DRAM (ids -> shifted populations -> the others; ncu source page)
    Vec<S, V> vv[Q];
    static_for<Q>([&](auto lc) {
        XLB_IDX(l, lc);
        using D = Dir<Q, l>;
        const long long ox = D::cx == 0 ? ox0 : (D::cx == 1 ? oxm : oxp);
        const int oy = D::cy == 0 ? oy0 : (D::cy == 1 ? oym : oyp);
        vv[l] = ld_vec<S, V>(pl + (ox + oy) + z0);
        pl += p.qs;
    });
    if (HAS_BC) classify();
    // phase 2: the z shift (neighbour lanes by shuffle, across the warp boundary by a predicated scalar load)
    pl = f0;
    static_for<Q>([&](auto lc) {
        XLB_IDX(l, lc);
        using D = Dir<Q, l>;
        const Vec<S, V>& v = vv[l];
        if (D::cz == 0) {
#pragma unroll
            for (int j = 0; j < V; ++j) f[l][j] = to_compute<T, S>(v.v[j]);
        } else {
            const long long ox = D::cx == 0 ? ox0 : (D::cx == 1 ? oxm : oxp);
            const int oy = D::cy == 0 ? oy0 : (D::cy == 1 ? oym : oyp);
            const S* row = pl + (ox + oy);
            if (D::cz == 1) {  // pull from z-1
                const S left = ld_if(row + z_lo, edge_lo, shfl_up1(v.v[V - 1]));
                f[l][0] = to_compute<T, S>(left);
#pragma unroll
                for (int j = 1; j < V; ++j) f[l][j] = to_compute<T, S>(v.v[j - 1]);
            } else {  // pull from z+1
                const S right = ld_if(row + z_hi, edge_hi, shfl_down1(v.v[0]));
#pragma unroll
                for (int j = 0; j < V - 1; ++j) f[l][j] = to_compute<T, S>(v.v[j + 1]);
                f[l][V - 1] = to_compute<T, S>(right);
            }
        }
        pl += p.qs;
    });
#else
    static_for<Q>([&](auto lc) {
        XLB_IDX(l, lc);
        using D = Dir<Q, l>;
        const long long ox = D::cx == 0 ? ox0 : (D::cx == 1 ? oxm : oxp);
        const int oy = D::cy == 0 ? oy0 : (D::cy == 1 ? oym : oyp);
        const S* row = pl + (ox + oy);
        const Vec<S, V> v = ld_pop<S, V>(row + z0, pol_ld);
        if (D::cz == 0) {
#pragma unroll
            for (int j = 0; j < V; ++j) f[l][j] = to_compute<T, S>(v.v[j]);
        } else if (D::cz == 1) {  // pull from z-1
            const S left = ld_if(row + z_lo, edge_lo, shfl_up1(v.v[V - 1]));
            f[l][0] = to_compute<T, S>(left);
#pragma unroll
            for (int j = 1; j < V; ++j) f[l][j] = to_compute<T, S>(v.v[j - 1]);
        } else {  // pull from z+1
            const S right = ld_if(row + z_hi, edge_hi, shfl_down1(v.v[0]));
#pragma unroll
            for (int j = 0; j < V - 1; ++j) f[l][j] = to_compute<T, S>(v.v[j + 1]);
            f[l][V - 1] = to_compute<T, S>(right);
        }
        pl += p.qs;
    });
#endif
#if XLB_TUNE_K1_PHASED == 2
    // mode 2: the original interleaved load / shift loop, but the ids are looked at only now, and inactive lanes (last block only)
    // leave just before the stores -- an exit here lets the compiler sink the loads that need no shuffle below it
    if (HAS_BC) classify();
#else
    if (!active) return;
#endif

    if (HAS_BC && slow) {
        // a BC kind that needs f_pre / missing_mask / neighbour gathers: per-cell generic engine
        if (active)
            for (int j = 0; j < V; ++j) cell_update_generic<T, S, Q, COLL>(p, x, y, z0 + j);
        return;
    }
    if constexpr (HAS_BC && HALF) {
        if (any_half) {
            // HalfwayBounceBackBC, STREAMING step (bc_halfway_bounce_back.py:116-134): where missing[l],
            // f_post[l] = f_pre[opp l] (+ 6 w_l c_l.u_wall) of the same cell.  Two phases of independent vector loads
            // (all mask bytes, then all opposite populations) so no load waits on another.
            uint32_t mb[V];
#pragma unroll
            for (int j = 0; j < V; ++j) mb[j] = 0;
            if (p.mbits) {  // one vector load of the packed words instead of q byte-plane loads
                const Vec<uint32_t, V> m = ld_vec<uint32_t, V>(p.mbits + cell);
#pragma unroll
                for (int j = 0; j < V; ++j) mb[j] = m.v[j];
            } else {
                static_for<Q>([&](auto lc) {
                    XLB_IDX(l, lc);
                    const Vec<uint8_t, V> m = ld_vec<uint8_t, V>(p.missing + l * p.mqs + cell);
#pragma unroll
                    for (int j = 0; j < V; ++j) mb[j] |= (m.v[j] ? 1u : 0u) << l;
                });
            }
#pragma unroll
            for (int j = 0; j < V; ++j)
                if (!k_half[j]) mb[j] = 0;
            static_for<Q>([&](auto lc) {
                XLB_IDX(l, lc);
                const Vec<S, V> g = ld_vec<S, V>(f0 + Dir<Q, l>::o * p.qs + cell);
#pragma unroll
                for (int j = 0; j < V; ++j)
                    if (mb[j] & (1u << l)) f[l][j] = to_compute<T, S>(g.v[j]) + static_cast<T>(p.bct[ids[j]].table[l]);
            });
        }
    }
    if (HAS_BC && any_bc) {
        // EquilibriumBC (STREAMING step): f_post_stream = feq(rho_bc, u_bc) constants
#pragma unroll
        for (int j = 0; j < V; ++j) {
            if (k_eq[j]) {
                const double* tab = p.bct[ids[j]].table;
#pragma unroll
                for (int l = 0; l < Q; ++l) f[l][j] = static_cast<T>(tab[l]);
            }
        }
    }
    const CollArgs<T> ca = coll_args<T>(p);
    // ---- collide my cells in place; FP32 runs two cells per instruction on the packed FP32 pipe (FFMA2 / FADD2 / FMUL2)
    if constexpr (XLB_PACKED(COLL) && std::is_same<T, float>::value && (V % 2 == 0)) {
        static_for<V / 2>([&](auto jc) {
            XLB_IDX(jp, jc);
            collide_cols<F2, T, Q, COLL, V, 2 * jp>(f, ca, &k_full[2 * jp], HAS_BC && (k_full[2 * jp] || k_full[2 * jp + 1]));
        });
    } else {
        static_for<V>([&](auto jc) {
            XLB_IDX(j, jc);
            collide_cols<T, T, Q, COLL, V, j>(f, ca, &k_full[j], HAS_BC && k_full[j]);
        });
    }
    // ---- one aligned vector store per population (+ the halo push of the populations that leave the slab here)
#if XLB_TUNE_K1_PHASED == 2
    if (!active) return;
#endif
    S* ql = f1 + cell;
    PeerDst<S> pd{nullptr, 0, 0};
    if constexpr (PUSH) pd = peer_dst<S, Q>(p, x, y, (long long)y * p.nz + z0);
    S* qp = pd.base;
    static_for<Q>([&](auto lc) {
        XLB_IDX(l, lc);
        Vec<S, V> o;
#pragma unroll
        for (int j = 0; j < V; ++j) o.v[j] = to_store<S, T>(f[l][j]);
        st_pop<S, V>(ql, o, pol_st);
        ql += p.qs;
        if constexpr (PUSH) {
            constexpr int ca = slab_c<Q>(Dir<Q, l>::cx, Dir<Q, l>::cy);
            if (ca != 0 && pd.c == ca) st_vec<S, V>(qp, o);
            qp += pd.qs;
        }
    });
}

}  // namespace xlb
