// xlb_step_aa.cuh -- the fused step on ONE population buffer (in-place "AA" pattern, Bailey et al. 2009): half the memory of
// the two-buffer scheme at the same bytes per cell update.  north_star: "optionally the in-place AA pattern to halve memory".
//
// Let S_t be the post-collision populations after step t.  The buffer A alternates between two representations:
//   R0 (normal, after an even number of steps):  A[l][x]            = S_t[l][x]
//   R1 (after an odd number of steps)         :  A[opp m][x + c_m]  = S_t[m][x]
// P step (R0 -> R1): cell x pulls f_l = A[l][x - c_l] like the two-buffer kernel, applies the boundary conditions, collides,
//   and writes result m to A[opp m][x + c_m] -- which is location (slot o = opp m, cell x - c_o), one of the locations this very
//   cell has just read, and no other cell reads or writes it: the update is in place without any hazard.
// L step (R1 -> R0): the pull of population l, S[l][x - c_l], sits at A[opp l][x]: every access is local to the cell and
//   aligned; the results go to A[m][x], the normal layout again.
// Per cell the arithmetic is the collide_cols of every other kernel, so after an even number of steps A equals the f buffer of
// the two-buffer scheme bit for bit (tests/test_gpu_aa.py).  Supported boundary conditions: none (periodic), FullwayBounceBack
// and EquilibriumBC -- the ones that act on the pulled populations in registers (the cavity of BASELINE configs[0..1], [3]);
// BCs that read f_pre or a neighbour (half-way, Zou-He, extrapolation) need the second buffer and are rejected on the host.
#pragma once
#include "xlb_step.cuh"

namespace xlb {

// predicated scalar store (no branch)
XLB_D void st_if(float* ptr, bool pred, float v) {
    asm volatile("{ .reg .pred q; setp.ne.s32 q, %2, 0; @q st.global.f32 [%0], %1; }" ::"l"(ptr), "f"(v), "r"((int)pred) : "memory");
}
XLB_D void st_if(double* ptr, bool pred, double v) {
    asm volatile("{ .reg .pred q; setp.ne.s32 q, %2, 0; @q st.global.f64 [%0], %1; }" ::"l"(ptr), "d"(v), "r"((int)pred) : "memory");
}
XLB_D void st_if(__half* ptr, bool pred, __half v) {
    asm volatile("{ .reg .pred q; setp.ne.s32 q, %2, 0; @q st.global.u16 [%0], %1; }" ::"l"(ptr), "h"(__half_as_ushort(v)), "r"((int)pred) : "memory");
}

// LOCAL = false: P step, LOCAL = true: L step
template <typename T, typename S, int Q, int COLL, int V, int BLOCK, int MINB, bool HAS_BC, bool LOCAL>
__global__ void __launch_bounds__(BLOCK, MINB) step_aa_kernel(const __grid_constant__ StepParams p) {
    S* __restrict__ A = static_cast<S*>(p.f1);
    const int nvec = p.n_rows * p.ntz;
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

    // BC state kept packed (two bit masks) so that it costs two registers across the collision, not 2 * V.  The ids are LOADED
    // here and first LOOKED AT after every population load has been issued: a branch on them up here would put one full DRAM
    // round trip in front of the loads (ncu source page of the FP16 kernel).
    Vec<uint8_t, V> idv;
#pragma unroll
    for (int j = 0; j < V; ++j) idv.v[j] = 0;
    if (HAS_BC) idv = ld_vec<uint8_t, V>(p.bc_mask + cell);

    // row offsets of the P step's neighbour accesses, indexed by c (0 -> same, +1 -> minus side, -1 -> plus side).  They are
    // needed before AND after the collision; holding them across it costs ~14 registers the collision has to spill for, so the
    // store phase recomputes them from (x, y, tz) behind an opaque barrier instead (a dozen integer instructions).
    int xm, xp, ym, yp, oy0, oym, oyp, z_lo, z_hi;
    long long ox0, oxm, oxp;
    bool edge_lo, edge_hi;
    auto offsets = [&](int xx, int yy, int tzz) {
        xm = p.x_periodic ? wrap1(xx - 1, p.nx) : xx - 1;
        xp = p.x_periodic ? wrap1(xx + 1, p.nx) : xx + 1;
        ym = wrap1(yy - 1, p.ny), yp = wrap1(yy + 1, p.ny);
        ox0 = xx * plane, oxm = xm * plane, oxp = xp * plane;
        oy0 = yy * p.nz, oym = ym * p.nz, oyp = yp * p.nz;
        edge_lo = (lane == 0) || (tzz == 0);
        edge_hi = (lane == 31) || (tzz == p.ntz - 1);
        z_lo = (tzz == 0) ? p.nz - 1 : tzz * V - 1;     // the element left of my vector (periodic)
        z_hi = (tzz == p.ntz - 1) ? 0 : tzz * V + V;    // the element right of my vector (periodic)
    };
    if constexpr (!LOCAL) offsets(x, y, tz);

    T f[Q][V];
    // both steps walk the SLOTS of the buffer in memory order with one running pointer (slot s holds population s in the P
    // step and population opp(s) in the L step), which keeps a single 64-bit address live instead of one per population
    const S* ps = A;
    static_for<Q>([&](auto sc) {
        XLB_IDX(sl, sc);
        if constexpr (LOCAL) {
            constexpr int l = Dir<Q, sl>::o;  // S[l][x - c_l] sits in slot opp l of this cell
            const Vec<S, V> v = ld_vec<S, V>(ps + cell);
#pragma unroll
            for (int j = 0; j < V; ++j) f[l][j] = to_compute<T, S>(v.v[j]);
        } else {
            constexpr int l = sl;
            using D = Dir<Q, l>;
            const long long ox = D::cx == 0 ? ox0 : (D::cx == 1 ? oxm : oxp);
            const int oy = D::cy == 0 ? oy0 : (D::cy == 1 ? oym : oyp);
            const S* row = ps + (ox + oy);
            const Vec<S, V> v = ld_vec<S, V>(row + z0);
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
        }
        ps += p.qs;
    });
    // inactive lanes only served the load shuffles; they stay alive for the store shuffles of the P step as well

    uint32_t full_mask = 0, eq_mask = 0;
    if (HAS_BC) {
        bool any = false;
#pragma unroll
        for (int j = 0; j < V; ++j) any |= idv.v[j] != 0;
        if (any) {
#pragma unroll
            for (int j = 0; j < V; ++j) {
                if (idv.v[j]) {
                    const int k = p.bct[idv.v[j]].kind;
                    full_mask |= (uint32_t)(k == XLB_BC_FULLWAY) << j;
                    eq_mask |= (uint32_t)(k == XLB_BC_EQUILIBRIUM) << j;
                }
            }
        }
    }
    if (HAS_BC && eq_mask) {
        // EquilibriumBC (STREAMING step): feq(rho_bc, u_bc) constants
#pragma unroll
        for (int j = 0; j < V; ++j) {
            if ((eq_mask >> j) & 1) {
                const double* tab = p.bct[idv.v[j]].table;
#pragma unroll
                for (int l = 0; l < Q; ++l) f[l][j] = static_cast<T>(tab[l]);
            }
        }
    }
    const CollArgs<T> ca = coll_args<T>(p);
    if constexpr (XLB_PACKED(COLL) && std::is_same<T, float>::value && (V % 2 == 0)) {
        static_for<V / 2>([&](auto jc) {
            XLB_IDX(jp, jc);
            const bool kf[2] = {(bool)((full_mask >> (2 * jp)) & 1), (bool)((full_mask >> (2 * jp + 1)) & 1)};
            collide_cols<F2, T, Q, COLL, V, 2 * jp>(f, ca, kf, HAS_BC && ((full_mask >> (2 * jp)) & 3) != 0);
        });
    } else {
        static_for<V>([&](auto jc) {
            XLB_IDX(j, jc);
            const bool kf[1] = {(bool)((full_mask >> j) & 1)};
            collide_cols<T, T, Q, COLL, V, j>(f, ca, kf, HAS_BC && kf[0]);
        });
    }

    if constexpr (!LOCAL) {
        int xx = x, yy = y, tzz = tz;
#ifndef XLB_TUNE_AA_KEEP
        asm volatile("" : "+r"(xx), "+r"(yy), "+r"(tzz));
#endif
        offsets(xx, yy, tzz);
    }
    S* qs_ptr = A;
    static_for<Q>([&](auto sc) {
        XLB_IDX(sl, sc);
        // slot sl receives: L step -- result sl of this cell; P step -- result m = opp(sl) of the cells z - c_m along this
        // thread's row shifted by c_m in x / y (the location the pull of population sl read)
        constexpr int m = LOCAL ? sl : Dir<Q, sl>::o;
        S out[V];
#pragma unroll
        for (int j = 0; j < V; ++j) out[j] = to_store<S, T>(f[m][j]);
        if constexpr (LOCAL) {
            if (active) {
                Vec<S, V> o;
#pragma unroll
                for (int j = 0; j < V; ++j) o.v[j] = out[j];
                st_vec<S, V>(qs_ptr + cell, o);
            }
        } else {
            using DO = Dir<Q, sl>;
            const long long ox = DO::cx == 0 ? ox0 : (DO::cx == 1 ? oxm : oxp);
            const int oy = DO::cy == 0 ? oy0 : (DO::cy == 1 ? oym : oyp);
            S* row = qs_ptr + (ox + oy);
            if (DO::cz == 0) {
                if (active) {
                    Vec<S, V> v;
#pragma unroll
                    for (int j = 0; j < V; ++j) v.v[j] = out[j];
                    st_vec<S, V>(row + z0, v);
                }
            } else if (DO::cz == 1) {
                // targets z - 1: element j of my vector takes my cell j + 1; element V - 1 takes the first cell of the lane to the
                // right; my first cell's result belongs to the vector on the left
                const S from_right = shfl_down1(out[0]);
                if (active) {
                    if (!edge_hi) {
                        Vec<S, V> v;
#pragma unroll
                        for (int j = 0; j < V - 1; ++j) v.v[j] = out[j + 1];
                        v.v[V - 1] = from_right;
                        st_vec<S, V>(row + z0, v);
                    } else {  // the lane to the right is in another warp (or the row ends): it stores that element itself
#pragma unroll
                        for (int j = 0; j < V - 1; ++j) row[z0 + j] = out[j + 1];
                    }
                    st_if(row + z_lo, edge_lo, out[0]);
                }
            } else {
                // targets z + 1: element j takes my cell j - 1; element 0 takes the last cell of the lane to the left
                const S from_left = shfl_up1(out[V - 1]);
                if (active) {
                    if (!edge_lo) {
                        Vec<S, V> v;
                        v.v[0] = from_left;
#pragma unroll
                        for (int j = 1; j < V; ++j) v.v[j] = out[j - 1];
                        st_vec<S, V>(row + z0, v);
                    } else {
#pragma unroll
                        for (int j = 1; j < V; ++j) row[z0 + j] = out[j - 1];
                    }
                    st_if(row + z_hi, edge_hi, out[V - 1]);
                }
            }
        }
        qs_ptr += p.qs;
    });
}

}  // namespace xlb
