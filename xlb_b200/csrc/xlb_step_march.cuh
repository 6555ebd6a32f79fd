// xlb_step_march.cuh -- FP16-storage step with register double buffering (experimental variant, XLB_TUNE_MARCH).
//
// step_vector_h2_kernel (xlb_step_h2.cuh) issues all loads of a thread's cells, waits, collides, stores and exits: at the
// 10-12 warps per SM its register budget allows, a warp's load phase and compute phase do not overlap (ncu: DRAM 64 %,
// issue slots 37 %, long-scoreboard stalls).  Here every thread keeps its (y, z) position and MARCHES along x for XC
// planes: the packed loads of plane x+1 are issued into a second register set before plane x is collided, so each warp
// always has a plane in flight while it computes.  Same addressing, shuffles, BC handling and arithmetic per cell as the
// h2 kernel (results are bit-identical); V = 4 cells per thread keeps two packed tiles (2 x 38 registers for D3Q19) plus
// the collision temporaries in ~180 registers.
// MEASURED SLOWER on B200 (47-49 k MLUPS against 68 k for step_vector_h2_kernel, profiles/sweep13_f16_march.log): kept as
// an off-by-default tuning knob and as the record of the experiment.
#pragma once
#include "xlb_step_h2.cuh"

namespace xlb {

// index of population L among the populations with c_z != 0 (the ones that need an edge element), -1 otherwise
template <int Q>
XLB_HD constexpr int shifted_index(int L) {
    if (Lat<Q>::c(L, 2) == 0) return -1;
    int n = 0;
    for (int l = 0; l < L; ++l) n += Lat<Q>::c(l, 2) != 0;
    return n;
}
template <int Q>
XLB_HD constexpr int n_shifted_pops() {
    int n = 0;
    for (int l = 0; l < Q; ++l) n += Lat<Q>::c(l, 2) != 0;
    return n;
}

template <int Q, int V>
struct MarchTile {
    PackedH2<Q, V / 2> pk;                  // loaded words, unshifted (become the packed, shifted populations in finish)
    uint32_t e[(n_shifted_pops<Q>() + 1) / 2];  // edge elements (16 bit each, two per register) for lanes 0 / 31 and row ends
    uint32_t idw[(V + 3) / 4];              // bc ids of the V cells
};

// BC classification, slow path, collision pair by pair, store -- the tail of step_vector_h2_kernel on a ready packed tile
template <int Q, int COLL, int V, bool HAS_BC, bool HALF>
XLB_D void h2_collide_store(const StepParams& p, PackedH2<Q, V / 2>& pk, const uint32_t (&idw)[(V + 3) / 4], int x, int y, int z0, long long cell) {
    using T = float;
    using S = __half;
    constexpr int W = V / 2;
    const S* __restrict__ f0 = static_cast<const S*>(p.f0);
    S* __restrict__ f1 = static_cast<S*>(p.f1);
    uint32_t m_full = 0, m_eq = 0, m_half = 0;
    bool any_bc = false, slow = false;
    auto id_of = [&](int j) -> int { return (idw[j >> 2] >> (8 * (j & 3))) & 0xff; };
    if (HAS_BC) {
#pragma unroll
        for (int k = 0; k < (V + 3) / 4; ++k) any_bc |= idw[k] != 0;
        if (any_bc) {
#pragma unroll
            for (int j = 0; j < V; ++j) {
                const int id = id_of(j);
                if (id) {
                    const int k = p.bct[id].kind;
                    m_full |= (k == XLB_BC_FULLWAY ? 1u : 0u) << j;
                    m_eq |= (k == XLB_BC_EQUILIBRIUM ? 1u : 0u) << j;
                    const bool here = p.cells == nullptr || !bc_is_listed(k);
                    if constexpr (HALF) {
                        m_half |= (k == XLB_BC_HALFWAY ? 1u : 0u) << j;
                        slow |= bc_is_complex(k) && k != XLB_BC_HALFWAY && here;
                    } else {
                        slow |= bc_is_complex(k) && here;
                    }
                }
            }
        }
    }
    if (HAS_BC && slow) {
        for (int j = 0; j < V; ++j) cell_update_generic<T, S, Q, COLL>(p, x, y, z0 + j);
        return;
    }
    const CollArgs<T> ca = coll_args<T>(p);
    constexpr bool FORCED = (COLL & XLB_COLL_FORCED) != 0;
    // always_inline: left to the inliner, this lambda is outlined and the packed tile is forced into local memory
    static_for<W>([&](auto jc) __attribute__((always_inline)) {
        XLB_IDX(jp, jc);
        constexpr uint32_t PAIR = 3u << (2 * jp);
        const bool pair_bc = HAS_BC && (((m_eq | m_half) & PAIR) != 0);
        const bool pair_full = HAS_BC && ((m_full & PAIR) != 0);
        bool fw[2] = {false, false};
        if (pair_full) {
            fw[0] = (m_full >> (2 * jp)) & 1u;
            fw[1] = (m_full >> (2 * jp + 1)) & 1u;
        }
        if (FORCED || !XLB_PACKED(COLL) || pair_bc) {
            T g[Q][2];
#pragma unroll
            for (int l = 0; l < Q; ++l) {
                const float2 t = h2_to_f2(pk.w[l][jp]);
                g[l][0] = t.x;
                g[l][1] = t.y;
            }
            if constexpr (HAS_BC && HALF) {
                if (m_half & PAIR) {  // HalfwayBounceBackBC, STREAMING step (bc_halfway_bounce_back.py:116-134)
                    uint32_t mb[2] = {0, 0};
                    static_for<Q>([&](auto lc) {
                        XLB_IDX(l, lc);
                        const Vec<uint8_t, 2> m = ld_vec<uint8_t, 2>(p.missing + l * p.mqs + cell + 2 * jp);
                        mb[0] |= (m.v[0] ? 1u : 0u) << l;
                        mb[1] |= (m.v[1] ? 1u : 0u) << l;
                    });
                    if (!((m_half >> (2 * jp)) & 1u)) mb[0] = 0;
                    if (!((m_half >> (2 * jp + 1)) & 1u)) mb[1] = 0;
                    const double* tab0 = p.bct[id_of(2 * jp)].table;
                    const double* tab1 = p.bct[id_of(2 * jp + 1)].table;
                    static_for<Q>([&](auto lc) {
                        XLB_IDX(l, lc);
                        const float2 pre = h2_to_f2(*reinterpret_cast<const uint32_t*>(f0 + Dir<Q, l>::o * p.qs + cell + 2 * jp));
                        if (mb[0] & (1u << l)) g[l][0] = pre.x + static_cast<T>(tab0[l]);
                        if (mb[1] & (1u << l)) g[l][1] = pre.y + static_cast<T>(tab1[l]);
                    });
                }
            }
            if (HAS_BC && (m_eq & PAIR)) {
#pragma unroll
                for (int k = 0; k < 2; ++k) {
                    if ((m_eq >> (2 * jp + k)) & 1u) {  // EquilibriumBC (STREAMING step)
                        const double* tab = p.bct[id_of(2 * jp + k)].table;
#pragma unroll
                        for (int l = 0; l < Q; ++l) g[l][k] = static_cast<T>(tab[l]);
                    }
                }
            }
            if constexpr (XLB_PACKED(COLL)) {
                collide_cols<F2, T, Q, COLL, 2, 0>(g, ca, fw, pair_full);
            } else {
                collide_cols<T, T, Q, COLL, 2, 0>(g, ca, &fw[0], fw[0]);
                collide_cols<T, T, Q, COLL, 2, 1>(g, ca, &fw[1], fw[1]);
            }
#pragma unroll
            for (int l = 0; l < Q; ++l) pk.w[l][jp] = f2_to_h2(g[l][0], g[l][1]);
        } else {
            collide_cols<F2, T, Q, COLL, V, 2 * jp>(pk, ca, fw, pair_full);
        }
    });
    S* ql = f1 + cell;
#pragma unroll
    for (int l = 0; l < Q; ++l) {
        st_words<W>(ql, pk.w[l]);
        ql += p.qs;
    }
}

// x-independent state of a marching thread
struct MarchCtx {
    long long plane;
    int y, z0, oy0, oym, oyp, z_lo, z_hi;
    bool edge_lo, edge_hi, active;
};

// all loads of plane x into tile t; nothing here waits for a loaded value
template <int Q, int V, bool HAS_BC>
XLB_D void march_issue(const StepParams& p, const MarchCtx& c, int x, MarchTile<Q, V>& t) {
    using S = __half;
    constexpr int W = V / 2;
    const int xm = p.x_periodic ? wrap1(x - 1, p.nx) : x - 1;
    const int xp = p.x_periodic ? wrap1(x + 1, p.nx) : x + 1;
    const long long ox0 = x * c.plane, oxm = xm * c.plane, oxp = xp * c.plane;
#pragma unroll
    for (int k = 0; k < (n_shifted_pops<Q>() + 1) / 2; ++k) t.e[k] = 0;
#pragma unroll
    for (int k = 0; k < (V + 3) / 4; ++k) t.idw[k] = 0;
    if (HAS_BC) {
        const uint8_t* m = p.bc_mask + (ox0 + c.oy0 + c.z0);
        if constexpr (V == 2) {
            t.idw[0] = *reinterpret_cast<const unsigned short*>(m);
        } else if constexpr (V == 4) {
            t.idw[0] = *reinterpret_cast<const uint32_t*>(m);
        } else {
            const uint2 w = *reinterpret_cast<const uint2*>(m);
            t.idw[0] = w.x;
            t.idw[1] = w.y;
        }
    }
    const S* pl = static_cast<const S*>(p.f0);
    static_for<Q>([&](auto lc) {
        XLB_IDX(l, lc);
        using D = Dir<Q, l>;
        const long long ox = D::cx == 0 ? ox0 : (D::cx == 1 ? oxm : oxp);
        const int oy = D::cy == 0 ? c.oy0 : (D::cy == 1 ? c.oym : c.oyp);
        const S* row = pl + (ox + oy);
        const Words<W> v = ld_words<W>(row + c.z0);
#pragma unroll
        for (int k = 0; k < W; ++k) t.pk.w[l][k] = v.w[k];
        constexpr int si = shifted_index<Q>(l);
        if constexpr (si >= 0) {
            const uint32_t el = D::cz == 1 ? ld_if_u16(row + c.z_lo, c.edge_lo, 0u) : ld_if_u16(row + c.z_hi, c.edge_hi, 0u);
            t.e[si / 2] |= el << (16 * (si & 1));
        }
        pl += p.qs;
    });
}

// shuffles + byte permutes turn the loaded words into the pulled populations, then BCs / collision / store
template <int Q, int COLL, int V, bool HAS_BC, bool HALF>
XLB_D void march_finish(const StepParams& p, const MarchCtx& c, int x, MarchTile<Q, V>& t) {
    constexpr int W = V / 2;
    static_for<Q>([&](auto lc) {
        XLB_IDX(l, lc);
        using D = Dir<Q, l>;
        constexpr int si = shifted_index<Q>(l);
        if constexpr (si >= 0) {
            uint32_t (&w)[W] = t.pk.w[l];
            const uint32_t el = (t.e[si / 2] >> (16 * (si & 1))) & 0xffffu;
            if (D::cz == 1) {  // pull from z-1: cell j takes element j-1
                const uint32_t prev = __shfl_up_sync(0xffffffffu, w[W - 1], 1);
                uint32_t carry = c.edge_lo ? el : (prev >> 16);  // low half = element to the left of word k
#pragma unroll
                for (int k = 0; k < W; ++k) {
                    const uint32_t cur = w[k];
                    w[k] = __byte_perm(carry, cur, 0x5410);
                    carry = cur >> 16;
                }
            } else {  // pull from z+1: cell j takes element j+1
                const uint32_t next = __shfl_down_sync(0xffffffffu, w[0], 1);
                const uint32_t right = c.edge_hi ? el : (next & 0xffffu);
#pragma unroll
                for (int k = 0; k < W - 1; ++k) w[k] = __byte_perm(w[k], w[k + 1], 0x5432);
                w[W - 1] = __byte_perm(w[W - 1], right, 0x5432);
            }
        }
    });
    if (c.active) h2_collide_store<Q, COLL, V, HAS_BC, HALF>(p, t.pk, t.idw, x, c.y, c.z0, x * c.plane + (long long)c.oy0 + c.z0);
}

template <int Q, int COLL, int V, int BLOCK, int MINB, bool HAS_BC, bool HALF, int XC>
__global__ void __launch_bounds__(BLOCK, MINB) step_march_h2_kernel(const __grid_constant__ StepParams p) {
    static_assert(V % 2 == 0 && V >= 2 && V <= 8, "V = 2, 4 or 8 fp16 cells per thread");
    const int nvec = p.ny * p.ntz;
    int i = blockIdx.x * BLOCK + threadIdx.x;
    MarchCtx c;
    c.active = i < nvec;
    if (!c.active) i = nvec - 1;  // keep the lane alive for the shuffles
    int y, tz;
    if (p.ntz_shift >= 0) {
        y = i >> p.ntz_shift;
        tz = i & (p.ntz - 1);
    } else {
        y = i / p.ntz;
        tz = i - y * p.ntz;
    }
    const int lane = threadIdx.x & 31;
    c.y = y;
    c.z0 = tz * V;
    c.plane = (long long)p.ny * p.nz;
    c.oy0 = y * p.nz;
    c.oym = wrap1(y - 1, p.ny) * p.nz;
    c.oyp = wrap1(y + 1, p.ny) * p.nz;
    c.edge_lo = (lane == 0) || (tz == 0);
    c.edge_hi = (lane == 31) || (tz == p.ntz - 1);
    c.z_lo = (tz == 0) ? p.nz - 1 : c.z0 - 1;
    c.z_hi = (tz == p.ntz - 1) ? 0 : c.z0 + V;
    const int xa = p.x_begin + blockIdx.y * XC;
    const int xb = min(xa + XC, p.x_end);
    if (xa >= xb) return;  // block-uniform
    MarchTile<Q, V> A, B;
    march_issue<Q, V, HAS_BC>(p, c, xa, A);
    for (int x = xa; x < xb; x += 2) {
        const bool hasB = x + 1 < xb;
        if (hasB) march_issue<Q, V, HAS_BC>(p, c, x + 1, B);
        march_finish<Q, COLL, V, HAS_BC, HALF>(p, c, x, A);
        if (!hasB) break;
        if (x + 2 < xb) march_issue<Q, V, HAS_BC>(p, c, x + 2, A);
        march_finish<Q, COLL, V, HAS_BC, HALF>(p, c, x + 1, B);
    }
}

}  // namespace xlb
