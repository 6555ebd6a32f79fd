// xlb_step_h2.cuh -- the fused step for FP16-stored populations (FP32FP16, precision_policy.py:57-97) with the
// populations held PACKED (two fp16 cells per 32-bit register) from the load to the store.
//
// Same addressing, boundary handling and per-cell arithmetic as step_vector_kernel (xlb_step.cuh); what differs is the
// register footprint.  step_vector_kernel widens all Q x V populations to FP32 right after the load (Q*V live
// registers: 76 for D3Q19, V = 4), which at the occupancy this bandwidth-bound kernel needs forces spills (ncu,
// profiles/r1q_ncu_details_f16.txt: 6 % of the executed instructions, 20 GB of local-memory traffic through L2 per
// 302 M-cell launch).  Here the loaded vectors stay as half2 words (Q*V/2 registers), the +-1 z shift of the pull is
// done on the packed words with byte permutes (PRMT), and cells are widened to FP32 only one PAIR at a time for the
// collision (two cells per FFMA2 / FADD2 / FMUL2) and narrowed again with cvt.rn.f16x2.f32 before the next pair.
// Results are bit-identical to the scalar / generic kernels (same FP32 operations, same round-to-nearest-even store).
// Default instance (B200 sweeps, profiles/sweep9..12_f16_*.log): V = 8 cells per thread (one LDG.128 per population),
// 64-thread blocks, pairs widened to registers; 68-69 k MLUPS on D3Q19 BGK = 0.80-0.82 of the measured HBM peak.
#pragma once
#include "xlb_step.cuh"

// 1: always widen a pair to FP32 registers before the collision (one convert per element instead of one per read);
// needs the register budget the V = 8 instance runs with (168-204 registers per thread), +2 % there
#ifndef XLB_TUNE_H2_WIDEN
#define XLB_TUNE_H2_WIDEN -1  // -1: widen where the register budget allows it (V = 8), profiles/sweep11_f16_*.log
#endif

// 1 (experiments only, results are WRONG): skip the collision -- same loads, z shifts and stores.  Separates "the access
// pattern cannot go faster" from "the collision's registers / latency hold the loads back" (VERDICT r1 item 5).
#ifndef XLB_TUNE_STREAM_ONLY
#define XLB_TUNE_STREAM_ONLY 0
#endif

// 1: the nine source-row addresses (c_x, c_y) of the pull are formed once per thread as byte pointers and every population's
// load adds only the warp-uniform l * q_stride to one of them (2 integer instructions per load instead of ~6)
#ifndef XLB_TUNE_H2_ROWPTR
#define XLB_TUNE_H2_ROWPTR 0
#endif

namespace xlb {

XLB_D uint32_t ld_if_u16(const void* ptr, bool pred, uint32_t keep) {
    unsigned short k = static_cast<unsigned short>(keep);
    asm("{ .reg .pred q; setp.ne.s32 q, %2, 0; @q ld.global.nc.u16 %0, [%1]; }" : "+h"(k) : "l"(ptr), "r"((int)pred));
    return k;
}

template <int W>
struct Words {
    uint32_t w[W];
};
template <int W>
XLB_D Words<W> ld_words(const void* p) {
    Words<W> r;
    if constexpr (W == 1) {
        r.w[0] = *reinterpret_cast<const uint32_t*>(p);
    } else if constexpr (W == 2) {
        const uint2 t = *reinterpret_cast<const uint2*>(p);
        r.w[0] = t.x;
        r.w[1] = t.y;
    } else {
        static_assert(W == 4, "1, 2 or 4 words");
        const uint4 t = *reinterpret_cast<const uint4*>(p);
        r.w[0] = t.x;
        r.w[1] = t.y;
        r.w[2] = t.z;
        r.w[3] = t.w;
    }
    return r;
}
template <int W>
XLB_D void st_words(void* p, const uint32_t (&w)[W]) {
    if constexpr (W == 1) {
        *reinterpret_cast<uint32_t*>(p) = w[0];
    } else if constexpr (W == 2) {
        *reinterpret_cast<uint2*>(p) = make_uint2(w[0], w[1]);
    } else {
        *reinterpret_cast<uint4*>(p) = make_uint4(w[0], w[1], w[2], w[3]);
    }
}

#ifndef XLB_TUNE_H2_PHASED
#define XLB_TUNE_H2_PHASED 1  // 1: ids -> all population loads -> classification -> shifts (one DRAM round trip instead of three)
#endif

template <int Q, int COLL, int V, int BLOCK, int MINB, bool HAS_BC, bool HALF, bool PUSH = false>
__global__ void __launch_bounds__(BLOCK, MINB) step_vector_h2_kernel(const __grid_constant__ StepParams p) {
    using T = float;
    using S = __half;
    static_assert(V % 2 == 0 && V >= 2 && V <= 8, "V = 2, 4 or 8 fp16 cells per thread");
    constexpr int W = V / 2;  // packed words per population
    const S* __restrict__ f0 = static_cast<const S*>(p.f0);
    S* __restrict__ f1 = static_cast<S*>(p.f1);
    const int nvec = p.n_rows * p.ntz;
    int i = blockIdx.x * BLOCK + threadIdx.x;
    const bool active = i < nvec;
    if (!active) i = nvec - 1;
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

    // ---- packed bc ids of my V cells; the per-cell kinds are kept as bit masks (one register each) because this
    // kernel lives on its register budget
    uint32_t idw[(V + 3) / 4];
    uint32_t m_full = 0, m_eq = 0, m_half = 0;
    bool any_bc = false, slow = false;
#pragma unroll
    for (int k = 0; k < (V + 3) / 4; ++k) idw[k] = 0;
    auto id_of = [&](int j) -> int { return (idw[j >> 2] >> (8 * (j & 3))) & 0xff; };
    if (HAS_BC) {
        if constexpr (V == 2) {
            idw[0] = *reinterpret_cast<const unsigned short*>(p.bc_mask + cell);
        } else if constexpr (V == 4) {
            idw[0] = *reinterpret_cast<const uint32_t*>(p.bc_mask + cell);
        } else {
            const uint2 t = *reinterpret_cast<const uint2*>(p.bc_mask + cell);
            idw[0] = t.x;
            idw[1] = t.y;
        }
    }
    auto classify = [&]() {
#pragma unroll
        for (int k = 0; k < (V + 3) / 4; ++k) any_bc |= idw[k] != 0;
        if (any_bc) {
            // Only the tagged cells are decoded, one per trip of a rolled loop: in a cavity two warps of every row touch a wall
            // cell at a z end, i.e. most warps come through here with ONE tagged cell in one or two lanes; decoding all V cells
            // under predicates cost those warps ~150 instructions (ncu, profiles/r2b_ncu_details_f16.txt), this costs ~25.
            uint32_t todo = 0;  // bit j: cell j is tagged
#pragma unroll
            for (int j = 0; j < V; ++j) todo |= (id_of(j) != 0 ? 1u : 0u) << j;
#pragma unroll 1
            while (todo) {
                const int j = __ffs(todo) - 1;
                todo &= todo - 1;
                const uint32_t word = (V > 4 && j >= 4) ? idw[(V + 3) / 4 - 1] : idw[0];  // a select, not an indexed (local-memory) access
                const int k = p.bct[(word >> (8 * (j & 3))) & 0xff].kind;
                m_full |= (k == XLB_BC_FULLWAY ? 1u : 0u) << j;
                m_eq |= (k == XLB_BC_EQUILIBRIUM ? 1u : 0u) << j;
                const bool here = p.cells == nullptr || !bc_is_listed(k);  // see step_vector_kernel
                if constexpr (HALF) {
                    m_half |= (k == XLB_BC_HALFWAY ? 1u : 0u) << j;
                    slow |= bc_is_complex(k) && k != XLB_BC_HALFWAY && here;
                } else {
                    slow |= bc_is_complex(k) && here;
                }
            }
        }
    };
#if !XLB_TUNE_H2_PHASED
    if (HAS_BC) classify();
#endif

    // ---- pull-stream: one aligned vector load per population, kept packed; the z shift is a byte permute
    const int xm = p.x_periodic ? wrap1(x - 1, p.nx) : x - 1;
    const int xp = p.x_periodic ? wrap1(x + 1, p.nx) : x + 1;
    const int ym = wrap1(y - 1, p.ny), yp = wrap1(y + 1, p.ny);
    const long long ox0 = x * plane, oxm = xm * plane, oxp = xp * plane;
    const int oy0 = y * p.nz, oym = ym * p.nz, oyp = yp * p.nz;
    const bool edge_lo = (lane == 0) || (tz == 0);
    const bool edge_hi = (lane == 31) || (tz == p.ntz - 1);
    const int z_lo = (tz == 0) ? p.nz - 1 : z0 - 1;
    const int z_hi = (tz == p.ntz - 1) ? 0 : z0 + V;
    PackedH2<Q, W> pk;
    uint32_t (&raw)[Q][W] = pk.w;
    const S* pl = f0;
#if XLB_TUNE_H2_ROWPTR
    const S* rowp[3][3];  // [c_x: 0, +1, -1][c_y: 0, +1, -1] -> population 0's source row
    {
        const long long oxs[3] = {ox0, oxm, oxp};
        const int oys[3] = {oy0, oym, oyp};
#pragma unroll
        for (int a = 0; a < 3; ++a)
#pragma unroll
            for (int b = 0; b < 3; ++b) rowp[a][b] = f0 + (oxs[a] + oys[b]);
    }
    long long poff = 0;  // l * q_stride: the same for every thread of the grid (uniform datapath)
#endif
#if XLB_TUNE_H2_PHASED
    // phase 1: issue every population's vector load before anything waits on one (the classification of tagged cells, the
    // shuffles and the exit of inactive lanes used to split the loads into three dependent round trips to DRAM: ids -> the
    // z-shifted populations -> the others; ncu source page, profiles/r2g_ncu_details_f16.txt)
    Words<W> vv[Q];
    static_for<Q>([&](auto lc) {
        XLB_IDX(l, lc);
        using D = Dir<Q, l>;
#if XLB_TUNE_H2_ROWPTR
        const S* row = rowp[D::cx == 0 ? 0 : (D::cx == 1 ? 1 : 2)][D::cy == 0 ? 0 : (D::cy == 1 ? 1 : 2)] + poff;
        poff += p.qs;
#else
        const long long ox = D::cx == 0 ? ox0 : (D::cx == 1 ? oxm : oxp);
        const int oy = D::cy == 0 ? oy0 : (D::cy == 1 ? oym : oyp);
        const S* row = pl + (ox + oy);
#endif
        vv[l] = ld_words<W>(row + z0);
        pl += p.qs;
    });
    if (HAS_BC) classify();
    // phase 2: the z shift (a byte permute; the element across the lane boundary comes by shuffle, across the warp boundary by
    // a predicated scalar load)
    pl = f0;
    static_for<Q>([&](auto lc) {
        XLB_IDX(l, lc);
        using D = Dir<Q, l>;
        const Words<W>& v = vv[l];
        if (D::cz == 0) {
#pragma unroll
            for (int k = 0; k < W; ++k) raw[l][k] = v.w[k];
        } else {
            const long long ox = D::cx == 0 ? ox0 : (D::cx == 1 ? oxm : oxp);
            const int oy = D::cy == 0 ? oy0 : (D::cy == 1 ? oym : oyp);
            const S* row = pl + (ox + oy);
            if (D::cz == 1) {  // pull from z-1: cell j takes element j-1; element -1 comes from the lane to the left
                const uint32_t prev = __shfl_up_sync(0xffffffffu, v.w[W - 1], 1);
                const uint32_t left = ld_if_u16(row + z_lo, edge_lo, prev >> 16);
                raw[l][0] = __byte_perm(left, v.w[0], 0x5410);
#pragma unroll
                for (int k = 1; k < W; ++k) raw[l][k] = __byte_perm(v.w[k - 1], v.w[k], 0x5432);
            } else {  // pull from z+1: cell j takes element j+1; element V comes from the lane to the right
                const uint32_t next = __shfl_down_sync(0xffffffffu, v.w[0], 1);
                const uint32_t right = ld_if_u16(row + z_hi, edge_hi, next & 0xffffu);
#pragma unroll
                for (int k = 0; k < W - 1; ++k) raw[l][k] = __byte_perm(v.w[k], v.w[k + 1], 0x5432);
                raw[l][W - 1] = __byte_perm(v.w[W - 1], right, 0x5432);
            }
        }
        pl += p.qs;
    });
#else
    static_for<Q>([&](auto lc) {
        XLB_IDX(l, lc);
        using D = Dir<Q, l>;
#if XLB_TUNE_H2_ROWPTR
        const S* row = rowp[D::cx == 0 ? 0 : (D::cx == 1 ? 1 : 2)][D::cy == 0 ? 0 : (D::cy == 1 ? 1 : 2)] + poff;
        poff += p.qs;
#else
        const long long ox = D::cx == 0 ? ox0 : (D::cx == 1 ? oxm : oxp);
        const int oy = D::cy == 0 ? oy0 : (D::cy == 1 ? oym : oyp);
        const S* row = pl + (ox + oy);
#endif
        const Words<W> v = ld_words<W>(row + z0);
        if (D::cz == 0) {
#pragma unroll
            for (int k = 0; k < W; ++k) raw[l][k] = v.w[k];
        } else if (D::cz == 1) {  // pull from z-1: cell j takes element j-1; element -1 comes from the lane to the left
            const uint32_t prev = __shfl_up_sync(0xffffffffu, v.w[W - 1], 1);
            const uint32_t left = ld_if_u16(row + z_lo, edge_lo, prev >> 16);
            raw[l][0] = __byte_perm(left, v.w[0], 0x5410);
#pragma unroll
            for (int k = 1; k < W; ++k) raw[l][k] = __byte_perm(v.w[k - 1], v.w[k], 0x5432);
        } else {  // pull from z+1: cell j takes element j+1; element V comes from the lane to the right
            const uint32_t next = __shfl_down_sync(0xffffffffu, v.w[0], 1);
            const uint32_t right = ld_if_u16(row + z_hi, edge_hi, next & 0xffffu);
#pragma unroll
            for (int k = 0; k < W - 1; ++k) raw[l][k] = __byte_perm(v.w[k], v.w[k + 1], 0x5432);
            raw[l][W - 1] = __byte_perm(v.w[W - 1], right, 0x5432);
        }
        pl += p.qs;
    });
#endif
    if (!active) return;

    if (HAS_BC && slow) {
        for (int j = 0; j < V; ++j) cell_update_generic<T, S, Q, COLL>(p, x, y, z0 + j);
        return;
    }
    const CollArgs<T> ca = coll_args<T>(p);
    // ---- one pair of cells at a time.  Pairs without streaming-step BC cells collide straight on the packed storage
    // (widened on access, narrowed once on the final set); pairs that hold an EquilibriumBC / half-way cell, and the
    // forced collisions (which read their own output), are widened to FP32 first so that every intermediate stays FP32.
    constexpr bool FORCED = (COLL & XLB_COLL_FORCED) != 0;
    auto do_pair = [&](auto jc, bool force_widen) {
        XLB_IDX(jp, jc);
        constexpr uint32_t PAIR = 3u << (2 * jp);
        const bool pair_bc = HAS_BC && (((m_eq | m_half) & PAIR) != 0);
        const bool pair_full = HAS_BC && ((m_full & PAIR) != 0);
        bool fw[2] = {false, false};
        if (pair_full) {
            fw[0] = (m_full >> (2 * jp)) & 1u;
            fw[1] = (m_full >> (2 * jp + 1)) & 1u;
        }
        constexpr bool WIDEN = XLB_TUNE_H2_WIDEN < 0 ? (V == 8) : (XLB_TUNE_H2_WIDEN != 0);
        if (FORCED || !XLB_PACKED(COLL) || pair_bc || WIDEN || force_widen) {
            T g[Q][2];
#pragma unroll
            for (int l = 0; l < Q; ++l) {
                const float2 t = h2_to_f2(raw[l][jp]);
                g[l][0] = t.x;
                g[l][1] = t.y;
            }
            if constexpr (HAS_BC && HALF) {
                if (m_half & PAIR) {
                    // HalfwayBounceBackBC, STREAMING step (bc_halfway_bounce_back.py:116-134), see step_vector_kernel
                    uint32_t mb[2] = {0, 0};
                    if (p.mbits) {
                        const Vec<uint32_t, 2> m = ld_vec<uint32_t, 2>(p.mbits + cell + 2 * jp);
                        mb[0] = m.v[0];
                        mb[1] = m.v[1];
                    } else {
                        static_for<Q>([&](auto lc) {
                            XLB_IDX(l, lc);
                            const Vec<uint8_t, 2> m = ld_vec<uint8_t, 2>(p.missing + l * p.mqs + cell + 2 * jp);
                            mb[0] |= (m.v[0] ? 1u : 0u) << l;
                            mb[1] |= (m.v[1] ? 1u : 0u) << l;
                        });
                    }
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
                    if ((m_eq >> (2 * jp + k)) & 1u) {  // EquilibriumBC (STREAMING step): feq(rho_bc, u_bc) constants
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
            for (int l = 0; l < Q; ++l) raw[l][jp] = f2_to_h2(g[l][0], g[l][1]);
        } else {
            collide_cols<F2, T, Q, COLL, V, 2 * jp>(pk, ca, fw, pair_full);
        }
    };
    if constexpr (!XLB_TUNE_STREAM_ONLY) static_for<W>([&](auto jc) { do_pair(jc, false); });
    // ---- one aligned vector store per population (+ the halo push of the populations that leave the slab here)
    S* ql = f1 + cell;
    PeerDst<S> pd{nullptr, 0, 0};
    if constexpr (PUSH) pd = peer_dst<S, Q>(p, x, y, (long long)y * p.nz + z0);  // see step_vector_kernel
    S* qp = pd.base;
    static_for<Q>([&](auto lc) {
        XLB_IDX(l, lc);
        st_words<W>(ql, raw[l]);
        ql += p.qs;
        if constexpr (PUSH) {
            constexpr int ca = slab_c<Q>(Dir<Q, l>::cx, Dir<Q, l>::cy);
            if (ca != 0 && pd.c == ca) st_words<W>(qp, raw[l]);
            qp += pd.qs;
        }
    });
}

}  // namespace xlb
