// xlb_step_bulk.cuh -- the fused step as a persistent, bulk-copy-staged pipeline (K1b).
//
// Same per-cell arithmetic as step_vector_kernel (xlb_step.cuh), different data movement: the q source rows of a
// tile (TZ = NT*C consecutive z cells of one (x, y) row) are fetched by the copy engine with 1-D bulk async copies
// (PTX cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes, SASS UBLKCP) into a STAGES-deep ring of
// shared-memory buffers, completion is signalled on an mbarrier per stage, and the block prefetches the tiles it
// will process next while it collides the current one.  What this buys over per-thread LDG:
//   * DRAM latency is hidden by the prefetch ring instead of by occupancy, so the register-hungry collisions
//     (D3Q27 KBC, FP16 storage) no longer stall on their own loads;
//   * no per-thread address arithmetic, shuffles or edge loads for the q pulled populations: lane l of warp 0 issues
//     the copy of population l (one row base + one 16-byte halo chunk where c_z != 0);
//   * the +-1 z shift of the pull is a shared-memory index offset; every thread owns C cells NT apart, so all
//     shared-memory reads are conflict-free 32-bit accesses and all global stores are fully coalesced.
// Requirements (checked on the host, otherwise the vector kernel runs): nz % TZ == 0, 16-byte aligned rows, no halo push
// (the edge-plane launch of a slab runs the vector kernel's PUSH instance).
#pragma once
#include "xlb_step.cuh"
#include "xlb_step_h2.cuh"

// 1: tiles are handed out dynamically (one atomicAdd on a global counter per tile) instead of by a static stride.  A tile
// that holds cells of the per-cell engine (Zou-He / extrapolation planes) costs ~10 plain tiles; with a static stride the
// launch ends when the unluckiest block does (BASELINE configs[2], 384x128x128: 256 of 888 blocks carry one such tile on top
// of 55 plain ones -> +16 % on the kernel), with the counter the other blocks take over the rest of its share.
#ifndef XLB_TUNE_DYN
#define XLB_TUNE_DYN 1
#endif

namespace xlb {

XLB_D uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

XLB_D void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
XLB_D void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
XLB_D void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
XLB_D void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// arrival counter of a stage: returns the previous value; acq_rel at CTA scope orders this warp's shared-memory reads of
// the stage before the refill that the last-arriving warp issues
XLB_D unsigned stage_arrive(unsigned* cnt) {
    unsigned old;
    asm volatile("atom.acq_rel.cta.shared::cta.add.u32 %0, [%1], 1;" : "=r"(old) : "r"(smem_u32(cnt)) : "memory");
    return old;
}
XLB_D void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "XLB_WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra XLB_DONE_%=;\n\t"
        "bra XLB_WAIT_%=;\n\t"
        "XLB_DONE_%=:\n\t}" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// global -> shared bulk copy, bytes % 16 == 0, both addresses 16-byte aligned; completes on `bar`
XLB_D void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)), "l"(src), "r"(bytes),
                 "r"(smem_u32(bar))
                 : "memory");
}

#ifndef XLB_TUNE_BULK_REVERSE
#define XLB_TUNE_BULK_REVERSE 0  // the backwards walk of odd steps (StepParams.reverse) pays in K1 / K1c / K1d; here it is neutral on wide lattices and costs configs[2] 1.7 % (profiles/r3a_sweep_bulk_reverse.log)
#endif
#ifndef XLB_TUNE_BULK_ROT
#define XLB_TUNE_BULK_ROT 1
#endif

template <typename S, int Q, int TZ, bool HAS_BC>
struct BulkLayout {
    static constexpr int HE = 16 / (int)sizeof(S);        // halo chunk, elements
    static constexpr int ROW = TZ + 2 * HE;               // elements per staged population row
    static constexpr int POP_BYTES = Q * ROW * (int)sizeof(S);
    static constexpr int MASK_BYTES = HAS_BC ? TZ : 0;
    static constexpr int STAGE_BYTES = ((POP_BYTES + MASK_BYTES + 127) / 128) * 128;
    XLB_HD static constexpr int n_shifted() {
        int n = 0;
        for (int l = 0; l < Q; ++l) n += Lat<Q>::c(l, 2) != 0;
        return n;
    }
    static constexpr uint32_t TX_BYTES = Q * TZ * sizeof(S) + n_shifted() * 16 + MASK_BYTES;
    static size_t smem_bytes(int stages) { return (size_t)stages * STAGE_BYTES + 128; }
};

// H2 (FP16 storage only): the staged rows are read as packed half2 words -- thread t owns the cell pairs
// (2t, 2t+1) + k*2*NT of the tile -- and stay packed in registers (PackedH2, see xlb_step_h2.cuh) through the collision.
template <typename T, typename S, int Q, int COLL, int C, int NT, int STAGES, int MINB, bool HAS_BC, bool HALF = false, bool H2 = false, bool MULTI = false>
__global__ void __launch_bounds__(NT, MINB) step_bulk_kernel(const __grid_constant__ StepParams p) {
    static_assert(!MULTI || (!H2 && C > 1), "multi-row tiles: C > 1 rows per tile, not for the packed-half2 consumer");
    static_assert(!H2 || (std::is_same<S, __half>::value && std::is_same<T, float>::value && C % 2 == 0), "H2: FP32 compute, FP16 storage, even C");
    constexpr int TZ = NT * C;
    using L = BulkLayout<S, Q, TZ, HAS_BC>;
    constexpr int HE = L::HE;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw + (size_t)STAGES * L::STAGE_BYTES);
    unsigned* arrived = reinterpret_cast<unsigned*>(full + STAGES);  // per stage: warps that have consumed it (monotonic)
    int* tile_of = reinterpret_cast<int*>(arrived + STAGES);         // per stage: the tile staged there (-1: no more tiles)
    int* tile_next = tile_of + STAGES;                                // per stage: the tile claimed for its NEXT refill (-1: none)
    unsigned* exited = reinterpret_cast<unsigned*>(tile_next + STAGES);  // warps that have left the tile loop
    const bool dyn = XLB_TUNE_DYN && p.sched != nullptr;
    constexpr unsigned NW = NT / 32;
    const S* __restrict__ f0 = static_cast<const S*>(p.f0);
    S* __restrict__ f1 = static_cast<S*>(p.f1);
    const int tid = threadIdx.x;
    const int lane = tid & 31;
    // multi-row tiles: on rows of exactly NT cells (nz == NT, e.g. the 128-wide BASELINE configs[2]) a tile is C whole,
    // consecutive rows -- thread t owns cell z = t of each -- so that the bulk copies are C rows long where the source rows are
    // adjacent in memory and the per-tile costs (copy issue, barrier, claim) are paid per C * NT cells like on wide rows
    // A template flag, not a run-time test: as a run-time test it cost the wide-row instances 2.5 % (profiles/r2v_bench_default.json).
    constexpr bool multi = MULTI;  // the host launches this instance only when p.nz == NT
    const int nseg = multi ? 1 : p.nz / TZ;
    const int tiles_per_plane = multi ? p.n_rows / C : p.n_rows * nseg;
    const int total = p.n_planes * tiles_per_plane;  // < 2^31, checked on the host
    const long long plane = (long long)p.ny * p.nz;

    // lattice vector of the population this lane issues copies for
    int lcx = 0, lcy = 0, lcz = 0;
    static_for<Q>([&](auto lc) {
        XLB_IDX(l, lc);
        if (lane == l) {
            lcx = Dir<Q, l>::cx;
            lcy = Dir<Q, l>::cy;
            lcz = Dir<Q, l>::cz;
        }
    });

    auto tile_coords = [&](int t, int& x, int& y, int& z0) {
        int xr = t / tiles_per_plane;
        const int r = t - xr * tiles_per_plane;
#if XLB_TUNE_BULK_ROT
        // claim order: the LAST plane first.  Inlet / outlet planes (x = 0, nx - 1) are the dear ones when their cells take the
        // per-cell engine in this kernel; in plain order the outlet's tiles were the last ones claimed and formed the tail.
        xr = xr == 0 ? p.n_planes - 1 : xr - 1;
#endif
#if XLB_TUNE_BULK_REVERSE
        // odd time steps walk the planes backwards (StepParams.reverse), keeping the two dear planes in front:
        // [n-1, 0, 1, ..., n-2] becomes [0, n-1, n-2, ..., 1]
        if (p.reverse) xr = p.n_planes - 1 - xr;
#endif
        const int yr = multi ? r * C : r / nseg;
        z0 = multi ? 0 : (r - yr * nseg) * TZ;
        y = p.y_begin + yr * p.y_stride;
        x = p.x_begin + xr * p.x_stride;
    };
    // one warp (all 32 lanes): fetch tile t into stage s
    // Halo of the z shift.  The packed-half2 consumer (H2) keeps the original scheme: one 16-byte chunk per shifted population
    // next to the segment.  Otherwise the chunk is MERGED into the population's main copy (one copy of TZ + HE elements that
    // starts HE elements early, or ends HE elements late) -- except at the two ends of a row, where the periodic neighbour is at
    // the other end: a separate 16-byte copy there, or, when the row is one single tile (nseg == 1, e.g. the 128-wide
    // BASELINE configs[2]), no halo copy at all: the neighbour is inside the staged row and the consumer wraps its index.
    // Per tile this is Q + 1 bulk copies instead of Q + n_shifted + 1 (28 instead of 46 for D3Q27).
    constexpr bool MERGE = !H2;
    const bool one_seg = MERGE && nseg == 1;
    const uint32_t tx_bytes = one_seg ? (uint32_t)(Q * TZ * sizeof(S) + L::MASK_BYTES) : L::TX_BYTES;
    auto issue = [&](int t, int s) {
        if (t >= total) return;
        int x, y, z0;
        tile_coords(t, x, y, z0);
        unsigned char* stage = smem_raw + (size_t)s * L::STAGE_BYTES;
        if (lane == 0) mbar_arrive_expect_tx(&full[s], tx_bytes);
        __syncwarp();
        if (lane < Q) {
            int xs = x - lcx;
            if (p.x_periodic) xs = wrap1(xs, p.nx);
            const int ys = wrap1(y - lcy, p.ny);
            const S* row = f0 + lane * p.qs + (xs * plane + (long long)ys * p.nz);
            S* dst = reinterpret_cast<S*>(stage) + lane * L::ROW;
            const bool first = z0 == 0, last = z0 + TZ == p.nz;
            if (multi && ys + C > p.ny) {  // the C source rows straddle the periodic y boundary: one copy per row
#pragma unroll
                for (int r = 0; r < C; ++r)
                    bulk_g2s(dst + HE + r * NT, f0 + lane * p.qs + (xs * plane + (long long)wrap1(y + r - lcy, p.ny) * p.nz), NT * sizeof(S), &full[s]);
            } else if (lcz == 0 || one_seg) {
                bulk_g2s(dst + HE, row + z0, TZ * sizeof(S), &full[s]);
            } else if (lcz == 1) {  // pull from z-1: HE elements left of the segment (periodic)
                if (MERGE && !first) {
                    bulk_g2s(dst, row + z0 - HE, (TZ + HE) * sizeof(S), &full[s]);
                } else {
                    bulk_g2s(dst + HE, row + z0, TZ * sizeof(S), &full[s]);
                    bulk_g2s(dst, row + (first ? p.nz - HE : z0 - HE), 16, &full[s]);
                }
            } else {  // pull from z+1: HE elements right of the segment (periodic)
                if (MERGE && !last) {
                    bulk_g2s(dst + HE, row + z0, (TZ + HE) * sizeof(S), &full[s]);
                } else {
                    bulk_g2s(dst + HE, row + z0, TZ * sizeof(S), &full[s]);
                    bulk_g2s(dst + HE + TZ, row + (last ? 0 : z0 + TZ), 16, &full[s]);
                }
            }
        } else if (HAS_BC && lane == Q) {
            bulk_g2s(stage + L::POP_BYTES, p.bc_mask + (x * plane + (long long)y * p.nz + z0), TZ, &full[s]);
        }
    };

    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full[s], 1);
            arrived[s] = 0;
            const long long t0 = (long long)blockIdx.x + (long long)s * gridDim.x;  // the first STAGES rounds are static
            tile_of[s] = t0 < total ? (int)t0 : -1;
            tile_next[s] = -1;
            if (dyn && t0 < total) {  // the tile this stage gets at its first refill
                const long long nt = (long long)STAGES * gridDim.x + atomicAdd(p.sched, 1u);
                tile_next[s] = nt < total ? (int)nt : -1;
            }
        }
        *exited = 0;
        mbar_fence_init();
    }
    __syncthreads();
    if (tid < 32) {
#pragma unroll
        for (int s = 0; s < STAGES; ++s) {
            if (tile_of[s] >= 0) issue(tile_of[s], s);
            else if (dyn && lane == 0) mbar_arrive(&full[s]);  // nothing to stage: complete the phase so that nobody waits on it
        }
    }
    // refill of stage s by the warp that consumed it last.  Static: the tile STAGES rounds ahead in the block's stride.  Dynamic:
    // the tile in tile_next[s], which was claimed from the global counter at the previous refill of this stage; the claim for
    // the refill after this one is issued right here (lane 0) but its result is only stored at the end of the iteration
    // (flush_claim), so the atomic's round trip to L2 overlaps the collision instead of delaying the copies.  tile_of[s] /
    // tile_next[s] are written before an arrive on the stage (release) and read after the wait / last arrival (acquire).
    unsigned claim_val = 0;
    int claim_stage = -1;
    auto refill = [&](int t_done, int s) {
        if (!dyn) {
            if ((long long)t_done + (long long)STAGES * gridDim.x < total) issue(t_done + STAGES * (int)gridDim.x, s);
            return;
        }
        const int nt = tile_next[s];
        if (nt >= 0) {
            if (lane == 0) tile_of[s] = nt;
            issue(nt, s);
            if (lane == 0) {
                claim_val = atomicAdd(p.sched, 1u);
                claim_stage = s;
            }
        } else if (lane == 0) {
            tile_of[s] = -1;
            mbar_arrive(&full[s]);
        }
    };
    auto flush_claim = [&]() {
        if (claim_stage >= 0) {
            const long long nt = (long long)STAGES * gridDim.x + claim_val;
            tile_next[claim_stage] = nt < total ? (int)nt : -1;
            claim_stage = -1;
        }
    };

    const CollArgs<T> ca = coll_args<T>(p);
    int it = 0;
    for (int t = blockIdx.x;; ++it) {
        const int s = it % STAGES;
        const uint32_t parity = (uint32_t)(it / STAGES) & 1u;
        if (dyn) {
            mbar_wait(&full[s], parity);
            t = tile_of[s];
            if (t < 0) break;
        } else {
            if (it > 0) t += gridDim.x;
            if (t >= total) break;
        }
        int x, y, z0;
        tile_coords(t, x, y, z0);
        const unsigned char* stage = smem_raw + (size_t)s * L::STAGE_BYTES;
        const S* pops = reinterpret_cast<const S*>(stage);
        if (!dyn) mbar_wait(&full[s], parity);
        if constexpr (H2) {
            // ---- staged rows -> packed registers.  Pair k of thread t = cells z0 + 2t + k*2NT (+1); a pull shift of +-1
            // straddles two 32-bit words of the staged row: two conflict-free LDS.32 and one byte permute.
            constexpr int W = C / 2;
            PackedH2<Q, W> pk;
            const uint32_t* words = reinterpret_cast<const uint32_t*>(stage);
            static_for<Q>([&](auto lc) {
                XLB_IDX(l, lc);
                constexpr int cz = Dir<Q, l>::cz;
                constexpr int base = (l * L::ROW + HE) / 2;  // word index of cell 0 of the tile in row l (ROW and HE are even)
#pragma unroll
                for (int k = 0; k < W; ++k) {
                    const int w = base + tid + k * NT;
                    if (cz == 0) pk.w[l][k] = words[w];
                    else if (cz == 1) pk.w[l][k] = __byte_perm(words[w - 1], words[w], 0x5432);  // elements 2t-1, 2t
                    else pk.w[l][k] = __byte_perm(words[w], words[w + 1], 0x5432);                  // elements 2t+1, 2t+2
                }
            });
            uint32_t idp[W];  // the two bc ids of each pair
            bool any_bc = false;
#pragma unroll
            for (int k = 0; k < W; ++k) {
                idp[k] = HAS_BC ? *reinterpret_cast<const unsigned short*>(stage + L::POP_BYTES + 2 * (tid + k * NT)) : 0u;
                any_bc |= idp[k] != 0;
            }
            __syncwarp();
            unsigned prev = 0;
            if (lane == 0) prev = stage_arrive(&arrived[s]);
            prev = __shfl_sync(0xffffffffu, prev, 0);
            if ((prev + 1) % NW == 0) refill(t, s);

            static_for<W>([&](auto kc) {
                XLB_IDX(k, kc);
                const int zc = z0 + 2 * (tid + k * NT);
                const long long cell0 = x * plane + (long long)y * p.nz + zc;
                bool fw[2] = {false, false}, eq[2] = {false, false}, hf[2] = {false, false};
                bool slow = false;
                int id2[2] = {(int)(idp[k] & 0xffu), (int)(idp[k] >> 8)};
                if (HAS_BC && idp[k] != 0) {
#pragma unroll
                    for (int j = 0; j < 2; ++j) {
                        if (id2[j]) {
                            const int kind = p.bct[id2[j]].kind;
                            fw[j] = kind == XLB_BC_FULLWAY;
                            eq[j] = kind == XLB_BC_EQUILIBRIUM;
                            const bool here = p.cells == nullptr || !bc_is_listed(kind);
                            if constexpr (HALF) {
                                hf[j] = kind == XLB_BC_HALFWAY;
                                slow |= bc_is_complex(kind) && !hf[j] && here;
                            } else {
                                slow |= bc_is_complex(kind) && here;
                            }
                        }
                    }
                }
                if (HAS_BC && slow) {
                    cell_update_generic<T, S, Q, COLL>(p, x, y, zc);
                    cell_update_generic<T, S, Q, COLL>(p, x, y, zc + 1);
                } else {
                    constexpr bool FORCED = (COLL & XLB_COLL_FORCED) != 0;
                    const bool pair_full = HAS_BC && (fw[0] || fw[1]);
                    if (FORCED || !XLB_PACKED(COLL) || (HAS_BC && (eq[0] || eq[1] || hf[0] || hf[1]))) {
                        T g[Q][2];
#pragma unroll
                        for (int l = 0; l < Q; ++l) {
                            const float2 v2 = h2_to_f2(pk.w[l][k]);
                            g[l][0] = v2.x;
                            g[l][1] = v2.y;
                        }
                        if constexpr (HAS_BC && HALF) {
                            if (hf[0] || hf[1]) {  // HalfwayBounceBackBC, STREAMING step (bc_halfway_bounce_back.py:116-134)
                                uint32_t mb[2] = {0, 0};
                                if (p.mbits) {
                                    const Vec<uint32_t, 2> m = ld_vec<uint32_t, 2>(p.mbits + cell0);
                                    mb[0] = m.v[0];
                                    mb[1] = m.v[1];
                                } else {
                                    static_for<Q>([&](auto lc) {
                                        XLB_IDX(l, lc);
                                        const Vec<uint8_t, 2> m = ld_vec<uint8_t, 2>(p.missing + l * p.mqs + cell0);
                                        mb[0] |= (m.v[0] ? 1u : 0u) << l;
                                        mb[1] |= (m.v[1] ? 1u : 0u) << l;
                                    });
                                }
                                if (!hf[0]) mb[0] = 0;
                                if (!hf[1]) mb[1] = 0;
                                const double* tab0 = p.bct[id2[0]].table;
                                const double* tab1 = p.bct[id2[1]].table;
                                static_for<Q>([&](auto lc) {
                                    XLB_IDX(l, lc);
                                    const float2 pre = h2_to_f2(*reinterpret_cast<const uint32_t*>(f0 + Dir<Q, l>::o * p.qs + cell0));
                                    if (mb[0] & (1u << l)) g[l][0] = pre.x + static_cast<T>(tab0[l]);
                                    if (mb[1] & (1u << l)) g[l][1] = pre.y + static_cast<T>(tab1[l]);
                                });
                            }
                        }
#pragma unroll
                        for (int j = 0; j < 2; ++j) {
                            if (HAS_BC && eq[j]) {  // EquilibriumBC (STREAMING step)
                                const double* tab = p.bct[id2[j]].table;
#pragma unroll
                                for (int l = 0; l < Q; ++l) g[l][j] = static_cast<T>(tab[l]);
                            }
                        }
                        if constexpr (XLB_PACKED(COLL)) {
                            collide_cols<F2, T, Q, COLL, 2, 0>(g, ca, fw, pair_full);
                        } else {
                            collide_cols<T, T, Q, COLL, 2, 0>(g, ca, &fw[0], fw[0]);
                            collide_cols<T, T, Q, COLL, 2, 1>(g, ca, &fw[1], fw[1]);
                        }
#pragma unroll
                        for (int l = 0; l < Q; ++l) pk.w[l][k] = f2_to_h2(g[l][0], g[l][1]);
                    } else {
                        collide_cols<F2, T, Q, COLL, C, 2 * k>(pk, ca, fw, pair_full);
                    }
                    uint32_t* ql = reinterpret_cast<uint32_t*>(f1 + cell0);
#pragma unroll
                    for (int l = 0; l < Q; ++l) {
                        *ql = pk.w[l][k];
                        ql += p.qs / 2;
                    }
                }
            });
            flush_claim();
            continue;
        }
        // ---- staged rows -> registers (pull shift = index offset)
        T f[Q][C];
        static_for<Q>([&](auto lc) {
            XLB_IDX(l, lc);
            constexpr int cz = Dir<Q, l>::cz;
#pragma unroll
            for (int j = 0; j < C; ++j) {
                int idx = tid + j * NT - cz;
                if (multi) {  // row j of the tile, periodic within the row
                    if (cz == 1 && tid == 0) idx = j * NT + NT - 1;
                    if (cz == -1 && tid == NT - 1) idx = j * NT;
                } else {
                    if (cz == 1 && j == 0) idx = (one_seg && idx < 0) ? TZ - 1 : idx;           // z = 0 pulls from z = nz - 1
                    if (cz == -1 && j == C - 1) idx = (one_seg && idx == TZ) ? 0 : idx;         // z = nz - 1 pulls from z = 0
                }
                f[l][j] = to_compute<T, S>(pops[l * L::ROW + HE + idx]);
            }
        });
        uint8_t ids[C];
        bool any_bc = false;
        if (HAS_BC) {
#pragma unroll
            for (int j = 0; j < C; ++j) {
                ids[j] = stage[L::POP_BYTES + tid + j * NT];
                any_bc |= ids[j] != 0;
            }
        }
        // No block-wide barrier: each warp counts itself off the stage once its reads are done, and the warp that
        // arrives last refills the stage with the tile STAGES iterations ahead.  Warps never wait for each other.
        __syncwarp();
        unsigned prev = 0;
        if (lane == 0) prev = stage_arrive(&arrived[s]);
        prev = __shfl_sync(0xffffffffu, prev, 0);
        if ((prev + 1) % NW == 0) refill(t, s);

        bool slow = false, any_half = false;
        bool k_full[C], k_eq[C], k_half[C];
#pragma unroll
        for (int j = 0; j < C; ++j) k_full[j] = k_eq[j] = k_half[j] = false;
        if (HAS_BC && any_bc) {
#pragma unroll
            for (int j = 0; j < C; ++j) {
                if (ids[j]) {
                    const int k = p.bct[ids[j]].kind;
                    k_full[j] = k == XLB_BC_FULLWAY;
                    k_eq[j] = k == XLB_BC_EQUILIBRIUM;
                    const bool here = p.cells == nullptr || !bc_is_listed(k);  // see step_vector_kernel
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
        const int zc = z0 + tid;
        if (HAS_BC && slow) {
            for (int j = 0; j < C; ++j) {
                if (multi) cell_update_generic<T, S, Q, COLL>(p, x, y + j, zc);
                else cell_update_generic<T, S, Q, COLL>(p, x, y, zc + j * NT);
            }
            flush_claim();
            continue;
        }
        if constexpr (HAS_BC && HALF) {
            if (any_half) {
                // HalfwayBounceBackBC, STREAMING step (bc_halfway_bounce_back.py:116-134), coalesced and in two phases of
                // independent loads: all mask bytes, then all opposite populations of the same cell
                const long long cell0 = x * plane + (long long)y * p.nz + zc;
                uint32_t mb[C];
#pragma unroll
                for (int j = 0; j < C; ++j) mb[j] = 0;
                if (p.mbits) {
#pragma unroll
                    for (int j = 0; j < C; ++j) mb[j] = p.mbits[cell0 + j * NT];
                } else {
                    static_for<Q>([&](auto lc) {
                        XLB_IDX(l, lc);
#pragma unroll
                        for (int j = 0; j < C; ++j) mb[j] |= (p.missing[l * p.mqs + cell0 + j * NT] ? 1u : 0u) << l;
                    });
                }
#pragma unroll
                for (int j = 0; j < C; ++j)
                    if (!k_half[j]) mb[j] = 0;
                static_for<Q>([&](auto lc) {
                    XLB_IDX(l, lc);
#pragma unroll
                    for (int j = 0; j < C; ++j) {
                        const T g = to_compute<T, S>(f0[Dir<Q, l>::o * p.qs + cell0 + j * NT]);
                        if (mb[j] & (1u << l)) f[l][j] = g + static_cast<T>(p.bct[ids[j]].table[l]);
                    }
                });
            }
        }
        if (HAS_BC && any_bc) {
#pragma unroll
            for (int j = 0; j < C; ++j) {
                if (k_eq[j]) {
                    const double* tab = p.bct[ids[j]].table;
#pragma unroll
                    for (int l = 0; l < Q; ++l) f[l][j] = static_cast<T>(tab[l]);
                }
            }
        }
        if constexpr (XLB_PACKED(COLL) && std::is_same<T, float>::value && (C % 2 == 0)) {
            static_for<C / 2>([&](auto jc) {
                XLB_IDX(jp, jc);
                collide_cols<F2, T, Q, COLL, C, 2 * jp>(f, ca, &k_full[2 * jp], HAS_BC && (k_full[2 * jp] || k_full[2 * jp + 1]));
            });
        } else {
            static_for<C>([&](auto jc) {
                XLB_IDX(j, jc);
                collide_cols<T, T, Q, COLL, C, j>(f, ca, &k_full[j], HAS_BC && k_full[j]);
            });
        }
        S* ql = f1 + (x * plane + (long long)y * p.nz + zc);
#pragma unroll
        for (int l = 0; l < Q; ++l) {
#pragma unroll
            for (int j = 0; j < C; ++j) ql[j * NT] = to_store<S, T>(f[l][j]);
            ql += p.qs;
        }
        flush_claim();
    }
    if (dyn) {
        // the launch's last warp to leave puts the scheduler words back to zero for the next launch that gets this slot: every
        // counter increment of a block precedes the exit of its last warp, every block's exit precedes the reset
        __syncwarp();
        if (lane == 0) {
            unsigned left;
            asm volatile("atom.acq_rel.cta.shared::cta.add.u32 %0, [%1], 1;" : "=r"(left) : "r"(smem_u32(exited)) : "memory");
            if (left == NW - 1) {
                __threadfence();
                if (atomicAdd(p.sched + 1, 1u) == gridDim.x - 1) {
                    __threadfence();
                    p.sched[0] = 0u;
                    p.sched[1] = 0u;
                }
            }
        }
    }
}

}  // namespace xlb
