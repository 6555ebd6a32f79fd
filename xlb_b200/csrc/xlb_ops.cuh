// xlb_ops.cuh -- stand-alone operator kernels (K3a-f) and the device-side indices masker (K4).
// One cell per thread, runtime-typed element access (ld_any / st_any): these are the API-surface
// operators of the reference (each can be called on its own, e.g. from post-processing or tests),
// not the timed hot path -- that is the fused kernel in xlb_step.cuh.
#pragma once
#include "xlb_math.cuh"

namespace xlb {

struct OpParams {
    const void* a;  // input 0
    const void* b;  // input 1
    void* o0;       // output 0
    void* o1;       // output 1
    const uint8_t* bc_mask;
    const uint8_t* missing;
    const XlbBcEntry* bct;
    int nx, ny, nz;
    int x_periodic;
    int dt_in, dt_out;
    int id, flag, flag2;
    long long qs, mqs, ncell;
    double omega;
    double les_cs;  // SmagorinskyLESBGK coefficient
};

XLB_D int wrap1o(int v, int n) { return v < 0 ? v + n : (v >= n ? v - n : v); }

#define XLB_CELL_PROLOGUE                                                      \
    const long long cell = (long long)blockIdx.x * blockDim.x + threadIdx.x;   \
    if (cell >= p.ncell) return;                                               \
    const int z = (int)(cell % p.nz);                                          \
    const int y = (int)((cell / p.nz) % p.ny);                                 \
    const int x = (int)(cell / ((long long)p.nz * p.ny));                      \
    (void)x; (void)y; (void)z;

// Stream (stream.py:29-62): pure shifted copy, bit-exact for any element type
template <int Q>
__global__ void __launch_bounds__(256) stream_kernel(const __grid_constant__ OpParams p) {
    XLB_CELL_PROLOGUE
    const int esz = p.dt_in == XLB_F64 ? 8 : (p.dt_in == XLB_F32 ? 4 : (p.dt_in == XLB_F16 ? 2 : 1));
    static_for<Q>([&](auto lc) {
        XLB_IDX(l, lc);
        using D = Dir<Q, l>;
        const int xs = wrap1o(x - D::cx, p.nx);
        const int ys = wrap1o(y - D::cy, p.ny);
        const int zs = wrap1o(z - D::cz, p.nz);
        const long long src = l * p.qs + ((long long)xs * p.ny + ys) * p.nz + zs;
        const long long dst = l * p.qs + cell;
        if (esz == 4)
            static_cast<uint32_t*>(p.o0)[dst] = static_cast<const uint32_t*>(p.a)[src];
        else if (esz == 8)
            static_cast<uint64_t*>(p.o0)[dst] = static_cast<const uint64_t*>(p.a)[src];
        else if (esz == 2)
            static_cast<uint16_t*>(p.o0)[dst] = static_cast<const uint16_t*>(p.a)[src];
        else
            static_cast<uint8_t*>(p.o0)[dst] = static_cast<const uint8_t*>(p.a)[src];
    });
}

// Macroscopic / ZeroMoment / FirstMoment
template <typename T, int Q>
__global__ void __launch_bounds__(256) macroscopic_kernel(const __grid_constant__ OpParams p) {
    XLB_CELL_PROLOGUE
    T f[Q][1];
#pragma unroll
    for (int l = 0; l < Q; ++l) f[l][0] = ld_any<T>(p.a, p.dt_in, l * p.qs + cell);
    T rho, u[3];
    moments<T, Q, 1, 0>(f, rho, u);
    if (p.o0) st_any<T>(p.o0, p.dt_out, cell, rho);
    if (p.o1) {
#pragma unroll
        for (int a = axis0<Q>(); a < 3; ++a) st_any<T>(p.o1, p.dt_out, (a - axis0<Q>()) * p.ncell + cell, u[a]);
    }
}

// SecondMoment: pi[k] = sum_l cc_lk f_l, k over (a<=b) of the lattice dimensions
template <typename T, int Q>
__global__ void __launch_bounds__(256) second_moment_kernel(const __grid_constant__ OpParams p) {
    XLB_CELL_PROLOGUE
    T pi[6] = {T(0), T(0), T(0), T(0), T(0), T(0)};
    static_for<Q>([&](auto lc) {
        XLB_IDX(l, lc);
        acc_second_moment<Q, l>(pi, ld_any<T>(p.a, p.dt_in, l * p.qs + cell));
    });
    int out = 0, k = 0;
#pragma unroll
    for (int a = 0; a < 3; ++a)
#pragma unroll
        for (int b = a; b < 3; ++b) {
            if (a >= axis0<Q>()) st_any<T>(p.o0, p.dt_out, (long long)(out++) * p.ncell + cell, pi[k]);
            ++k;
        }
}

// QuadraticEquilibrium
template <typename T, int Q>
__global__ void __launch_bounds__(256) equilibrium_kernel(const __grid_constant__ OpParams p) {
    XLB_CELL_PROLOGUE
    const T rho = ld_any<T>(p.a, p.dt_in, cell);
    T u[3] = {T(0), T(0), T(0)};
#pragma unroll
    for (int a = axis0<Q>(); a < 3; ++a) u[a] = ld_any<T>(p.b, p.dt_in, (a - axis0<Q>()) * p.ncell + cell);
    const T usqr = usqr_of<T>(u, axis0<Q>());
    static_for<Q>([&](auto lc) {
        XLB_IDX(l, lc);
        st_any<T>(p.o0, p.dt_out, l * p.qs + cell, feq_dir<T, Q, l>(rho, u, usqr));
    });
}

// BGK / KBC with feq given (bgk.py:27-32, kbc.py:40-79)
template <typename T, int Q, int COLL>
__global__ void __launch_bounds__(128) collide_kernel(const __grid_constant__ OpParams p) {
    XLB_CELL_PROLOGUE
    T f[Q], fe[Q];
#pragma unroll
    for (int l = 0; l < Q; ++l) {
        f[l] = ld_any<T>(p.a, p.dt_in, l * p.qs + cell);
        fe[l] = ld_any<T>(p.b, p.dt_in, l * p.qs + cell);
    }
    const T omega = static_cast<T>(p.omega);
    if (COLL == XLB_BGK) {
#pragma unroll
        for (int l = 0; l < Q; ++l) st_any<T>(p.o0, p.dt_out, l * p.qs + cell, f[l] - omega * (f[l] - fe[l]));
    } else if (COLL == XLB_SMAGORINSKY_LES_BGK) {  // smagorinsky_les_bgk.py:44-66, same operation order
        T pi[6] = {T(0), T(0), T(0), T(0), T(0), T(0)};
        static_for<Q>([&](auto lc) {
            XLB_IDX(l, lc);
            acc_second_moment<Q, l>(pi, f[l] - fe[l]);
        });
        T dsum = pi[3] * pi[3] + pi[5] * pi[5], osum = pi[4] * pi[4];
        if (axis0<Q>() == 0) {
            dsum = pi[0] * pi[0] + pi[3] * pi[3] + pi[5] * pi[5];
            osum = pi[1] * pi[1] + pi[2] * pi[2] + pi[4] * pi[4];
        }
        const T strain = dsum + T(2.0) * osum;
        const T tau0 = T(1.0) / omega;
        const T cs = static_cast<T>(p.les_cs);
        const T tau = T(0.5) * (tau0 + sqrt(tau0 * tau0 + T(36.0) * (cs * cs) * sqrt(strain)));
        const T omega_eff = T(1.0) / tau;
#pragma unroll
        for (int l = 0; l < Q; ++l) st_any<T>(p.o0, p.dt_out, l * p.qs + cell, f[l] - omega_eff * (f[l] - fe[l]));
    } else {
        T pi[6] = {T(0), T(0), T(0), T(0), T(0), T(0)};
        static_for<Q>([&](auto lc) {
            XLB_IDX(l, lc);
            acc_second_moment<Q, l>(pi, f[l] - fe[l]);
        });
        const T beta = T(0.5) * omega;
        const T inv_beta = T(1.0) / beta;
        T sp1 = T(0), sp2 = T(0);
        static_for<Q>([&](auto lc) {
            XLB_IDX(l, lc);
            const T ds = kbc_delta_s<T, Q, l>(pi);
            const T dh = (f[l] - fe[l]) - ds;
            const T temp = dh / fe[l];
            sp1 += temp * ds;
            sp2 += temp * dh;
        });
        const T gamma = inv_beta - (T(2.0) - inv_beta) * sp1 / (T(1e-32) + sp2);
        static_for<Q>([&](auto lc) {
            XLB_IDX(l, lc);
            const T ds = kbc_delta_s<T, Q, l>(pi);
            const T dh = (f[l] - fe[l]) - ds;
            st_any<T>(p.o0, p.dt_out, l * p.qs + cell, f[l] - beta * (T(2.0) * ds + gamma * dh));
        });
    }
}

// One BC kind applied to one tagged cell: f = f_post (updated in place), g = f_pre, JAX semantics of bc_*.py.
template <typename T, int Q>
XLB_D void bc_apply_cell(int kind, const XlbBcEntry& e, T (&f)[Q][1], const T (&g)[Q], uint32_t miss, int x, int y, int z) {
    if (kind == XLB_BC_FULLWAY) {
        static_for<Q>([&](auto lc) {
            XLB_IDX(l, lc);
            f[l][0] = g[Dir<Q, l>::o];
        });
    } else if (kind == XLB_BC_EQUILIBRIUM) {
#pragma unroll
        for (int l = 0; l < Q; ++l) f[l][0] = static_cast<T>(e.table[l]);
    } else if (kind == XLB_BC_DONOTHING) {
#pragma unroll
        for (int l = 0; l < Q; ++l) f[l][0] = g[l];
    } else if (kind == XLB_BC_HALFWAY || kind == XLB_BC_EXTRAPOLATION) {
        static_for<Q>([&](auto lc) {
            XLB_IDX(l, lc);
            constexpr int o = Dir<Q, l>::o;
            if (miss & (1u << l)) f[l][0] = (kind == XLB_BC_HALFWAY) ? g[o] + static_cast<T>(e.table[l]) : g[o];
        });
    } else if (kind == XLB_BC_ZOUHE_VELOCITY || kind == XLB_BC_ZOUHE_PRESSURE || kind == XLB_BC_REGULARIZED_VELOCITY || kind == XLB_BC_REGULARIZED_PRESSURE) {
        zouhe_apply<T, Q>(f, miss, e, kind == XLB_BC_ZOUHE_VELOCITY || kind == XLB_BC_REGULARIZED_VELOCITY, x, y, z, kind >= XLB_BC_REGULARIZED_VELOCITY);
    }
}

// BoundaryCondition.__call__ / assemble_auxiliary_data for one BC id, whole grid, like the
// generic Warp BC kernel (boundary_condition.py:146-180) with the JAX semantics of bc_*.py.
template <typename T, int Q>
__global__ void __launch_bounds__(128) apply_bc_kernel(const __grid_constant__ OpParams p) {
    XLB_CELL_PROLOGUE
    const bool mine = p.bc_mask[cell] == p.id;
    if (!mine) {
        if (p.o0 != p.b) {
#pragma unroll
            for (int l = 0; l < Q; ++l) st_any<T>(p.o0, p.dt_out, l * p.qs + cell, ld_any<T>(p.b, p.dt_in, l * p.qs + cell));
        }
        return;
    }
    const XlbBcEntry& e = p.bct[p.id];
    const int kind = e.kind;
    T f[Q][1], g[Q];  // f = f_post, g = f_pre
    uint32_t miss = 0;
#pragma unroll
    for (int l = 0; l < Q; ++l) {
        f[l][0] = ld_any<T>(p.b, p.dt_in, l * p.qs + cell);
        g[l] = ld_any<T>(p.a, p.dt_in, l * p.qs + cell);
        miss |= (p.missing[l * p.mqs + cell] ? 1u : 0u) << l;
    }
    if (p.flag) {  // assemble_auxiliary_data: f_pre = post-stream, f_post = post-collision
        if (kind == XLB_BC_EXTRAPOLATION) {
            const T cs = T(1.0) / sqrt(T(3.0));
            const int xn = wrap1o(x - e.normal[0], p.nx), yn = wrap1o(y - e.normal[1], p.ny), zn = wrap1o(z - e.normal[2], p.nz);
            const long long ncell = ((long long)xn * p.ny + yn) * p.nz + zn;
            static_for<Q>([&](auto lc) {
                XLB_IDX(l, lc);
                if (miss & (1u << l)) {
                    const T nb = ld_any<T>(p.a, p.dt_in, l * p.qs + ncell);
                    f[Dir<Q, l>::o][0] = cs * nb + (T(1.0) - cs) * g[l];
                }
            });
        }
    } else {
        bc_apply_cell<T, Q>(kind, e, f, g, miss, x, y, z);
    }
#pragma unroll
    for (int l = 0; l < Q; ++l) st_any<T>(p.o0, p.dt_out, l * p.qs + cell, f[l][0]);
}

// MomentumTransfer (momentum_transfer.py:174-211) with FetchPopulations (:63-70) for the dense stream-then-collide
// stepper: on cells tagged with the no-slip BC whose centre is not solid, f_post_stream = BC(stream(f_0)) and the
// momentum exchanged along every missing direction l is phi_l = f_0[opp l] + f_post_stream[l], pointing along c_opp(l).
// Per-thread sums in FP64, warp shuffle + one atomicAdd per warp into force[3] (kernel axes).  p.id = BC id,
// [p.flag, p.flag2) = plane range that contributes (ghost planes of a slab are skipped).
template <typename T, int Q>
__global__ void __launch_bounds__(128) momentum_transfer_kernel(const __grid_constant__ OpParams p) {
    const long long cell = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    double m[3] = {0.0, 0.0, 0.0};
    if (cell < p.ncell && p.bc_mask[cell] == p.id && !p.missing[cell]) {
        const int z = (int)(cell % p.nz);
        const int y = (int)((cell / p.nz) % p.ny);
        const int x = (int)(cell / ((long long)p.nz * p.ny));
        if (x >= p.flag && x < p.flag2) {
            const XlbBcEntry& e = p.bct[p.id];
            T f[Q][1], g[Q];
            uint32_t miss = 0;
            static_for<Q>([&](auto lc) {
                XLB_IDX(l, lc);
                using D = Dir<Q, l>;
                g[l] = ld_any<T>(p.a, p.dt_in, l * p.qs + cell);
                const int xs = p.x_periodic ? wrap1o(x - D::cx, p.nx) : x - D::cx;  // ghost planes hold the neighbours
                const int ys = wrap1o(y - D::cy, p.ny);
                const int zs = wrap1o(z - D::cz, p.nz);
                f[l][0] = ld_any<T>(p.a, p.dt_in, l * p.qs + ((long long)xs * p.ny + ys) * p.nz + zs);
                miss |= (p.missing[l * p.mqs + cell] ? 1u : 0u) << l;
            });
            bc_apply_cell<T, Q>(e.kind, e, f, g, miss, x, y, z);
            static_for<Q>([&](auto lc) {
                XLB_IDX(l, lc);
                using D = Dir<Q, l>;
                if (miss & (1u << l)) {
                    const double phi = static_cast<double>(g[D::o] + f[l][0]);
                    m[0] -= D::cx * phi;  // c_opp(l) = -c_l
                    m[1] -= D::cy * phi;
                    m[2] -= D::cz * phi;
                }
            });
        }
    }
#pragma unroll
    for (int a = 0; a < 3; ++a) {
#pragma unroll
        for (int s = 16; s > 0; s >>= 1) m[a] += __shfl_down_sync(0xffffffffu, m[a], s);
    }
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int a = 0; a < 3; ++a)
            if (m[a] != 0.0) atomicAdd(static_cast<double*>(p.o0) + a, m[a]);
    }
}

// ------------------------------------------------------------------ indices masker (K4)
struct MaskParams {
    const int32_t* idx;  // [d][n] global coordinates
    long long n;
    int origin[3];  // global coordinate of local cell (0,0,0), kernel axes
    int gshape[3];  // global domain extents, kernel axes
    int nx, ny, nz;
    int id, mark_solid, pad;
    uint8_t* bc_mask;
    uint8_t* solid;
    const uint8_t* solid_in;
    uint8_t* missing;
    long long mqs;
};

// missing_mask planes -> one word per cell (XlbCollisionParams.missing_bits)
template <int Q>
__global__ void __launch_bounds__(256) pack_missing_kernel(const uint8_t* __restrict__ missing, long long mqs, long long ncell, uint32_t* __restrict__ bits) {
    const long long cell = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (cell >= ncell) return;
    uint32_t m = 0;
#pragma unroll
    for (int l = 0; l < Q; ++l) m |= (missing[l * mqs + cell] ? 1u : 0u) << l;
    bits[cell] = m;
}

// indices_boundary_masker.py:107-128: tag bc ids (optionally on the padded set index + c_l) and solid cells
template <int Q>
__global__ void __launch_bounds__(256) mask_tag_kernel(const __grid_constant__ MaskParams p) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long total = p.pad ? p.n * Q : p.n;
    if (t >= total) return;
    const long long i = p.pad ? t / Q : t;
    const int l = p.pad ? (int)(t % Q) : 0;
    constexpr int A0 = axis0<Q>();
    int g[3] = {0, 0, 0};
#pragma unroll
    for (int a = A0; a < 3; ++a) g[a] = p.idx[(long long)(a - A0) * p.n + i];
    if (p.mark_solid && l == 0 && p.solid) {
        const int x = g[0] - p.origin[0], y = g[1] - p.origin[1], z = g[2] - p.origin[2];
        if (x >= 0 && x < p.nx && y >= 0 && y < p.ny && z >= 0 && z < p.nz) p.solid[((long long)x * p.ny + y) * p.nz + z] = 1;
    }
    if (p.id >= 0 && p.bc_mask) {
        int c[3] = {0, 0, 0};
        if (p.pad) {
            // runtime direction: select through the constexpr table
            static_for<Q>([&](auto mc) {
                XLB_IDX(m, mc);
                if (m == l) {
                    c[0] = Dir<Q, m>::cx;
                    c[1] = Dir<Q, m>::cy;
                    c[2] = Dir<Q, m>::cz;
                }
            });
        }
        const int gx = g[0] + c[0], gy = g[1] + c[1], gz = g[2] + c[2];
        // padded cells outside the global domain live in the pad ring of the reference and are cropped (:137-142)
        if (gx < 0 || gx >= p.gshape[0] || gy < 0 || gy >= p.gshape[1] || gz < 0 || gz >= p.gshape[2]) return;
        const int x = gx - p.origin[0], y = gy - p.origin[1], z = gz - p.origin[2];
        if (x >= 0 && x < p.nx && y >= 0 && y < p.ny && z >= 0 && z < p.nz) p.bc_mask[((long long)x * p.ny + y) * p.nz + z] = (uint8_t)p.id;
    }
}

// indices_boundary_masker.py:79-96,134: the all-True pad ring + solid cells, pulled along c_l
template <int Q>
__global__ void __launch_bounds__(256) mask_missing_kernel(const __grid_constant__ MaskParams p) {
    const long long cell = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long ncell = (long long)p.nx * p.ny * p.nz;
    if (cell >= ncell) return;
    const int z = (int)(cell % p.nz);
    const int y = (int)((cell / p.nz) % p.ny);
    const int x = (int)(cell / ((long long)p.nz * p.ny));
    static_for<Q>([&](auto lc) {
        XLB_IDX(l, lc);
        using D = Dir<Q, l>;
        const int xs = x - D::cx, ys = y - D::cy, zs = z - D::cz;
        const int gx = xs + p.origin[0], gy = ys + p.origin[1], gz = zs + p.origin[2];
        uint8_t m;
        if (gx < 0 || gx >= p.gshape[0] || gy < 0 || gy >= p.gshape[1] || gz < 0 || gz >= p.gshape[2])
            m = 1;  // pull source is outside the global domain
        else if (xs < 0 || xs >= p.nx || ys < 0 || ys >= p.ny || zs < 0 || zs >= p.nz)
            m = 0;  // outside this block but inside the domain: only reachable from ghost planes, value unused
        else
            m = p.solid_in ? p.solid_in[((long long)xs * p.ny + ys) * p.nz + zs] : 0;
        p.missing[l * p.mqs + cell] = m;
    });
}

// ------------------------------------------------------------------ BC cell list (XlbCollisionParams.bc_cells)
struct ListParams {
    const uint8_t* bc_mask;
    const XlbBcEntry* bct;
    long long first, last;  // linear cell range [first, last)
    long long ncell;        // cells of the whole block (bytes of bc_mask)
    int kind_mask;
    long long* cells;
    long long capacity;
    unsigned long long* count;
};

// one thread per 16 cells (one 128-bit load of bc ids); hits are appended with one atomic per warp
__global__ void __launch_bounds__(256) list_bc_cells_kernel(const __grid_constant__ ListParams p) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long c0 = (p.first & ~15ll) + t * 16;  // bc_mask rows are 16-byte aligned (checked on the host)
    uint32_t hit = 0;
    if (c0 < p.last) {
        uint4 w = make_uint4(0, 0, 0, 0);
        if (c0 + 16 <= p.ncell) {
            w = *reinterpret_cast<const uint4*>(p.bc_mask + c0);
        } else {  // ragged end of the block
            uint32_t tail[4] = {0, 0, 0, 0};
            for (int j = 0; c0 + j < p.ncell; ++j) tail[j >> 2] |= (uint32_t)p.bc_mask[c0 + j] << (8 * (j & 3));
            w = make_uint4(tail[0], tail[1], tail[2], tail[3]);
        }
        const uint32_t words[4] = {w.x, w.y, w.z, w.w};
        if (w.x | w.y | w.z | w.w) {
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                const int id = (words[j >> 2] >> (8 * (j & 3))) & 0xff;
                const long long c = c0 + j;
                if (id != 0 && c >= p.first && c < p.last && ((p.kind_mask >> p.bct[id].kind) & 1)) hit |= 1u << j;
            }
        }
    }
    const int n = __popc(hit);
    // warp-aggregated append: exclusive prefix of n over the lanes, one atomicAdd by the last lane
    const int lane = threadIdx.x & 31;
    int incl = n;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += v;
    }
    const int total = __shfl_sync(0xffffffffu, incl, 31);
    if (total == 0) return;
    unsigned long long base = 0;
    if (lane == 31) base = atomicAdd(p.count, (unsigned long long)total);
    base = __shfl_sync(0xffffffffu, base, 31);
    long long slot = (long long)base + (incl - n);
    if (p.cells == nullptr) return;
    while (hit) {
        const int j = __ffs(hit) - 1;
        hit &= hit - 1;
        if (slot < p.capacity) p.cells[slot] = c0 + j;
        ++slot;
    }
}

}  // namespace xlb
