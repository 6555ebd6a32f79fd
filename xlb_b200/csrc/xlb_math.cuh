// xlb_math.cuh -- per-cell LBM arithmetic on register arrays f[Q][V] (column J = one cell).
//
// Operation order follows the reference's JAX branches so that results agree to rounding:
//   rho = sum_l f_l                 zero_moment.py:14-17   (index order 0..Q-1)
//   u   = (sum_l c_l f_l) / rho     first_moment.py:14-18
//   feq = rho*w*(1 + cu*(1+0.5cu) - usqr), cu = 3 c.u, usqr = 1.5 u.u   quadratic_equilibrium.py:23-30
//   BGK: f - omega*(f-feq)          bgk.py:27-32
//   KBC: kbc.py:40-174
// Every loop over directions is a compile-time loop (static_for): lattice constants are immediates.
#pragma once
#include "xlb_common.cuh"

namespace xlb {

// compile-time constants of direction L of lattice Q
template <int Q, int L>
struct Dir {
    static constexpr int cx = Lat<Q>::c(L, 0), cy = Lat<Q>::c(L, 1), cz = Lat<Q>::c(L, 2);
    static constexpr int o = opp<Q>(L);
    static constexpr bool main = is_main<Q>(L);
    static constexpr double w = Lat<Q>::w(L);
};

// s (+)= c * v for a compile-time c in {-1,0,1}; `first` tracks whether s holds a value yet
template <int C, typename T>
XLB_D void acc_signed(T& s, bool& first, T v) {
    if (C == 1) {
        s = first ? v : s + v;
        first = false;
    } else if (C == -1) {
        s = first ? -v : s - v;
        first = false;
    }
}

template <typename T, int Q, int V, int J>
XLB_D void moments(const T (&f)[Q][V], T& rho, T (&u)[3]) {
    T r = T(0), sx = T(0), sy = T(0), sz = T(0);
    bool f0 = true, fx = true, fy = true, fz = true;
    static_for<Q>([&](auto lc) {
        XLB_IDX(l, lc);
        using D = Dir<Q, l>;
        const T v = f[l][J];
        acc_signed<1>(r, f0, v);
        acc_signed<D::cx>(sx, fx, v);
        acc_signed<D::cy>(sy, fy, v);
        acc_signed<D::cz>(sz, fz, v);
    });
    rho = r;
    u[0] = (axis0<Q>() > 0) ? T(0) : sx / r;
    u[1] = sy / r;
    u[2] = sz / r;
}

template <typename T>
XLB_D T usqr_of(const T (&u)[3], int a0) {
    // 1.5 * sum_a u_a^2 (a in lattice dimension order)
    T s = (a0 == 0) ? u[0] * u[0] + u[1] * u[1] : u[1] * u[1];
    s += u[2] * u[2];
    return T(1.5) * s;
}

// equilibrium of direction L
template <typename T, int Q, int L>
XLB_D T feq_dir(T rho, const T (&u)[3], T usqr) {
    using D = Dir<Q, L>;
    T s = T(0);
    bool first = true;
    acc_signed<D::cx>(s, first, u[0]);
    acc_signed<D::cy>(s, first, u[1]);
    acc_signed<D::cz>(s, first, u[2]);
    const T cu = T(3.0) * s;
    return rho * T(D::w) * (T(1.0) + cu * (T(1.0) + T(0.5) * cu) - usqr);
}

// pi[k] (+)= cc_Lk * v over the six kernel-axis pairs xx,xy,xz,yy,yz,zz (second_moment.py:35-55)
template <int Q, int L, typename T>
XLB_D void acc_second_moment(T (&pi)[6], T v) {
    using D = Dir<Q, L>;
    constexpr int cc[6] = {D::cx * D::cx, D::cx * D::cy, D::cx * D::cz, D::cy * D::cy, D::cy * D::cz, D::cz * D::cz};
#pragma unroll
    for (int k = 0; k < 6; ++k) {
        if (cc[k] == 1) pi[k] += v;
        if (cc[k] == -1) pi[k] -= v;
    }
}

// KBC shear part delta_s of direction L given the six kernel-axis moments of fneq.
template <typename T, int Q, int L>
XLB_D T kbc_delta_s(const T (&pi)[6]) {
    if (Q == 27) {  // kbc.py:96-145, delta_s = shear
        const T Nxz = pi[0] - pi[5];
        const T Nyz = pi[3] - pi[5];
        switch (L) {
            case 9: case 18: return (T(2.0) * Nxz - Nyz) / T(6.0);
            case 3: case 6: return (-Nxz + T(2.0) * Nyz) / T(6.0);
            case 1: case 2: return (-Nxz - Nyz) / T(6.0);
            case 12: case 24: return pi[1] / T(4.0);
            case 21: case 15: return -pi[1] / T(4.0);
            case 10: case 20: return pi[2] / T(4.0);
            case 19: case 11: return -pi[2] / T(4.0);
            case 8: case 4: return pi[4] / T(4.0);
            case 7: case 5: return -pi[4] / T(4.0);
            default: return T(0);
        }
    } else {  // D2Q9: kbc.py:147-174 then /4 (kbc.py:60-61); user xx,xy,yy = kernel yy,yz,zz
        const T N = pi[3] - pi[5];
        switch (L) {
            case 3: case 6: return N / T(4.0);
            case 1: case 2: return -N / T(4.0);
            case 8: case 7: return pi[4] / T(4.0);
            case 4: case 5: return -pi[4] / T(4.0);
            default: return T(0);
        }
    }
}

// ---------------------------------------------------------------------------- fused-kernel arithmetic
// The fused step evaluates the same formulas as the reference but shares work between a direction L
// and its opposite O (c_O = -c_L, w_O = w_L):
//     feq_L = rw*(1 - usqr) + rw*cu + 0.5*rw*cu^2,   feq_O = rw*(1 - usqr) - rw*cu + 0.5*rw*cu^2,   rw = rho*w_L
// and u = (sum c f) * (1/rho).  This differs from quadratic_equilibrium.py:25-30 / first_moment.py:16-18 only
// in rounding (a few ulp of the compute type) -- the same freedom XLA takes when it fuses and contracts the
// JAX expression -- and is covered by the parity tolerances (1e-6 one step / 1e-5 N steps FP32, 1e-13 / 1e-12 FP64).
XLB_D float fast_div(float a, float b) { return __fdividef(a, b); }  // MUFU.RCP * a, <= 2 ulp
XLB_D double fast_div(double a, double b) { return a / b; }
XLB_D float rcp_of(float x) { return __frcp_rn(x); }
XLB_D double rcp_of(double x) { return 1.0 / x; }

template <typename T, int Q, int V, int J>
XLB_D void moments_fused(const T (&f)[Q][V], T& rho, T (&u)[3]) {
    T r = T(0), sx = T(0), sy = T(0), sz = T(0);
    bool f0 = true, fx = true, fy = true, fz = true;
    static_for<Q>([&](auto lc) {
        XLB_IDX(l, lc);
        using D = Dir<Q, l>;
        const T v = f[l][J];
        acc_signed<1>(r, f0, v);
        acc_signed<D::cx>(sx, fx, v);
        acc_signed<D::cy>(sy, fy, v);
        acc_signed<D::cz>(sz, fz, v);
    });
    rho = r;
    const T inv = rcp_of(r);
    u[0] = (axis0<Q>() > 0) ? T(0) : sx * inv;
    u[1] = sy * inv;
    u[2] = sz * inv;
}

// cu = 3 c_L.u from the pre-scaled velocity u3 = 3u
template <typename T, int Q, int L>
XLB_D T cu_of(const T (&u3)[3]) {
    using D = Dir<Q, L>;
    T s = T(0);
    bool first = true;
    acc_signed<D::cx>(s, first, u3[0]);
    acc_signed<D::cy>(s, first, u3[1]);
    acc_signed<D::cz>(s, first, u3[2]);
    return s;
}

// symmetric (even in c) and antisymmetric parts of feq for the pair (L, opp L); base = 1 - usqr
template <typename T, int Q, int L>
XLB_D void feq_pair(T rho, const T (&u3)[3], T base, T& even, T& odd) {
    const T rw = rho * T(Dir<Q, L>::w);
    const T cu = cu_of<T, Q, L>(u3);
    even = rw * (T(0.5) * cu * cu + base);
    odd = rw * cu;
}

// Collide column J in place.  If fullway is set the COLLISION-step FullwayBounceBackBC result
// f_post[l] = f_post_stream[opp l] (bc_fullway_bounce_back.py:50-56) is selected instead.
template <typename T, int Q, int COLL, int V, int J>
XLB_D void collide_column(T (&f)[Q][V], T omega, bool fullway) {
    T rho, u[3];
    moments_fused<T, Q, V, J>(f, rho, u);
    const T base = T(1.0) - usqr_of<T>(u, axis0<Q>());
    const T u3[3] = {T(3.0) * u[0], T(3.0) * u[1], T(3.0) * u[2]};
    if (COLL == XLB_BGK) {
        static_for<Q>([&](auto lc) {
            XLB_IDX(l, lc);
            constexpr int o = Dir<Q, l>::o;
            if (o == l) {  // rest population: cu = 0
                const T fl = f[l][J];
                const T nl = fl - omega * (fl - rho * T(Dir<Q, l>::w) * base);
                f[l][J] = fullway ? fl : nl;
            } else if (o > l) {
                T even, odd;
                feq_pair<T, Q, l>(rho, u3, base, even, odd);
                const T fl = f[l][J], fo = f[o][J];
                const T nl = fl - omega * (fl - (even + odd));
                const T no = fo - omega * (fo - (even - odd));
                f[l][J] = fullway ? fo : nl;
                f[o][J] = fullway ? fl : no;
            }
        });
    } else {
        // ---- KBC (kbc.py:40-79).  Pi = sum cc fneq only sees the even part of fneq; delta_s is even too.
        T pi[6] = {T(0), T(0), T(0), T(0), T(0), T(0)};
        static_for<Q>([&](auto lc) {
            XLB_IDX(l, lc);
            constexpr int o = Dir<Q, l>::o;
            if (o > l) {
                T even, odd;
                feq_pair<T, Q, l>(rho, u3, base, even, odd);
                acc_second_moment<Q, l>(pi, (f[l][J] + f[o][J]) - T(2.0) * even);
            }
        });
        const T beta = T(0.5) * omega;
        const T inv_beta = T(1.0) / beta;
        T sp1 = T(0), sp2 = T(0);
        static_for<Q>([&](auto lc) {
            XLB_IDX(l, lc);
            constexpr int o = Dir<Q, l>::o;
            if (o == l) {
                const T fe = rho * T(Dir<Q, l>::w) * base;
                const T dh = f[l][J] - fe;  // delta_s of the rest direction is 0
                sp2 += fast_div(dh, fe) * dh;
            } else if (o > l) {
                T even, odd;
                feq_pair<T, Q, l>(rho, u3, base, even, odd);
                const T ds = kbc_delta_s<T, Q, l>(pi);
                const T fel = even + odd, feo = even - odd;
                const T dhl = (f[l][J] - fel) - ds, dho = (f[o][J] - feo) - ds;
                const T tl = fast_div(dhl, fel), to = fast_div(dho, feo);
                sp1 += (tl + to) * ds;
                sp2 += tl * dhl + to * dho;
            }
        });
        const T gamma = inv_beta - (T(2.0) - inv_beta) * fast_div(sp1, T(1e-32) + sp2);
        static_for<Q>([&](auto lc) {
            XLB_IDX(l, lc);
            constexpr int o = Dir<Q, l>::o;
            if (o == l) {
                const T fl = f[l][J];
                const T dh = fl - rho * T(Dir<Q, l>::w) * base;
                const T nl = fl - beta * (gamma * dh);
                f[l][J] = fullway ? fl : nl;
            } else if (o > l) {
                T even, odd;
                feq_pair<T, Q, l>(rho, u3, base, even, odd);
                const T ds = kbc_delta_s<T, Q, l>(pi);
                const T fl = f[l][J], fo = f[o][J];
                const T dhl = (fl - (even + odd)) - ds, dho = (fo - (even - odd)) - ds;
                const T nl = fl - beta * (T(2.0) * ds + gamma * dhl);
                const T no = fo - beta * (T(2.0) * ds + gamma * dho);
                f[l][J] = fullway ? fo : nl;
                f[o][J] = fullway ? fl : no;
            }
        });
    }
}

// ZouHeBC (bc_zouhe.py:165-304) on one cell: f = post-stream populations (column 0), miss = missing
// bits of the cell, e = BC row (prescribed values), velocity type or pressure type.
template <typename T, int Q>
XLB_D void zouhe_apply(T (&f)[Q][1], uint32_t miss, const XlbBcEntry& e, bool velocity_type, int x, int y, int z) {
    int nrm[3] = {0, 0, 0};
    T s_mid = T(0), s_known = T(0);
    static_for<Q>([&](auto lc) {
        XLB_IDX(l, lc);
        using D = Dir<Q, l>;
        const bool m = (miss >> l) & 1u;
        const bool k = (miss >> D::o) & 1u;  // known_mask = missing[opp]            :165-169
        if (D::main && m) {                   // normals = -sum_main c_l missing_l    :171-176
            nrm[0] -= D::cx;
            nrm[1] -= D::cy;
            nrm[2] -= D::cz;
        }
        s_mid += (!(m || k)) ? f[l][0] : T(0);
        s_known += k ? f[l][0] : T(0);
    });
    const T fsum = s_mid + T(2.0) * s_known;  // :248, :263
    T rho, vel[3] = {T(0), T(0), T(0)};
    constexpr int A0 = axis0<Q>();
    const long long pvo = (long long)x * e.pv_stride[1] + (long long)y * e.pv_stride[2] + (long long)z * e.pv_stride[3];
    if (velocity_type) {
        T un = T(0);
#pragma unroll
        for (int a = A0; a < 3; ++a) {
            vel[a] = ld_any<T>(e.pv, e.pv_dtype, pvo + (a - A0) * e.pv_stride[0]);
            un += static_cast<T>(nrm[a]) * vel[a];
        }
        rho = fsum / (T(1.0) + un);  // :256-266
    } else {
        rho = ld_any<T>(e.pv, e.pv_dtype, pvo);
        const T un = T(-1.0) + fsum / rho;  // :241-254
#pragma unroll
        for (int a = A0; a < 3; ++a) vel[a] = un * static_cast<T>(nrm[a]);
    }
    const T usqr = usqr_of<T>(vel, A0);
    T g[Q];
#pragma unroll
    for (int l = 0; l < Q; ++l) g[l] = f[l][0];
    static_for<Q>([&](auto lc) {
        XLB_IDX(l, lc);
        constexpr int o = Dir<Q, l>::o;
        if (miss & (1u << l)) {  // bounceback_nonequilibrium :279-288
            f[l][0] = g[o] + feq_dir<T, Q, l>(rho, vel, usqr) - feq_dir<T, Q, o>(rho, vel, usqr);
        }
    });
}

}  // namespace xlb
