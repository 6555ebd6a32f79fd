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
        if (cc[k] == 1) pi[k] = pi[k] + v;
        if (cc[k] == -1) pi[k] = pi[k] - v;
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

// signed index into the distinct delta_s values used by collide_cols: 0 -> delta_s = 0, +k -> ds[k], -k -> -ds[k]
template <int Q, int L>
XLB_HD constexpr int kbc_ds_index() {
    if (Q == 27) {
        switch (L) {
            case 9: case 18: return 1;
            case 3: case 6: return 2;
            case 1: case 2: return 3;
            case 12: case 24: return 4;
            case 21: case 15: return -4;
            case 10: case 20: return 5;
            case 19: case 11: return -5;
            case 8: case 4: return 6;
            case 7: case 5: return -6;
            default: return 0;
        }
    } else if (Q == 9) {
        switch (L) {
            case 3: case 6: return 1;
            case 1: case 2: return -1;
            case 8: case 7: return 2;
            case 4: case 5: return -2;
            default: return 0;
        }
    }
    return 0;
}

// ---------------------------------------------------------------------------- fused-kernel arithmetic
// The fused step evaluates the reference's formulas with work shared between a direction L and its opposite O
// (c_O = -c_L, w_O = w_L):
//     feq_L = even + odd, feq_O = even - odd,  odd = rw*cu,  even = rw*(1 - usqr) + 0.5*odd*cu,  rw = rho*w_L, cu = 3 c_L.u
//     BGK :  f - omega*(f - feq)           = (1-omega)*f + omega*feq
//     KBC :  f - beta*(2 ds + gamma*dh)     = (1 - beta*gamma)*f + beta*gamma*feq - beta*(2-gamma)*ds
// and u = (sum c f)*(1/rho).  This differs from quadratic_equilibrium.py:25-30 / bgk.py:29-32 / kbc.py:70-79 only in
// rounding (a few ulp of the compute type) -- the freedom XLA takes when it fuses the JAX expression -- and is covered
// by the parity tolerances.  The translation units are compiled with -fmad=false and every multiply-add below is an
// explicit fma, so all kernel variants (vector / scalar / generic) produce bit-identical results.
//
// Number types: the arithmetic is written once over N = double, float, or F2 = two FP32 cells packed in a 64-bit
// register pair.  F2 maps onto Blackwell's packed FP32 pipe (SASS FADD2 / FMUL2 / FFMA2, sm_100+), which halves the
// number of issue slots of the FP32 collision.
struct F2 {
    float2 v;
};
XLB_D F2 operator+(F2 a, F2 b) { return {__fadd2_rn(a.v, b.v)}; }
XLB_D F2 operator-(F2 a) { return {make_float2(-a.v.x, -a.v.y)}; }
XLB_D F2 operator-(F2 a, F2 b) { return {__fadd2_rn(a.v, make_float2(-b.v.x, -b.v.y))}; }
XLB_D F2 operator*(F2 a, F2 b) { return {__fmul2_rn(a.v, b.v)}; }
XLB_D F2 fma_(F2 a, F2 b, F2 c) { return {__ffma2_rn(a.v, b.v, c.v)}; }
XLB_D float fma_(float a, float b, float c) { return fmaf(a, b, c); }
XLB_D double fma_(double a, double b, double c) { return fma(a, b, c); }

template <typename N>
struct Num {
    using scalar = N;
    static constexpr int lanes = 1;
};
template <>
struct Num<F2> {
    using scalar = float;
    static constexpr int lanes = 2;
};
template <typename N>
XLB_D N splat(typename Num<N>::scalar s) { return s; }
template <>
XLB_D F2 splat<F2>(float s) { return {make_float2(s, s)}; }

// a / b for b in the normal range (equilibria, entropic sums): one MUFU.RCP + one multiply, <= 2 ulp.  __fdividef would
// add a range-scaling sequence (FSETP + 2 predicated FMUL) per call that is never needed here.
XLB_D float rcp_approx(float b) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(b));
    return r;
}
XLB_D float fast_div(float a, float b) { return a * rcp_approx(b); }
XLB_D double fast_div(double a, double b) { return a / b; }
XLB_D F2 fast_div(F2 a, F2 b) { return a * F2{make_float2(rcp_approx(b.v.x), rcp_approx(b.v.y))}; }
XLB_D float sqrt_of(float x) { return __fsqrt_rn(x); }
XLB_D double sqrt_of(double x) { return sqrt(x); }
XLB_D F2 sqrt_of(F2 x) { return {make_float2(__fsqrt_rn(x.v.x), __fsqrt_rn(x.v.y))}; }
// 1 / x for the densities and relaxation times (x of order one).  XLB_TUNE_DIET_RCP: MUFU.RCP + one Newton step (two FMAs, packed
// for F2; <= 1 ulp, no range check) instead of the IEEE-rounded __frcp_rn sequence (~8 instructions with a slow-path branch per
// lane).  Every kernel variant goes through this function, so they stay bit-identical to each other.
#ifndef XLB_TUNE_DIET_RCP
#define XLB_TUNE_DIET_RCP 1
#endif
// BGK / LES branch experiments (profiles/r2k_sweep_diet.log: both together cost the FP16 kernel 3.7 %, the FP32 one 1.2 %):
// bit 0 = Horner form of the equilibrium pair, bit 1 = pair-sum moments
#ifndef XLB_TUNE_DIET_BGK
#define XLB_TUNE_DIET_BGK 0
#endif
#if XLB_TUNE_DIET_RCP
XLB_D float rcp_of(float x) {
    const float r = rcp_approx(x);
    return fmaf(r, fmaf(-x, r, 1.0f), r);
}
XLB_D F2 rcp_of(F2 x) {
    const F2 r{make_float2(rcp_approx(x.v.x), rcp_approx(x.v.y))};
    const F2 e = fma_(-x, r, F2{make_float2(1.0f, 1.0f)});
    return fma_(r, e, r);
}
#else
XLB_D float rcp_of(float x) { return __frcp_rn(x); }
XLB_D F2 rcp_of(F2 x) { return {make_float2(__frcp_rn(x.v.x), __frcp_rn(x.v.y))}; }
#endif
XLB_D double rcp_of(double x) { return 1.0 / x; }
// lane-wise select: flag[k] ? a : b
XLB_D float pick(const bool* flag, float a, float b) { return flag[0] ? a : b; }
XLB_D double pick(const bool* flag, double a, double b) { return flag[0] ? a : b; }
XLB_D F2 pick(const bool* flag, F2 a, F2 b) { return {make_float2(flag[0] ? a.v.x : b.v.x, flag[1] ? a.v.y : b.v.y)}; }

// columns J .. J+lanes-1 of f[l] as one number, and back
template <typename N, typename T, int Q, int V>
XLB_D N col_get(const T (&f)[Q][V], int l, int J) {
    if constexpr (Num<N>::lanes == 2) {
        return N{make_float2(f[l][J], f[l][J + 1])};
    } else {
        return f[l][J];
    }
}
template <typename N, typename T, int Q, int V>
XLB_D void col_set(T (&f)[Q][V], int l, int J, N x) {
    if constexpr (Num<N>::lanes == 2) {
        f[l][J] = x.v.x;
        f[l][J + 1] = x.v.y;
    } else {
        f[l][J] = x;
    }
}

// swap the populations l and o of column J (one lane of a packed pair)
template <typename T, int Q, int V>
XLB_D void col_swap(T (&f)[Q][V], int l, int o, int J) {
    const T t = f[l][J];
    f[l][J] = f[o][J];
    f[o][J] = t;
}

// FP16-stored populations kept packed: word w[l][J/2] holds cells J (low half) and J+1 (high half) of population l.
// col_get widens a pair to FP32 on access, col_set narrows with round-to-nearest-even (cvt.rn.f16x2.f32, the store
// rounding of precision_policy.py:105-110) -- so a collision may run on this storage only if it sets every element
// exactly once, after its last read (BGK / LES / KBC without the forcing pass).
template <int Q, int W>
struct PackedH2 {
    uint32_t w[Q][W];
};
XLB_D float2 h2_to_f2(uint32_t w) {
    __half2 h;
    *reinterpret_cast<uint32_t*>(&h) = w;
    return __half22float2(h);
}
XLB_D uint32_t f2_to_h2(float a, float b) {
    const __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<const uint32_t*>(&h);
}
template <typename N, int Q, int W>
XLB_D N col_get(const PackedH2<Q, W>& f, int l, int J) {
    static_assert(Num<N>::lanes == 2, "packed fp16 storage is read one pair at a time");
    return N{h2_to_f2(f.w[l][J >> 1])};
}
template <typename N, int Q, int W>
XLB_D void col_set(PackedH2<Q, W>& f, int l, int J, N x) {
    static_assert(Num<N>::lanes == 2, "packed fp16 storage is written one pair at a time");
    f.w[l][J >> 1] = f2_to_h2(x.v.x, x.v.y);
}
template <int Q, int W>
XLB_D void col_swap(PackedH2<Q, W>& f, int l, int o, int J) {
    const uint32_t a = f.w[l][J >> 1], b = f.w[o][J >> 1];
    const uint32_t sel = (J & 1) ? 0x7610u : 0x3254u;  // take the other word's high / low half
    f.w[l][J >> 1] = __byte_perm(a, b, sel);
    f.w[o][J >> 1] = __byte_perm(b, a, sel);
}

// weight class of direction L (0 rest, 1 axis, 2 face diagonal, 3 cube diagonal)
template <int Q, int L>
XLB_HD constexpr int wclass() {
    return (Lat<Q>::c(L, 0) != 0) + (Lat<Q>::c(L, 1) != 0) + (Lat<Q>::c(L, 2) != 0);
}
template <int Q>
XLB_HD constexpr double wclass_weight(int k) {
    for (int l = 0; l < Q; ++l)
        if ((Lat<Q>::c(l, 0) != 0) + (Lat<Q>::c(l, 1) != 0) + (Lat<Q>::c(l, 2) != 0) == k) return Lat<Q>::w(l);
    return 0.0;
}
template <int Q>
XLB_HD constexpr int n_wclass() { return Q == 27 ? 4 : 3; }

// cu = 3 c_L.u from the pre-scaled velocity u3 = 3u
template <typename N, int Q, int L>
XLB_D N cu_of(const N (&u3)[3]) {
    using D = Dir<Q, L>;
    N s = u3[2];
    bool first = true;
    acc_signed<D::cx>(s, first, u3[0]);
    acc_signed<D::cy>(s, first, u3[1]);
    acc_signed<D::cz>(s, first, u3[2]);
    return s;
}

// scaled equilibrium of the pair (L, opp L): even = k*feq_even, odd = k*feq_odd where the scale k is folded into
// the per-weight-class constants a = k*rho*w, ah = 0.5*a, ab = a*(1 - usqr)
template <typename N, int Q, int L>
XLB_D void feq_pair(const N (&u3)[3], const N* a, const N* ab, N& even, N& odd) {
    using S = typename Num<N>::scalar;
    constexpr int k = wclass<Q, L>();
    const N cu = cu_of<N, Q, L>(u3);
    odd = a[k] * cu;
    even = fma_(splat<N>(S(0.5)) * odd, cu, ab[k]);
}

template <typename N, typename S, int Q>
XLB_D void wclass_constants(N rho, N base, N scale, N* a, N* ab) {
    const N sr = scale * rho;
    static_for<n_wclass<Q>()>([&](auto kc) {
        XLB_IDX(k, kc);
        constexpr double w = wclass_weight<Q>(k);  // forced compile-time evaluation
        a[k] = sr * splat<N>(static_cast<S>(w));
        ab[k] = a[k] * base;
    });
}

// Collide columns J .. J+lanes-1 in place.  fullway[k] selects the COLLISION-step FullwayBounceBackBC result
// f_post[l] = f_post_stream[opp l] (bc_fullway_bounce_back.py:50-56) for lane k: the populations are swapped with
// their opposites up front and the relaxation is switched off for that lane (omega = 0, rho = 1, 1/rho = 0 keep
// every intermediate finite), so the fast path carries no per-direction select.
//
// COLL = model | XLB_COLL_FORCED: BGK (bgk.py:27-32), KBC (kbc.py:40-174), SmagorinskyLESBGK (smagorinsky_les_bgk.py:44-66:
// BGK with a per-cell omega from the non-equilibrium stress), optionally followed by the exact-difference body force
// (forced_collision.py:51-57, exact_difference_force.py:52-76).
template <typename T>
struct CollArgs {
    T omega;
    T les_c;      // 36 cs^2
    T force3[3];  // 3 F, kernel axes
};
template <typename T>
XLB_D CollArgs<T> coll_args(const StepParams& p) {
    CollArgs<T> a;
    a.omega = static_cast<T>(p.omega);
    a.les_c = static_cast<T>(p.les_c);
    a.force3[0] = static_cast<T>(p.force3[0]);
    a.force3[1] = static_cast<T>(p.force3[1]);
    a.force3[2] = static_cast<T>(p.force3[2]);
    return a;
}

template <typename N, typename T, int Q, int COLL, int V, int J, typename FS>
XLB_D void collide_cols(FS& f, const CollArgs<T>& ca, const bool* fullway, bool any_fullway) {
    using S = typename Num<N>::scalar;
    constexpr int LN = Num<N>::lanes;
    constexpr int MODEL = XLB_COLL_MODEL(COLL);
    constexpr bool FORCED = (COLL & XLB_COLL_FORCED) != 0;
    const T omega_s = ca.omega;
    if (any_fullway) {
#pragma unroll
        for (int k = 0; k < LN; ++k) {
            if (fullway[k]) {
                static_for<Q>([&](auto lc) {
                    XLB_IDX(l, lc);
                    constexpr int o = Dir<Q, l>::o;
                    if (o > l) col_swap(f, l, o, J + k);
                });
            }
        }
    }
    // ---- moments (zero_moment.py:14-17, first_moment.py:14-18)
    N rho = splat<N>(S(0)), sx = rho, sy = rho, sz = rho;
    bool b0 = true, bx = true, by = true, bz = true;
    N pi[6];  // KBC / LES: sum_l cc_l f_l, accumulated from the pair sums (cc is even in c)
#pragma unroll
    for (int k = 0; k < 6; ++k) pi[k] = splat<N>(S(0));
    if (MODEL != XLB_BGK || (XLB_TUNE_DIET_BGK & 2)) {
        // sums and differences of a direction and its opposite: c is odd, cc is even, so rho and the second moment take the
        // sums, the momentum the differences (XLB_TUNE_DIET: BGK too -- 39 instead of 45 additions on D3Q19)
        static_for<Q>([&](auto lc) {
            XLB_IDX(l, lc);
            using D = Dir<Q, l>;
            constexpr int o = D::o;
            if (o == l) {
                acc_signed<1>(rho, b0, col_get<N>(f, l, J));
            } else if (o > l) {
                const N fl = col_get<N>(f, l, J), fo = col_get<N>(f, o, J);
                const N sum = fl + fo, dif = fl - fo;
                acc_signed<1>(rho, b0, sum);
                acc_signed<D::cx>(sx, bx, dif);
                acc_signed<D::cy>(sy, by, dif);
                acc_signed<D::cz>(sz, bz, dif);
                if (MODEL != XLB_BGK) acc_second_moment<Q, l>(pi, sum);
            }
        });
    } else {
        static_for<Q>([&](auto lc) {
            XLB_IDX(l, lc);
            using D = Dir<Q, l>;
            const N v = col_get<N>(f, l, J);
            acc_signed<1>(rho, b0, v);
            acc_signed<D::cx>(sx, bx, v);
            acc_signed<D::cy>(sy, by, v);
            acc_signed<D::cz>(sz, bz, v);
        });
    }
    N inv = rcp_of(rho);
    N omega = splat<N>(static_cast<S>(omega_s));
    if (any_fullway) {
        rho = pick(fullway, splat<N>(S(1)), rho);
        inv = pick(fullway, splat<N>(S(0)), inv);
        omega = pick(fullway, splat<N>(S(0)), omega);
    }
    N u3[3];
    const N three_inv = splat<N>(S(3)) * inv;
    u3[0] = (axis0<Q>() > 0) ? splat<N>(S(0)) : sx * three_inv;
    u3[1] = sy * three_inv;
    u3[2] = sz * three_inv;
    // base = 1 - usqr = 1 - 1.5 u.u = 1 - (u3.u3)/6
    N uu = u3[2] * u3[2];
    uu = fma_(u3[1], u3[1], uu);
    if (axis0<Q>() == 0) uu = fma_(u3[0], u3[0], uu);
    const N base = fma_(splat<N>(S(-1.0 / 6.0)), uu, splat<N>(S(1)));
    N a[4], ab[4];
    if (MODEL != XLB_KBC) {
        if (MODEL == XLB_SMAGORINSKY_LES_BGK) {
            // Pi_neq = sum cc (f - feq) = sum cc f - rho (u u + I/3)  (exact second moment of the quadratic equilibrium);
            // strain = sum diag^2 + 2 sum offdiag^2;  tau = (tau0 + sqrt(tau0^2 + 36 cs^2 sqrt(strain))) / 2;  omega = 1/tau
            const N r9 = rho * splat<N>(S(1.0 / 9.0));
            const N third = rho * splat<N>(S(1.0 / 3.0));
            const N ry = r9 * u3[1], rz = r9 * u3[2];
            const N pyy = fma_(-ry, u3[1], pi[3]) - third, pyz = fma_(-ry, u3[2], pi[4]), pzz = fma_(-rz, u3[2], pi[5]) - third;
            N dsum = fma_(pzz, pzz, pyy * pyy), osum = pyz * pyz;
            if (axis0<Q>() == 0) {
                const N rx = r9 * u3[0];
                const N pxx = fma_(-rx, u3[0], pi[0]) - third, pxy = fma_(-rx, u3[1], pi[1]), pxz = fma_(-rx, u3[2], pi[2]);
                dsum = fma_(pxx, pxx, dsum);
                osum = fma_(pxy, pxy, osum);
                osum = fma_(pxz, pxz, osum);
            }
            const N strain = fma_(splat<N>(S(2)), osum, dsum);
            const S tau0 = S(1) / static_cast<S>(omega_s);
            const N tau = splat<N>(S(0.5)) * (splat<N>(tau0) + sqrt_of(fma_(splat<N>(static_cast<S>(ca.les_c)), sqrt_of(strain), splat<N>(tau0 * tau0))));
            omega = rcp_of(tau);
            if (any_fullway) omega = pick(fullway, splat<N>(S(0)), omega);
        }
        // out = (1-omega) f + omega feq, with omega folded into the class constants
        wclass_constants<N, S, Q>(rho, base, omega, a, ab);
        const N keep = splat<N>(S(1)) - omega;
#if XLB_TUNE_DIET_BGK & 1
        N ah[4];  // 0.5 a: feq_l = ab + a cu + ah cu^2 in Horner form, four FMAs per direction pair instead of five operations
#pragma unroll
        for (int k = 0; k < n_wclass<Q>(); ++k) ah[k] = splat<N>(S(0.5)) * a[k];
#endif
        static_for<Q>([&](auto lc) {
            XLB_IDX(l, lc);
            constexpr int o = Dir<Q, l>::o;
            if (o == l) {
                col_set<N>(f, l, J, fma_(keep, col_get<N>(f, l, J), ab[0]));
            } else if (o > l) {
#if XLB_TUNE_DIET_BGK & 1
                constexpr int k = wclass<Q, l>();
                const N cu = cu_of<N, Q, l>(u3);
                const N fel = fma_(cu, fma_(ah[k], cu, a[k]), ab[k]);    // w rho omega (1 - usqr + cu + cu^2 / 2)
                const N feo = fma_(cu, fma_(ah[k], cu, -a[k]), ab[k]);   // the same with -cu
                col_set<N>(f, l, J, fma_(keep, col_get<N>(f, l, J), fel));
                col_set<N>(f, o, J, fma_(keep, col_get<N>(f, o, J), feo));
#else
                N even, odd;
                feq_pair<N, Q, l>(u3, a, ab, even, odd);
                col_set<N>(f, l, J, fma_(keep, col_get<N>(f, l, J), even + odd));
                col_set<N>(f, o, J, fma_(keep, col_get<N>(f, o, J), even - odd));
#endif
            }
        });
    } else {
        // ---- KBC (kbc.py:40-79).  Pi_neq = sum cc (f - feq) = sum cc f - rho (u u + I/3): the second moment of the
        // quadratic equilibrium is exact on D3Q27 / D2Q9, so no pass over feq is needed for it.
        wclass_constants<N, S, Q>(rho, base, splat<N>(S(1)), a, ab);
        {
            const N r9 = rho * splat<N>(S(1.0 / 9.0));  // u_a u_b = u3_a u3_b / 9
            const N rx = r9 * u3[0], ry = r9 * u3[1], rz = r9 * u3[2];
            pi[0] = fma_(-rx, u3[0], pi[0]);  // the isotropic part rho/3 drops out: only differences of the diagonal are used
            pi[1] = fma_(-rx, u3[1], pi[1]);
            pi[2] = fma_(-rx, u3[2], pi[2]);
            pi[3] = fma_(-ry, u3[1], pi[3]);
            pi[4] = fma_(-ry, u3[2], pi[4]);
            pi[5] = fma_(-rz, u3[2], pi[5]);
        }
        // the distinct values of delta_s (kbc.py:96-174); index 0 = directions with delta_s = 0
        N ds[7];
        ds[0] = splat<N>(S(0));
        if (Q == 27) {
            const N Nxz = pi[0] - pi[5], Nyz = pi[3] - pi[5];
            const N sixth = splat<N>(S(1.0 / 6.0)), quarter = splat<N>(S(0.25));
            ds[1] = (splat<N>(S(2)) * Nxz - Nyz) * sixth;   // l = 9, 18   (+-x)
            ds[2] = (splat<N>(S(2)) * Nyz - Nxz) * sixth;   // l = 3, 6    (+-y)
            ds[3] = -(Nxz + Nyz) * sixth;                    // l = 1, 2    (+-z)
            ds[4] = pi[1] * quarter;                         // xy: + for (12, 24), - for (21, 15)
            ds[5] = pi[2] * quarter;                         // xz: + for (10, 20), - for (19, 11)
            ds[6] = pi[4] * quarter;                         // yz: + for (8, 4),   - for (7, 5)
        } else {  // D2Q9 (kernel axes y, z): delta_s = shear / 4
            const N quarter = splat<N>(S(0.25));
            ds[1] = (pi[3] - pi[5]) * quarter;               // l = 3, 6 (+), l = 1, 2 (-)
            ds[2] = pi[4] * quarter;                         // l = 8, 7 (+), l = 4, 5 (-)
            ds[3] = ds[4] = ds[5] = ds[6] = splat<N>(S(0));
        }
        auto ds_of = [&](auto lc) -> N {
            XLB_IDX(l, lc);
            constexpr int idx = kbc_ds_index<Q, l>();
            if constexpr (idx == 0) {
                return ds[0];
            } else if constexpr (idx > 0) {
                return ds[idx];
            } else {
                return -ds[-idx];
            }
        };
        const N beta = splat<N>(S(0.5)) * omega;
        const N inv_beta = rcp_of(splat<N>(S(0.5)) * splat<N>(static_cast<S>(omega_s)));
        N sp1 = splat<N>(S(0)), sp2 = splat<N>(S(0));
        static_for<Q>([&](auto lc) {
            XLB_IDX(l, lc);
            constexpr int o = Dir<Q, l>::o;
            if (o == l) {
                const N fe = ab[0];
                const N dh = col_get<N>(f, l, J) - fe;  // delta_s of the rest direction is 0
                sp2 = fma_(fast_div(dh, fe), dh, sp2);
            } else if (o > l) {
                N even, odd;
                feq_pair<N, Q, l>(u3, a, ab, even, odd);
                const N fel = even + odd, feo = even - odd;
                N dhl = col_get<N>(f, l, J) - fel, dho = col_get<N>(f, o, J) - feo;
                if constexpr (kbc_ds_index<Q, l>() != 0) {
                    const N d = ds_of(lc);
                    dhl = dhl - d;
                    dho = dho - d;
                    const N tl = fast_div(dhl, fel), to = fast_div(dho, feo);
                    sp1 = fma_(tl + to, d, sp1);
                    sp2 = fma_(tl, dhl, sp2);
                    sp2 = fma_(to, dho, sp2);
                } else {
                    sp2 = fma_(fast_div(dhl, fel), dhl, sp2);
                    sp2 = fma_(fast_div(dho, feo), dho, sp2);
                }
            }
        });
        const N gamma = inv_beta - (splat<N>(S(2)) - inv_beta) * fast_div(sp1, splat<N>(S(1e-32)) + sp2);
        // out = A f + B feq - C ds,  A = 1 - beta gamma, B = beta gamma, C = beta (2 - gamma)
        const N B = beta * gamma;
        const N A = splat<N>(S(1)) - B;
        const N C = beta * (splat<N>(S(2)) - gamma);
#pragma unroll
        for (int k = 0; k < n_wclass<Q>(); ++k) {  // fold B into the class constants
            a[k] = B * a[k];
            ab[k] = B * ab[k];
        }
        static_for<Q>([&](auto lc) {
            XLB_IDX(l, lc);
            constexpr int o = Dir<Q, l>::o;
            if (o == l) {
                col_set<N>(f, l, J, fma_(A, col_get<N>(f, l, J), ab[0]));
            } else if (o > l) {
                N even, odd;
                feq_pair<N, Q, l>(u3, a, ab, even, odd);
                constexpr int idx = kbc_ds_index<Q, l>();
                if constexpr (idx > 0) even = fma_(-C, ds[idx], even);
                if constexpr (idx < 0) even = fma_(C, ds[-idx], even);
                col_set<N>(f, l, J, fma_(A, col_get<N>(f, l, J), even + odd));
                col_set<N>(f, o, J, fma_(A, col_get<N>(f, o, J), even - odd));
            }
        });
    }
    if constexpr (FORCED) {
        // ForcedCollision (forced_collision.py:51-57): rho', u' = moments of the post-collision populations, then
        // ExactDifference (exact_difference_force.py:72-75): f += feq(rho', u' + F) - feq(rho, u) with the PRE-collision
        // equilibrium.  Full-way lanes get rho' = rho = 1, u' + F = u = 0, so their difference is exactly zero.
        N rho2 = splat<N>(S(0)), tx = rho2, ty = rho2, tz = rho2;
        bool c0 = true, cx = true, cy = true, cz = true;
        static_for<Q>([&](auto lc) {
            XLB_IDX(l, lc);
            using D = Dir<Q, l>;
            constexpr int o = D::o;
            if (o == l) {
                acc_signed<1>(rho2, c0, col_get<N>(f, l, J));
            } else if (o > l) {
                const N fl = col_get<N>(f, l, J), fo = col_get<N>(f, o, J);
                const N dif = fl - fo;
                acc_signed<1>(rho2, c0, fl + fo);
                acc_signed<D::cx>(tx, cx, dif);
                acc_signed<D::cy>(ty, cy, dif);
                acc_signed<D::cz>(tz, cz, dif);
            }
        });
        N inv2 = rcp_of(rho2);
        N F3[3] = {splat<N>(static_cast<S>(ca.force3[0])), splat<N>(static_cast<S>(ca.force3[1])), splat<N>(static_cast<S>(ca.force3[2]))};
        if (any_fullway) {
            rho2 = pick(fullway, splat<N>(S(1)), rho2);
            inv2 = pick(fullway, splat<N>(S(0)), inv2);
#pragma unroll
            for (int k = 0; k < 3; ++k) F3[k] = pick(fullway, splat<N>(S(0)), F3[k]);
        }
        const N three_inv2 = splat<N>(S(3)) * inv2;
        N v3[3];
        v3[0] = (axis0<Q>() > 0) ? splat<N>(S(0)) : fma_(tx, three_inv2, F3[0]);
        v3[1] = fma_(ty, three_inv2, F3[1]);
        v3[2] = fma_(tz, three_inv2, F3[2]);
        N vv = v3[2] * v3[2];
        vv = fma_(v3[1], v3[1], vv);
        if (axis0<Q>() == 0) vv = fma_(v3[0], v3[0], vv);
        const N base2 = fma_(splat<N>(S(-1.0 / 6.0)), vv, splat<N>(S(1)));
        N a1[4], ab1[4], a2[4], ab2[4];
        wclass_constants<N, S, Q>(rho, base, splat<N>(S(1)), a1, ab1);
        wclass_constants<N, S, Q>(rho2, base2, splat<N>(S(1)), a2, ab2);
        static_for<Q>([&](auto lc) {
            XLB_IDX(l, lc);
            constexpr int o = Dir<Q, l>::o;
            if (o == l) {
                col_set<N>(f, l, J, col_get<N>(f, l, J) + (ab2[0] - ab1[0]));
            } else if (o > l) {
                N e1, o1, e2, o2;
                feq_pair<N, Q, l>(u3, a1, ab1, e1, o1);
                feq_pair<N, Q, l>(v3, a2, ab2, e2, o2);
                const N de = e2 - e1, dd = o2 - o1;
                col_set<N>(f, l, J, col_get<N>(f, l, J) + (de + dd));
                col_set<N>(f, o, J, col_get<N>(f, o, J) + (de - dd));
            }
        });
    }
}

// scalar entry point used by the per-cell generic engine
template <typename T, int Q, int COLL, int V, int J>
XLB_D void collide_column(T (&f)[Q][V], const CollArgs<T>& ca, bool fullway) {
    const bool flags[1] = {fullway};
    collide_cols<T, T, Q, COLL, V, J>(f, ca, flags, fullway);
}

// ZouHeBC (bc_zouhe.py:165-304) on one cell: f = post-stream populations (column 0), miss = missing
// bits of the cell, e = BC row (prescribed values), velocity type or pressure type.
template <typename T, int Q>
XLB_D void zouhe_apply(T (&f)[Q][1], uint32_t miss, const XlbBcEntry& e, bool velocity_type, int x, int y, int z, bool regularize = false) {
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
        if (e.pv_mode == 1) {  // per-index profile of the Warp convention: one inflow speed per cell, u = -value * normal (:332-337)
            const T speed = ld_any<T>(e.pv, e.pv_dtype, pvo);
#pragma unroll
            for (int a = A0; a < 3; ++a) {
                vel[a] = -speed * static_cast<T>(nrm[a]);
                un += static_cast<T>(nrm[a]) * vel[a];
            }
        } else {
#pragma unroll
            for (int a = A0; a < 3; ++a) {
                vel[a] = ld_any<T>(e.pv, e.pv_dtype, pvo + (a - A0) * e.pv_stride[0]);
                un += static_cast<T>(nrm[a]) * vel[a];
            }
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
    if (regularize) {
        // RegularizedBC.regularize_fpop (bc_regularized.py:78-114): f = feq + 4.5 w_l (Q_l : Pi_neq),
        // Pi_neq = sum_l c c (f - feq) (second_moment.py:35-55), Q_l = c c - I/3 with off-diagonals doubled
        // (velocity_set.py:139-153)
        T pi[6] = {T(0), T(0), T(0), T(0), T(0), T(0)};
        static_for<Q>([&](auto lc) {
            XLB_IDX(l, lc);
            acc_second_moment<Q, l>(pi, f[l][0] - feq_dir<T, Q, l>(rho, vel, usqr));
        });
        static_for<Q>([&](auto lc) {
            XLB_IDX(l, lc);
            using D = Dir<Q, l>;
            constexpr int cc[6] = {D::cx * D::cx, D::cx * D::cy, D::cx * D::cz, D::cy * D::cy, D::cy * D::cz, D::cz * D::cz};
            constexpr bool diag[6] = {true, false, false, true, false, true};
            T qp = T(0);
#pragma unroll
            for (int k = 0; k < 6; ++k) {
                // components that involve the unused kernel axis of a 2-D lattice vanish (pi = 0 there, cc = 0)
                if (A0 > 0 && (k == 0 || k == 1 || k == 2)) continue;
                const T qi = diag[k] ? T(cc[k] - 1.0 / 3.0) : T(2.0 * cc[k]);
                qp += qi * pi[k];
            }
            f[l][0] = feq_dir<T, Q, l>(rho, vel, usqr) + T(4.5) * T(D::w) * qp;
        });
    }
}

}  // namespace xlb
