// xlb_mesh.cuh -- ray-cast mesh voxelisation (MeshMaskerRay, xlb/operator/boundary_masker/ray.py:43-77, and the shared
// post-pass resolve_out_of_bound_kernel, mesh_boundary_masker.py:157-179).
//
// The reference casts, from every cell centre, one ray per lattice direction through warp.mesh_query_ray (BVH in the
// third-party warp-lang package).  Here the loop is turned inside out: one warp per TRIANGLE walks the cells of the
// triangle's bounding box grown by one cell (no other cell's unit-length rays can reach it) and tests the q-1 rays of each
// cell against that one triangle -- no acceleration structure to build, perfectly parallel over triangles, and every
// write is idempotent (bc id, missing bit) or an atomic minimum (closest hit), so the result does not depend on the
// order in which triangles are processed.  The intersection is the watertight test of Woop, Benthin and Wald (JCGT 2013),
// the algorithm behind mesh_query_ray, in FP32 with the FP64 fallback for vanishing edge functions; all arithmetic uses
// explicit round-to-nearest intrinsics (no FMA contraction) so the CPU oracle (oracle/mesh_masker.py) matches bit for bit.
#pragma once
#include "xlb_math.cuh"

namespace xlb {

struct MeshParams {
    const float* tri;  // [ntri][3 vertices][3 coords], kernel axes, lattice units, GLOBAL coordinates
    long long ntri;
    int nx, ny, nz;  // local block
    int origin[3];   // global coordinate of local cell (0,0,0)
    int gshape[3];   // global domain extents
    int id;
    uint8_t* bc_mask;
    uint8_t* missing;
    long long mqs;
    unsigned long long* dist_key;  // optional [q][ncell]: (bits of t) << 32 | bits of the distance weight, atomicMin
};

// constants of the watertight test for a lattice direction (dir = c / |c|)
template <int Q, int L>
struct RayDir {
    using D = Dir<Q, L>;
    static constexpr int n2 = D::cx * D::cx + D::cy * D::cy + D::cz * D::cz;
    XLB_HD static constexpr int comp(int a) { return a == 0 ? D::cx : (a == 1 ? D::cy : D::cz); }
    XLB_HD static constexpr int iabs(int v) { return v < 0 ? -v : v; }
    // max_dim of |dir|: the first component with the largest magnitude (all non-zero components are equal)
    static constexpr int kz = iabs(comp(0)) == 1 ? 0 : (iabs(comp(1)) == 1 ? 1 : 2);
    static constexpr int kx0 = (kz + 1) % 3, ky0 = (kz + 2) % 3;
    static constexpr int kx = comp(kz) < 0 ? ky0 : kx0;  // swap to preserve the winding when dir[kz] < 0
    static constexpr int ky = comp(kz) < 0 ? kx0 : ky0;
};

template <int Q, int L>
XLB_D bool ray_triangle(const float (&o)[3], const float (&va)[3], const float (&vb)[3], const float (&vc)[3], float& t, float& u, float& v) {
    using R = RayDir<Q, L>;
    const float len = __fsqrt_rn((float)R::n2);
    const float dkz = __fdiv_rn((float)R::comp(R::kz), len);
    const float sx = __fdiv_rn(__fdiv_rn((float)R::comp(R::kx), len), dkz);
    const float sy = __fdiv_rn(__fdiv_rn((float)R::comp(R::ky), len), dkz);
    const float sz = __fdiv_rn(1.0f, dkz);
    float A[3], B[3], C[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        A[a] = __fsub_rn(va[a], o[a]);
        B[a] = __fsub_rn(vb[a], o[a]);
        C[a] = __fsub_rn(vc[a], o[a]);
    }
    const float ax = __fsub_rn(A[R::kx], __fmul_rn(sx, A[R::kz])), ay = __fsub_rn(A[R::ky], __fmul_rn(sy, A[R::kz]));
    const float bx = __fsub_rn(B[R::kx], __fmul_rn(sx, B[R::kz])), by = __fsub_rn(B[R::ky], __fmul_rn(sy, B[R::kz]));
    const float cx = __fsub_rn(C[R::kx], __fmul_rn(sx, C[R::kz])), cy = __fsub_rn(C[R::ky], __fmul_rn(sy, C[R::kz]));
    float U = __fsub_rn(__fmul_rn(cx, by), __fmul_rn(cy, bx));
    float V = __fsub_rn(__fmul_rn(ax, cy), __fmul_rn(ay, cx));
    float W = __fsub_rn(__fmul_rn(bx, ay), __fmul_rn(by, ax));
    if (U == 0.0f || V == 0.0f || W == 0.0f) {  // FP64 fallback of the edge functions (Woop et al., section 3.3)
        U = (float)__dsub_rn(__dmul_rn((double)cx, (double)by), __dmul_rn((double)cy, (double)bx));
        V = (float)__dsub_rn(__dmul_rn((double)ax, (double)cy), __dmul_rn((double)ay, (double)cx));
        W = (float)__dsub_rn(__dmul_rn((double)bx, (double)ay), __dmul_rn((double)by, (double)ax));
    }
    if ((U < 0.0f || V < 0.0f || W < 0.0f) && (U > 0.0f || V > 0.0f || W > 0.0f)) return false;
    const float det = __fadd_rn(__fadd_rn(U, V), W);
    if (det == 0.0f) return false;
    const float az = __fmul_rn(sz, A[R::kz]), bz = __fmul_rn(sz, B[R::kz]), cz = __fmul_rn(sz, C[R::kz]);
    const float T = __fadd_rn(__fadd_rn(__fmul_rn(U, az), __fmul_rn(V, bz)), __fmul_rn(W, cz));
    const float rcp = __fdiv_rn(1.0f, det);
    t = __fmul_rn(T, rcp);
    u = __fmul_rn(U, rcp);
    v = __fmul_rn(V, rcp);
    return t >= 0.0f && t < len;
}

template <int Q>
__global__ void __launch_bounds__(128) mesh_ray_kernel(const __grid_constant__ MeshParams p) {
    const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (warp >= p.ntri) return;
    float va[3], vb[3], vc[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        va[a] = p.tri[warp * 9 + a];
        vb[a] = p.tri[warp * 9 + 3 + a];
        vc[a] = p.tri[warp * 9 + 6 + a];
    }
    // cells (global coordinates) whose rays can reach the triangle: its bounding box grown by one cell, clipped to my block
    int lo[3], ext[3];
    const int n[3] = {p.nx, p.ny, p.nz};
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float mn = fminf(va[a], fminf(vb[a], vc[a])), mx = fmaxf(va[a], fmaxf(vb[a], vc[a]));
        int l0 = (int)floorf(mn) - 1, h0 = (int)ceilf(mx) + 1;  // [l0, h0)
        l0 = max(l0, p.origin[a]);
        h0 = min(h0, p.origin[a] + n[a]);
        lo[a] = l0;
        ext[a] = max(h0 - l0, 0);
    }
    const long long ncand = (long long)ext[0] * ext[1] * ext[2];
    const long long ncell = (long long)p.nx * p.ny * p.nz;
    for (long long k = lane; k < ncand; k += 32) {
        const int cz = (int)(k % ext[2]);
        const int cy = (int)((k / ext[2]) % ext[1]);
        const int cx = (int)(k / ((long long)ext[2] * ext[1]));
        const int g[3] = {lo[0] + cx, lo[1] + cy, lo[2] + cz};
        const float o[3] = {__fadd_rn((float)g[0], 0.5f), __fadd_rn((float)g[1], 0.5f), __fadd_rn((float)g[2], 0.5f)};  // helper_functions_masker.py:47-52
        const long long cell = ((long long)(g[0] - p.origin[0]) * p.ny + (g[1] - p.origin[1])) * p.nz + (g[2] - p.origin[2]);
        bool tagged = false;
        static_for<Q>([&](auto lc) {
            XLB_IDX(l, lc);
            using D = Dir<Q, l>;
            if constexpr (D::o != l) {  // skip the rest direction (ray.py:55-57)
                float t, u, v;
                if (ray_triangle<Q, l>(o, va, vb, vc, t, u, v)) {
                    tagged = true;
                    p.missing[D::o * p.mqs + cell] = 1;  // ray.py:66
                    if (p.dist_key) {
                        // hit position from the barycentrics (wp.mesh_eval_position), distance to the centre, / |c_l| (ray.py:72-75)
                        const float w = __fsub_rn(__fsub_rn(1.0f, u), v);
                        float d2 = 0.0f;
#pragma unroll
                        for (int a = 0; a < 3; ++a) {
                            const float pos = __fadd_rn(__fadd_rn(__fmul_rn(u, va[a]), __fmul_rn(v, vb[a])), __fmul_rn(w, vc[a]));
                            const float dv = __fsub_rn(pos, o[a]);
                            d2 = a == 0 ? __fmul_rn(dv, dv) : __fadd_rn(d2, __fmul_rn(dv, dv));
                        }
                        const float len = __fsqrt_rn((float)RayDir<Q, l>::n2);
                        const float wgt = __fdiv_rn(__fsqrt_rn(d2), len);
                        const unsigned long long key = ((unsigned long long)__float_as_uint(__fadd_rn(t, 0.0f)) << 32) | __float_as_uint(wgt);  // -0 -> +0
                        atomicMin(p.dist_key + (l * ncell + cell), key);
                    }
                }
            }
        });
        if (tagged) p.bc_mask[cell] = (uint8_t)p.id;  // ray.py:65
    }
}

// mesh_boundary_masker.py:157-179: cells tagged with ``id`` also miss every direction that pulls from outside the domain
template <int Q>
__global__ void __launch_bounds__(256) mesh_resolve_oob_kernel(const __grid_constant__ MeshParams p) {
    const long long cell = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long ncell = (long long)p.nx * p.ny * p.nz;
    if (cell >= ncell) return;
    if (p.bc_mask[cell] != (uint8_t)p.id) return;
    const int z = (int)(cell % p.nz) + p.origin[2];
    const int y = (int)((cell / p.nz) % p.ny) + p.origin[1];
    const int x = (int)(cell / ((long long)p.nz * p.ny)) + p.origin[0];
    static_for<Q>([&](auto lc) {
        XLB_IDX(l, lc);
        using D = Dir<Q, l>;
        if constexpr (l != 0) {
            const int px = x - D::cx, py = y - D::cy, pz = z - D::cz;
            if (px < 0 || px >= p.gshape[0] || py < 0 || py >= p.gshape[1] || pz < 0 || pz >= p.gshape[2]) p.missing[l * p.mqs + cell] = 1;
        }
    });
}

}  // namespace xlb
