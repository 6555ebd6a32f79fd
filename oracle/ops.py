"""Oracle per-cell operators (TEST INFRASTRUCTURE -- see oracle/__init__.py).

NumPy restatement, whole-array and in the same operation order, of the JAX branches of
Stream, ZeroMoment/FirstMoment/Macroscopic, SecondMoment, QuadraticEquilibrium, BGK and KBC.

dtype handling follows JAX's promotion rules, which differ from NumPy's in one place that
matters here: ``int32 (x) float32 -> float32`` (NumPy would give float64).  Integer lattice
tables are therefore cast to the compute dtype before every contraction.  Python-float
literals are weakly typed in both systems (NumPy >= 2, NEP 50).
"""

import numpy as np

_POLICIES = {
    # precision_policy.py:57-97   name -> (compute, store)
    "FP64FP64": (np.float64, np.float64),
    "FP64FP32": (np.float64, np.float32),
    "FP64FP16": (np.float64, np.float16),
    "FP32FP32": (np.float32, np.float32),
    "FP32FP16": (np.float32, np.float16),
}


def policy_dtypes(policy):
    """precision_policy.py:65-97 -> (compute numpy dtype, store numpy dtype)."""
    c, s = _POLICIES[policy]
    return np.dtype(c), np.dtype(s)


def stream(lat, f):
    """Stream.jax_implementation, stream.py:29-62: f_out[l] = roll(f[l], c_l) (periodic pull)."""
    out = np.empty_like(f)
    axes = tuple(range(lat.d))
    for l in range(lat.q):
        out[l] = np.roll(f[l], tuple(int(v) for v in lat.c[:, l]), axis=axes)
    return out


def zero_moment(f):
    """ZeroMoment.jax_implementation, zero_moment.py:14-17."""
    return np.sum(f, axis=0, keepdims=True)


def first_moment(lat, f, rho):
    """FirstMoment.jax_implementation, first_moment.py:14-18."""
    c = lat.c.astype(f.dtype)
    return np.tensordot(c, f, axes=(-1, 0)) / rho


def macroscopic(lat, f):
    """Macroscopic.jax_implementation, macroscopic.py:21-26."""
    rho = zero_moment(f)
    u = first_moment(lat, f, rho)
    return rho, u


def second_moment(lat, fneq):
    """SecondMoment.jax_implementation, second_moment.py:35-55: Pi = cc^T . fneq."""
    cc = lat.cc.astype(fneq.dtype)
    return np.tensordot(cc, fneq, axes=(0, 0))


def quadratic_equilibrium(lat, rho, u, dtype=None):
    """QuadraticEquilibrium.jax_implementation, quadratic_equilibrium.py:23-30.

    ``dtype`` is the compute dtype of the operator (dtype of ``velocity_set.w``); the result
    dtype follows JAX promotion of (rho, w, u).
    """
    dtype = np.dtype(dtype) if dtype is not None else np.result_type(rho.dtype, u.dtype)
    c = lat.c.astype(u.dtype)  # int32 (x) float -> float dtype of u
    cu = 3.0 * np.tensordot(c, u, axes=(0, 0))
    usqr = 1.5 * np.sum(np.square(u), axis=0, keepdims=True)
    w = lat.w.astype(dtype).reshape((-1,) + (1,) * (u.ndim - 1))
    feq = rho * w * (1.0 + cu * (1.0 + 0.5 * cu) - usqr)
    return feq


def bgk(f, feq, omega):
    """BGK.jax_implementation, bgk.py:27-32."""
    fneq = f - feq
    fout = f - f.dtype.type(omega) * fneq
    return fout


def _decompose_shear_d3q27(lat, fneq):
    """KBC.decompose_shear_d3q27_jax, kbc.py:96-145."""
    Pi = second_moment(lat, fneq)
    Nxz = Pi[0] - Pi[5]
    Nyz = Pi[3] - Pi[5]
    s = np.zeros_like(fneq)
    s[9] = (2.0 * Nxz - Nyz) / 6.0
    s[18] = (2.0 * Nxz - Nyz) / 6.0
    s[3] = (-Nxz + 2.0 * Nyz) / 6.0
    s[6] = (-Nxz + 2.0 * Nyz) / 6.0
    s[1] = (-Nxz - Nyz) / 6.0
    s[2] = (-Nxz - Nyz) / 6.0
    s[12] = Pi[1] / 4.0
    s[24] = Pi[1] / 4.0
    s[21] = -Pi[1] / 4.0
    s[15] = -Pi[1] / 4.0
    s[10] = Pi[2] / 4.0
    s[20] = Pi[2] / 4.0
    s[19] = -Pi[2] / 4.0
    s[11] = -Pi[2] / 4.0
    s[8] = Pi[4] / 4.0
    s[4] = Pi[4] / 4.0
    s[7] = -Pi[4] / 4.0
    s[5] = -Pi[4] / 4.0
    return s


def _decompose_shear_d2q9(lat, fneq):
    """KBC.decompose_shear_d2q9_jax, kbc.py:147-174."""
    Pi = second_moment(lat, fneq)
    N = Pi[0] - Pi[2]
    s = np.zeros_like(fneq)
    s[3] = N
    s[6] = N
    s[2] = -N
    s[1] = -N
    s[8] = Pi[1]
    s[4] = -Pi[1]
    s[5] = -Pi[1]
    s[7] = Pi[1]
    return s


def kbc(lat, f, feq, omega, epsilon=1e-32):
    """KBC.jax_implementation, kbc.py:40-79 (+ compute_entropic_scalar_products :81-94)."""
    cdt = f.dtype.type
    fneq = f - feq
    if lat.d == 2 and lat.q == 9:
        shear = _decompose_shear_d2q9(lat, fneq)
        delta_s = shear / 4.0
    elif lat.d == 3 and lat.q == 27:
        shear = _decompose_shear_d3q27(lat, fneq)
        delta_s = shear
    else:
        raise NotImplementedError("Velocity set not supported: {}".format(lat))  # kbc.py:65-66
    beta = cdt(0.5) * cdt(omega)
    inv_beta = 1.0 / beta
    delta_h = fneq - delta_s
    temp = delta_h / feq
    sp1 = np.sum(temp * delta_s, axis=0)
    sp2 = np.sum(temp * delta_h, axis=0)
    gamma = inv_beta - (2.0 - inv_beta) * sp1 / (epsilon + sp2)
    fout = f - beta * (2.0 * delta_s + gamma[None, ...] * delta_h)
    return fout


def smagorinsky_les_bgk(lat, f, feq, omega, smagorinsky_coef=0.17):
    """SmagorinskyLESBGK.jax_implementation, smagorinsky_les_bgk.py:44-66."""
    dt = f.dtype.type
    fneq = f - feq
    pi_neq = np.tensordot(lat.cc.astype(f.dtype), fneq, axes=(0, 0))
    if lat.d == 3:
        diag = pi_neq[(0, 3, 5), ...]
        offdiag = pi_neq[(1, 2, 4), ...]
    else:
        diag = pi_neq[(0, 2), ...]
        offdiag = pi_neq[(1,), ...]
    strain = np.sum(diag * diag, axis=0) + dt(2.0) * np.sum(offdiag * offdiag, axis=0)
    tau0 = dt(1.0) / dt(omega)
    cs = dt(smagorinsky_coef)
    tau = dt(0.5) * (tau0 + np.sqrt(tau0 * tau0 + dt(36.0) * (cs * cs) * np.sqrt(strain)))
    omega_eff = dt(1.0) / tau
    fout = f - omega_eff[None, ...] * fneq
    return fout


def exact_difference(lat, f_postcollision, feq, rho, u, force_vector, dtype=None):
    """ExactDifference.jax_implementation, exact_difference_force.py:52-76: f += feq(rho, u + F) - feq(rho, u)."""
    force = np.asarray(force_vector)
    delta_u = np.broadcast_to(force.reshape((-1,) + (1,) * (u.ndim - 1)), u.shape)
    feq_force = quadratic_equilibrium(lat, rho, u + delta_u, dtype=dtype)
    return f_postcollision + (feq_force - feq)


def forced_collision(lat, collision, f, feq, omega, force_vector, dtype=None):
    """ForcedCollision.jax_implementation, forced_collision.py:51-57: inner collision, macroscopic of the result,
    then the forcing operator with the PRE-collision equilibrium."""
    fout = collide(lat, collision, f, feq, omega)
    rho, u = macroscopic(lat, fout)
    return exact_difference(lat, fout, feq, rho, u, force_vector, dtype=dtype)


def collide(lat, collision, f, feq, omega, force_vector=None, dtype=None):
    if force_vector is not None:
        return forced_collision(lat, collision, f, feq, omega, force_vector, dtype=dtype)
    if collision == "BGK":
        return bgk(f, feq, omega)
    if collision == "KBC":
        return kbc(lat, f, feq, omega)
    if collision == "SmagorinskyLESBGK":
        return smagorinsky_les_bgk(lat, f, feq, omega)
    raise ValueError(collision)
