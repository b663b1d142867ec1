"""CPU restatement of PBEMethods::MOM's per-cell path (methodOfMoments), SURVEY 8 row f3.

TEST INFRASTRUCTURE ONLY -- imported by tests/ (never by the product).

PARITY UNPINNED: the reference holds no test for MOM (`MOMTests.C` is orphaned) and its
third-party arithmetic is absent here (GSL `gsl_integration_qagiu`, Boost.Math
`gamma_distribution` / `pdf`).  Stand-ins, both restating published algorithms:
  * integrate() (src/PBESystem/Integrator.H:12-27: QAGIU, epsabs 1e-7, epsrel 0, limit 1000)
    -> scipy.integrate.quad on a semi-infinite range = QUADPACK QAGI, the routine GSL's QAGIU
    re-implements (same x = a + (1-t)/t transform, 15-point Gauss-Kronrod, epsilon extrapolation);
  * boost::math::pdf(gamma_distribution(k, theta), x) -> x^(k-1) e^(-x/theta) / (Gamma(k) theta^k).

Two flavours:
  mom_source_cell        reference-shaped: nested adaptive integrals, one per (term, moment index),
                         exactly the lambdas of MOM.C:259-311 and MOM.C:313-362;
  mom_source_cell_fixed  the fixed-depth rule the device kernel uses (same transform and 15-point
                         rule on a fixed, graded panel set; coalescence folded over unordered pairs,
                         breakup with the order of integration exchanged), vectorised in NumPy --
                         pins the device arithmetic tightly, while the first flavour pins the
                         rule's (and the two identities') accuracy against the reference's shape.
"""
from __future__ import annotations

import math
import warnings

import numpy as np

from . import numpy_restatement as npo

SMALL = 1.0e-15  # OpenFOAM SMALL, double precision

_XGK = np.array([0.991455371120812639206854697526329, 0.949107912342758524526189684047851,
                 0.864864423359769072789712788640926, 0.741531185599394439863864773280788,
                 0.586087235467691130294144838258730, 0.405845151377397166906606412076961,
                 0.207784955007898467600689403773245])
_WGK = np.array([0.022935322010529224963732008058970, 0.063092092629978553290700663189204,
                 0.104790010322250183839876322541518, 0.140653259715525918745189590510238,
                 0.169004726639267902826583426598550, 0.190350578064785409913256402421014,
                 0.204432940075298892414161999234649, 0.209482141084727828012999174891714])
GK15_X = np.concatenate([-_XGK, [0.0], _XGK[::-1]])
GK15_W = np.concatenate([_WGK[:7], [_WGK[7]], _WGK[:7][::-1]])


def integrate(f, lower):
    """Integrator.H:12-27."""
    from scipy.integrate import quad
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        return quad(f, lower, math.inf, epsabs=1e-7, epsrel=0.0, limit=1000)[0]


def gamma_params(m0, m1, m2):
    """MOM.C:151-161 -> (c0, k, theta)."""
    k = m1 ** 2 / (m2 * m0 - m1 ** 2 + SMALL)
    theta = (m0 * m2 - m1 ** 2) / max(m0 * m1, SMALL)
    return m0, k, theta


def gamma_pdf(k, theta, x):
    """boost::math::pdf(gamma_distribution<>(k, theta), x) for x > 0 (works on arrays)."""
    if not (k > 0 and theta > 0):
        return np.full(np.shape(x), np.nan) if np.ndim(x) else math.nan
    return np.exp((k - 1) * np.log(x) - x / theta - math.lgamma(k) - k * math.log(theta))


def _S(spec, xa, xb, eps, rho, nu):
    return npo._coal(spec, np.asarray(xa, float), np.asarray(xb, float), np.asarray(eps, float), rho, nu)


def _g(spec, xi, eps, rho):
    xi = np.asarray(xi, float)
    return npo._break(spec, xi, rho, np.asarray(eps, float) * np.ones_like(xi))


def mom_source_cell(spec, m, eps, rho, nu):
    """momentSourceTerm(k), k = 0..2, of one cell: MOM.C:244-254 = coalescence (MOM.C:259-311)
    + breakup (MOM.C:313-362), nested adaptive integrals as in the reference."""
    m0, k, theta = gamma_params(*m)
    n = lambda x: m0 * float(gamma_pdf(k, theta, x))  # noqa: E731
    out = np.zeros(3)
    for mi in range(3):
        src = 0.0
        if spec["coalescence"] != "noCoalescence":
            death = integrate(lambda xi: xi ** mi * n(xi) * integrate(
                lambda xp: float(_S(spec, xi, xp, eps, rho, nu)) * n(xp), 0.0), 0.0)
            birth = integrate(lambda xi: n(xi) * integrate(
                lambda xp: float(_S(spec, xi, xp, eps, rho, nu)) * (xp + xi) ** mi * n(xp), 0.0), 0.0)
            src += birth / 2.0 - death
        if spec["breakup"] != "noBreakup":
            def integrand(xi):
                death = float(_g(spec, xi, eps, rho)) * n(xi)
                birth = integrate(lambda xp: float(npo.daughter_beta(xi, xp))
                                  * float(_g(spec, xp, eps, rho)) * n(xp), xi)
                return xi ** mi * (birth - death)
            src += integrate(integrand, 0.0)
        out[mi] = src
    return out


GRADED_LEVELS = 15  # PBE_MOM_GRADED_LEVELS
TAIL_LEVELS = 5     # PBE_MOM_TAIL_LEVELS
ORIGIN_POWER = 8    # PBE_MOM_ORIGIN_POWER


def fixed_rule(panels):
    """Nodes x = u/(1-u) and weights w/(1-u)^2 (u = 1 - t) of the device kernel's rule
    (openpbe_b200/csrc/pbe_mom.cu mom_node): the 15-point rule on [h, 1-h) in panels - 2 equal
    panels (h = 1/panels), on [1-h, 1) bisected TAIL_LEVELS times towards u = 1, on [a, h) bisected
    GRADED_LEVELS times towards u = 0 (a = h 2^-GRADED_LEVELS), and on [0, a) under u = a s^ORIGIN_POWER."""
    h = 1.0 / panels
    lo = [(j + 1) * h for j in range(panels - 2)]
    hi = [(j + 2) * h for j in range(panels - 2)]
    for l in range(TAIL_LEVELS + 1):
        lo.append(1.0 - math.ldexp(h, -l))
        hi.append(1.0 - math.ldexp(h, -(l + 1)) if l < TAIL_LEVELS else 1.0)
    for l in range(GRADED_LEVELS):
        hi.append(math.ldexp(h, -l))
        lo.append(math.ldexp(h, -(l + 1)))
    lo, hi = np.array(lo), np.array(hi)
    uc, uh = 0.5 * (lo + hi), 0.5 * (hi - lo)
    u = (uc[:, None] + uh[:, None] * GK15_X[None, :]).ravel()
    wu = (GK15_W[None, :] * uh[:, None]).ravel()
    a = math.ldexp(h, -GRADED_LEVELS)
    s = 0.5 + 0.5 * GK15_X
    u = np.concatenate([u, a * s ** ORIGIN_POWER])
    wu = np.concatenate([wu, 0.5 * GK15_W * a * ORIGIN_POWER * s ** (ORIGIN_POWER - 1)])
    t = 1.0 - u
    return u / t, wu / (t * t)


def mom_source_cell_fixed(spec, m, eps, rho, nu, panels=12):
    m0, k, theta = gamma_params(*m)
    x, w = fixed_rule(panels)
    f = w * m0 * gamma_pdf(k, theta, x)
    out = np.zeros(3)
    if spec["coalescence"] != "noCoalescence":
        # symmetric kernels: birth / 2 - death folds to (-1/2, 0, xa xb) per pair
        tt = f[:, None] * f[None, :] * _S(spec, x[:, None], x[None, :], eps, rho, nu)
        out[0] += -0.5 * np.sum(tt)
        out[2] += np.sum(tt * x[:, None] * x[None, :])
    if spec["breakup"] != "noBreakup":
        # order of integration exchanged: int moment(xi', k) g n dxi' (uniformBinaryBreakup.C:62-69)
        gn = _g(spec, x, eps, rho) * f
        for mi in range(3):
            out[mi] += np.sum((npo.daughter_moment(x, mi) - x ** mi) * gn)
    return out


def mom_step_cell(S, m, dt):
    """MOM.C:198-233 with zero flux: scrub, m += dt S, clamp, d."""
    S = np.where(np.isnan(S), 0.0, S)
    mn = np.maximum(np.asarray(m, float) + dt * S, SMALL)
    return S, mn, (6.0 / math.pi * mn[1] / mn[0]) ** (1.0 / 3.0)


def closed_form_constant_binary(m, c):
    """Exact sources of the gamma closure for constantCoalescence c + binaryBreakup (g = xi^2) +
    uniformBinaryBreakup: coalescence (-c m0^2 / 2, 0, c m1^2); breakup (2/(k+1) - 1) m_{k+2}."""
    m0, k, theta = gamma_params(*m)
    m1 = m0 * k * theta
    m2 = m1 * (k + 1) * theta
    m3 = m2 * (k + 2) * theta
    m4 = m3 * (k + 3) * theta
    coal = np.array([-0.5 * c * m0 * m0, 0.0, c * m1 * m1])
    br = np.array([m2, 0.0, -m4 / 3.0])
    return coal, br
