"""Implicit transport of the class / moment fields -- CPU restatement of what MOC::solveWithFVM (MOC.C:186-198) and
QMOM::solveWithFVM (QMOM.C:201-209) ask OpenFOAM for:

    fvScalarMatrix eqn(fvm::ddt(psi) + fvm::div(phi, psi) == S);  eqn.relax();  eqn.solve();

TEST INFRASTRUCTURE ONLY (tests/, __graft_entry__.smoke(), bench.py's cpu legs).

PARITY UNPINNED: fvMatrix, the ddt / convection schemes and the linear solvers are OpenFOAM's (third-party, absent
from /root/reference).  This file restates OpenFOAM 3.0.x's assembly --

  * EulerDdtScheme::fvmDdt: diag = rDeltaT V, source = rDeltaT psi0 V;
  * gaussConvectionScheme::fvmDiv: lower = -w phi, upper = lower + phi, diag = -(sum of the column's lower / upper)
    (negSumDiag); patches: internalCoeffs = phi_b valueInternalCoeffs, boundaryCoeffs = -phi_b valueBoundaryCoeffs, i.e.
    zero-gradient adds phi_b to the diagonal, fixed-value subtracts phi_b psi_b from the source;
  * `== S`: source += S V;
  * fvMatrix::relax(alpha): D += |internalCoeffs|; D = max(|D|, sum|off-diagonals|); D /= alpha; D -= internalCoeffs;
    source += (D - D0) psi0;
  * lduMatrix::solver::normFactor: sum(|Ax - xbar A1| + |b - xbar A1|) + 1e-20, xbar = average(x); residual =
    sum|b - Ax| / normFactor

-- and solves the matrix DIRECTLY (scipy.sparse): the answer every convergent solver tends to.  The reference's own
MOC.C / QMOM.C reach the same matrix through oracle/ref_shim (tests/test_implicit.py compares the two).
"""
from __future__ import annotations

import numpy as np
import scipy.sparse as sp
import scipy.sparse.linalg as spla


def limited_weights(mesh: dict, scheme: str, phi, psi, psi_b=None):
    """limitedScheme::weights(vf) on the internal faces: limiter CDweights + (1 - limiter) pos(faceFlux), with
    gaussGrad (linear) gradients, NVDTVD::r and the limiter of `Gauss vanLeer` ((r + |r|) / (1 + |r|)) or
    `Gauss limitedLinear 1` (max(min(2 r, 1), 0)).  mesh needs Sf, C, weights."""
    own = np.asarray(mesh["owner"], dtype=np.int64)
    nei = np.asarray(mesh["neighbour"], dtype=np.int64)
    bown = np.asarray(mesh["boundary_owner"], dtype=np.int64)
    V, Sf, C, wl = (np.asarray(mesh[k], dtype=np.float64) for k in ("vcell", "Sf", "C", "weights"))
    ni = own.size
    phi = np.asarray(phi, dtype=np.float64)
    P, N = psi[own], psi[nei]
    ssf = wl * (P - N) + N  # linear face values
    grad = np.zeros((V.size, 3))
    np.add.at(grad, own, Sf[:ni] * ssf[:, None])
    np.subtract.at(grad, nei, Sf[:ni] * ssf[:, None])
    pv = psi[bown] if psi_b is None else np.asarray(psi_b, dtype=np.float64)
    np.add.at(grad, bown, Sf[ni:] * pv[:, None])
    grad /= V[:, None]
    d = C[nei] - C[own]
    up = np.where(phi[:ni] > 0, own, nei)
    gradcf = np.einsum("ij,ij->i", d, grad[up])
    gradf = N - P
    sgn = lambda a: np.where(a >= 0, 1.0, -1.0)
    big = np.abs(gradcf) >= 1000 * np.abs(gradf)
    with np.errstate(divide="ignore", invalid="ignore"):
        r = np.where(big, 2 * 1000 * sgn(gradcf) * sgn(gradf) - 1, 2 * (gradcf / np.where(big, 1.0, gradf)) - 1)
    if scheme in ("Gauss vanLeer", "vanLeer"):
        lim = (r + np.abs(r)) / (1 + np.abs(r))
    elif scheme in ("Gauss limitedLinear 1", "limitedLinear 1"):
        lim = np.maximum(np.minimum(2.0 * r, 1.0), 0.0)
    else:
        raise ValueError(scheme)
    return lim * wl + (1.0 - lim) * (phi[:ni] >= 0)


def assemble(mesh: dict, phi, dt, psi0, *, S=None, psi_b=None, relaxation=0.0, scheme="Gauss upwind"):
    """(A [csr], b) of one field.  mesh: owner, neighbour, boundary_owner, vcell (as tests/meshes.py builds them);
    phi [internal + boundary faces]; psi_b [boundary faces] fixed patch values or None = zero-gradient patches."""
    own = np.asarray(mesh["owner"], dtype=np.int64)
    nei = np.asarray(mesh["neighbour"], dtype=np.int64)
    bown = np.asarray(mesh["boundary_owner"], dtype=np.int64)
    V = np.asarray(mesh["vcell"], dtype=np.float64)
    nc, ni = V.size, own.size
    phi = np.asarray(phi, dtype=np.float64)
    rdt = 1.0 / dt
    diag = rdt * V
    b = rdt * psi0 * V
    if S is not None:
        b = b + S * V
    if scheme in ("Gauss upwind", "upwind") or ni == 0:
        w = (phi[:ni] >= 0).astype(np.float64)  # upwind: pos(faceFlux)
    else:
        w = limited_weights(mesh, scheme, phi, np.asarray(psi0, dtype=np.float64), psi_b)
    lower = -w * phi[:ni]
    upper = lower + phi[:ni]
    diag = diag.copy()
    np.subtract.at(diag, own, lower)
    np.subtract.at(diag, nei, upper)
    ic = np.zeros(nc)
    icmag = np.zeros(nc)
    if psi_b is None:
        np.add.at(ic, bown, phi[ni:])
        np.add.at(icmag, bown, np.abs(phi[ni:]))
    else:
        np.subtract.at(b, bown, phi[ni:] * psi_b)
    if relaxation > 0.0:
        sum_off = np.zeros(nc)
        np.add.at(sum_off, own, np.abs(upper))
        np.add.at(sum_off, nei, np.abs(lower))
        D0 = diag
        D = np.maximum(np.abs(D0 + icmag), sum_off) / relaxation - ic
        b = b + (D - D0) * psi0
        diag = D
    diag = diag + ic  # addBoundaryDiag
    rows = np.concatenate([np.arange(nc), own, nei])
    cols = np.concatenate([np.arange(nc), nei, own])
    vals = np.concatenate([diag, upper, lower])
    A = sp.csr_matrix((vals, (rows, cols)), shape=(nc, nc))
    return A, b


def norm_factor(A, b, x):
    xbar = x.mean()
    pA = xbar * np.asarray(A.sum(axis=1)).ravel()
    return np.abs(A @ x - pA).sum() + np.abs(b - pA).sum() + 1.0e-20


def residual(A, b, x, x0=None):
    """OpenFOAM's normalised residual of x (normalised with the INITIAL field x0's factor, as the solvers do)."""
    return np.abs(b - A @ x).sum() / norm_factor(A, b, x if x0 is None else x0)


def solve(mesh: dict, phi, dt, psi, *, S=None, psi_b=None, relaxation=0.0, scheme="Gauss upwind"):
    """All fields psi [K][ncells] -> the matrices' exact solutions [K][ncells]."""
    psi = np.asarray(psi, dtype=np.float64)
    out = np.empty_like(psi)
    for k in range(psi.shape[0]):
        A, b = assemble(mesh, phi, dt, psi[k], S=None if S is None else S[k], psi_b=None if psi_b is None else psi_b[k],
                        relaxation=relaxation, scheme=scheme)
        out[k] = spla.spsolve(A.tocsc(), b)
    return out


def jacobi_decomposed(part: dict, dt, psi, S, exchange, allreduce, nc_global, tol=1e-13, max_iter=2000):
    """The decomposed solve as the device runs it, restated for ONE rank of a partition (sharding.partition_mesh):
    processor-patch faces are coupled -- w phi_b on the diagonal, (1 - w) phi_b times the value across the face off
    it -- and every Jacobi sweep (x += z, z <- -D^-1 (R z + C z_halo)) starts with a halo exchange of z and ends with
    the global residual sum.  exchange(field [K][nlocal]) -> [K][ncoupled] values across the coupled faces;
    allreduce(array) -> the sum over the ranks.  Zero-gradient physical patches.  Returns (x, iterations, residuals)."""
    own = np.asarray(part["owner"], dtype=np.int64)
    nei = np.asarray(part["neighbour"], dtype=np.int64)
    bown = np.asarray(part["boundary_owner"], dtype=np.int64)
    V = np.asarray(part["vcell"], dtype=np.float64)
    phi = np.asarray(part["phi"], dtype=np.float64)
    ni = own.size
    nphys, ncpl = part["coupled"]
    w = (phi[:ni] >= 0).astype(np.float64)
    lower = -w * phi[:ni]
    upper = lower + phi[:ni]
    diag = V / dt
    diag = diag.copy()
    np.subtract.at(diag, own, lower)
    np.subtract.at(diag, nei, upper)
    np.add.at(diag, bown[:nphys], phi[ni:ni + nphys])
    phb = phi[ni + nphys:ni + nphys + ncpl]
    lowb = -(phb >= 0).astype(np.float64) * phb
    cfb = lowb + phb  # multiplies the value across the face
    cells_b = bown[nphys:nphys + ncpl]
    np.subtract.at(diag, cells_b, lowb)

    def offd(f, halo):  # [K][n] -> R f + C halo
        out = np.zeros_like(f)
        for k in range(f.shape[0]):
            np.add.at(out[k], own, upper * f[k][nei])
            np.add.at(out[k], nei, lower * f[k][own])
            np.add.at(out[k], cells_b, cfb * halo[k])
        return out

    a1 = diag.copy()
    np.add.at(a1, own, upper)
    np.add.at(a1, nei, lower)
    np.add.at(a1, cells_b, cfb)
    x = np.array(psi, dtype=np.float64, copy=True)
    b = x * V / dt + (0.0 if S is None else S * V)
    Ax = diag * x + offd(x, exchange(x))
    xbar = allreduce(x.sum(axis=1)) / nc_global
    pA = xbar[:, None] * a1[None, :]
    norm = allreduce((np.abs(Ax - pA) + np.abs(b - pA)).sum(axis=1)) + 1.0e-20
    r = b - Ax
    res0 = allreduce(np.abs(r).sum(axis=1)) / norm
    z = r / diag
    res, it = res0, 0
    while it < max_iter and np.any(res >= tol):
        halo = exchange(z)
        x += z
        rz = -offd(z, halo)
        res = allreduce(np.abs(rz).sum(axis=1)) / norm
        z = rz / diag
        it += 1
    return x, it, (res0, res)
