"""Radiosity solve on a ray-traced view-factor matrix (last step).

Reference: emissive_losses/emissive_losses.py:5-83 (``radiosity_RTVF``).  The view factors
come from ``tracer_b200.view_factors`` (device traces); this solve is an n x n dense system
with n = number of cavity elements (tens), so it stays on the host like the reference's.

Behaviour kept from the reference, including where it is surprising:
* when ``inc_radiation`` is given, only the flux rows are assembled; rows that carry a
  temperature keep the identity row and a zero right-hand side (:47-52);
* ``T`` is updated in place for the flux rows (:76);
* the returned tuple starts with the assembled system: (AA, bb, J, E, T, q, Q)."""
import numpy as N

SIGMA = 5.6677e-8          # the reference's Stefan-Boltzmann constant (emissive_losses.py:28)


def radiosity_RTVF(VF, areas, eps, T=None, inc_radiation=None, q_net=None):
    """Solves [AA][J] = [bb] for the radiosities J (W/m2) of n grey diffuse elements.

    VF - (n, n) view factors; areas, eps - per element; T - wall temperatures (K), NaN where
    the flux is imposed instead; inc_radiation - imposed flux (W/m2), NaN where the temperature
    is imposed; q_net - net heat removed from the element, subtracted from the right-hand side.

    Returns AA, bb, J, E (black-body emission, W/m2), T (K), q (net flux, W/m2), Q (W)."""
    n = N.shape(VF)[0]
    if len(eps) != len(areas):
        raise AttributeError
    if T is None and inc_radiation is None:
        raise AttributeError
    if inc_radiation is not None and T is not None:
        no_t, no_f = N.isnan(T), N.isnan(inc_radiation)
        if (no_t & no_f).any():
            raise AttributeError('At least one element has no boundary condition for radiosity')
        if (~no_t & ~no_f).any():
            raise AttributeError('At least one element has two boundary condition definitions for radiosity')

    AA = N.eye(n)
    bb = N.zeros(n)
    if inc_radiation is not None:
        rows = ~N.isnan(inc_radiation)
        bb[rows] += inc_radiation[rows]
        AA[rows] -= VF[rows]
    else:
        rows = ~N.isnan(T)
        bb[rows] += (eps * SIGMA)[rows] * T[rows] ** 4.
        AA[rows] -= VF[rows] * (1. - N.vstack(eps[rows]))
    if q_net is not None:
        rows = ~N.isnan(q_net)
        bb[rows] -= q_net[rows]
    if N.isnan(bb).any():
        raise AttributeError('Wrong right hand side')
    if N.isnan(AA).any():
        raise AttributeError('NaN in the radiosity matrix')

    J = N.linalg.solve(AA, bb)

    q = N.zeros(n)
    if T is None:
        T = N.full(n, N.nan)
    for i in range(n):
        if not N.isnan(T[i]):
            Ei = SIGMA * T[i] ** 4.
            if eps[i] != 1.:
                q[i] = eps[i] / (1. - eps[i]) * (Ei - J[i])
            else:
                q[i] = Ei - N.sum(VF[i, :] * J)
        elif not N.isnan(inc_radiation[i]):
            q[i] = bb[i]
            T[i] = (1. / SIGMA * (J[i] + (1. - eps[i]) / eps[i] * q[i])) ** 0.25
    E = SIGMA * T ** 4.
    Q = areas * q
    return AA, bb, J, E, T, q, Q
