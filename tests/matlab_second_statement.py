"""NumPy transliterations of the MATLAB prototypes the reference ships beside its Fortran -- the only reference-held SECOND
statement of the physics (there is no Fortran compiler here, and the reference has no tests):
  matlab/unused/dens.m + denssub.m     the Ngo diffusive-equilibrium density with plasmapause knee and Gaussian ducts
  matlab/stix_parameters.m, dispersion_relation.m, solve_dispersion_relation.m, bmodel.m
Test infrastructure: tests/test_matlab_second_statement.py checks the oracle's transcription of the Fortran against these
at 1e-6 (the two languages differ in float32 literals, the Earth radius of physconst.m, and branch conventions at exact
boundaries; none of that reaches 1e-6 away from the boundaries)."""
import numpy as np


def denssub(z, rbase, therm, num, rzero, scbot, ane0, alpha0, knee=None, ducts=None):
    """matlab/unused/denssub.m:1-242 (ani only).  knee = (LK, EXPK, DDK, RCONSN, SCR); ducts = list of dicts."""
    d2r = np.pi / 180
    r0 = 6370.0
    cosz2, sinz2 = np.cos(z[1]), np.sin(z[1])
    sinz22 = sinz2 * sinz2
    rb7370 = rbase / 7370
    sh = [0, 0, 1.1506 * therm * rb7370 * rb7370, 0, 0]
    sh[3] = sh[2] / 4
    sh[4] = sh[3] / 4
    gph = rbase * (1.0 - rbase / z[0])
    exnor = [0, 0, np.exp(-gph / sh[2]), 0, 0]
    exnor[3] = exnor[2] ** 4
    exnor[4] = exnor[3] ** 4
    q = 0.0
    qi = [0.0] * 5
    for i in range(2, num + 1):
        qi[i] = alpha0[i] * exnor[i]
        q = q + qi[i]
    alpha = [0.0] * 5
    for i in range(2, num + 1):
        alpha[i] = qi[i] / q
    ani1 = ane0 * np.sqrt(q)
    arg = min((z[0] - rzero) / scbot, 13.0)
    anli = 1.0 - np.exp(-(arg * arg))
    l = z[0] / (r0 * sinz22)
    ani1 = ani1 * anli
    if knee is not None:
        LK, EXPK, DDK, RCONSN, SCR = knee
        deltal = l - LK
        if deltal > 0:
            d2 = DDK ** 2
            argl = min(deltal * deltal / (d2 * 2.0), 80)
            f = np.exp(-argl)
            argr = min((z[0] - RCONSN) / SCR, 12.5)
            fr = np.exp(-argr * argr)
            trm = (RCONSN / z[0]) ** EXPK
            trmodl = trm + (1 - trm) * fr
            ani1 = ani1 * (f + trmodl * (1 - f))
    if ducts:
        latitu = 90.0 - z[1] / d2r
        for D in ducts:
            deltal = l - D["L0"]
            if deltal * D["SIDEDU"] < 0:
                deltal = 0
            argl = (deltal * deltal) / (2 * D["DD"] ** 2)
            delnl = D["DEF"] * np.exp(-argl)
            skip, frduct = False, 0.0
            if (latitu >= 0 and z[0] > D["RDUCUN"]) or (latitu < 0 and z[0] > D["RDUCUS"]):
                delr = z[0] - (D["RDUCUN"] if latitu >= 0 else D["RDUCUS"])
                arglr = delr * delr / ((D["HDUCUN"] if latitu >= 0 else D["HDUCUS"]) ** 2)
                frduct = np.exp(-arglr)
            elif (latitu >= 0 and z[0] < D["RDUCLN"]) or (latitu < 0 and z[0] < D["RDUCLS"]):
                delr = z[0] - (D["RDUCLN"] if latitu >= 0 else D["RDUCLS"])
                arglr = delr * delr / ((D["HDUCLN"] if latitu >= 0 else D["HDUCLS"]) ** 2)
                frduct = np.exp(-arglr)
            else:
                skip = True
            if not skip:
                delnl = delnl * frduct
            ani1 = ani1 * (1 + delnl)
    return [ani1] + [ani1 * alpha[i] for i in range(2, 5)]


def dens(L, lam, dsrrng, dsrlat, dsdens, therm, rbase, rzero, scbot, alpha0, knee=None, ducts=None):
    """matlab/unused/dens.m:63-160: ane0 reset so that ani(1) = dsdens at the set point, then ani at r = r0 L cos^2(lam)."""
    d2r = np.pi / 180
    r0 = 6370.0
    ane0 = 2562.0
    z = [dsrrng * r0, d2r * (90.0 - dsrlat)]
    ani = denssub(z, rbase, therm, 4, rzero, scbot, ane0, alpha0, knee, ducts)
    ane0 = dsdens * (ane0 / ani[0])
    out = np.zeros((len(L), 4))
    for i in range(len(L)):
        r = r0 * L[i] * np.cos(lam[i] * d2r) ** 2
        out[i] = denssub([r, d2r * (90.0 - lam[i])], rbase, therm, 4, rzero, scbot, ane0, alpha0, knee, ducts)
    return out


EPS0 = 8.854187817e-12
MU0 = np.pi * 4e-7
C = np.sqrt(1 / EPS0 / MU0)


def stix_parameters(w, qs, Ns, ms, nus, B0mag):
    """matlab/stix_parameters.m (nus = 0 on the path)"""
    wps2 = (Ns * qs ** 2 / ms / EPS0) * (w / (w + 1j * nus))
    wcs = ((qs * B0mag) / ms) * (w / (w + 1j * nus))
    R = 1 - np.sum(wps2 / (w * (w + wcs)))
    L = 1 - np.sum(wps2 / (w * (w - wcs)))
    P = 1 - np.sum(wps2 / w ** 2)
    return 0.5 * (R + L), 0.5 * (R - L), P, R, L


def dispersion_relation(n, w, qs, Ns, ms, nus, B0):
    """matlab/dispersion_relation.m; returns F and its three terms (for a cancellation-aware comparison)"""
    nmag2 = n @ n
    cos2phi = ((n @ B0) * (n @ B0)) / (nmag2 * (B0 @ B0))
    sin2phi = 1.0 - cos2phi
    S, D, P, R, L = stix_parameters(w, qs, Ns, ms, nus, np.sqrt(B0 @ B0))
    A = S * sin2phi + P * cos2phi
    B = R * L * sin2phi + P * S * (1 + cos2phi)
    return (A * nmag2 ** 2 - B * nmag2 + R * L * P).real, (abs(A * nmag2 ** 2) + abs(B * nmag2) + abs(R * L * P))


def solve_dispersion_relation(n, w, qs, Ns, ms, nus, B0):
    """matlab/solve_dispersion_relation.m (no handedness swap in the prototype)"""
    cos2phi = ((n @ B0) * (n @ B0)) / ((n @ n) * (B0 @ B0))
    sin2phi = 1.0 - cos2phi
    S, D, P, R, L = stix_parameters(w, qs, Ns, ms, nus, np.sqrt(B0 @ B0))
    A = S * sin2phi + P * cos2phi
    B = R * L * sin2phi + P * S * (1 + cos2phi)
    disc = np.sqrt(complex(B ** 2 - 4 * A * R * L * P))
    n1, n2 = np.sqrt((B + disc) / (2 * A)), np.sqrt((B - disc) / (2 * A))
    return w * n1 / C, w * n2 / C


def bmodel(R, theta_rad):
    """matlab/bmodel.m: R in Earth radii, polar angle in radians -> (Brad, Btheta)"""
    Bor3 = (.312 / 10000) * R ** (-3.0)
    return -2 * Bor3 * np.cos(theta_rad), -Bor3 * np.sin(theta_rad)
