// Optics callables on device: given the incoming direction, the surface normal
// at the hit and up to three random draws, produce the outgoing direction and
// energy.  Mirrors the semantics (not the structure) of optics.reflections
// (optics.py:145-157) and the optics classes of optics_callables.py:93-627.
#pragma once
#include "geom.cuh"

namespace trc {

// optics.reflections: d - 2 (d.n) n
__device__ __forceinline__ V3 reflect(V3 d, V3 n) {
    double k = dot(d, n);
    return mk(d.x - 2.0 * (k * n.x), d.y - 2.0 * (k * n.y), d.z - 2.0 * (k * n.z));
}

// numpy.round(x, 14) as used by general_axis_rotation (spatial_geometry.py:18).  numpy divides by 1e14; the
// multiplication by 1e-14 differs from that by at most one unit in the last place (1e-16 relative), far
// inside the 1e-9 parity bar, and saves a float64 division.
__device__ __forceinline__ double round14(double x) { return rint(x * 1e14) * 1e-14; }

// rotate_z_to_normal (ray_trace_utils/vector_manipulations.py:56-74): the
// minimal rotation taking +z to `n`, applied to `e`.  Axis = z x n normalised
// (x-hat when degenerate), angle = acos(n.z), Rodrigues matrix with sin/cos
// rounded to 14 decimals, identity when the angle is exactly 0.
__device__ __forceinline__ V3 rotate_z_to_normal(V3 e, V3 n) {
    double ax = -n.y, ay = n.x;
    double inv_len = 1.0 / sqrt(ax * ax + ay * ay);
    ax *= inv_len; ay *= inv_len;
    if (!(ax == ax)) { ax = 1.0; ay = 0.0; }
    // angle = acos(n.z); the reference then takes sin and cos of it.  cos(acos z) = z and
    // sin(acos z) = sqrt((1 - z)(1 + z)) agree with that to a few 1e-16 (then both are rounded to 14 decimals),
    // without three float64 transcendentals.  acos(z) == 0 exactly when z >= 1.
    if (n.z >= 1.0) return e;
    double s = round14(sqrt((1.0 - n.z) * (1.0 + n.z))), c = round14(n.z), w = 1.0 - c;
    // R = outer(a,a) w + I c + [a]x s with a = (ax, ay, 0)
    double r00 = ax * ax * w + c, r01 = ax * ay * w,     r02 = ay * s;
    double r10 = ay * ax * w,     r11 = ay * ay * w + c, r12 = -ax * s;
    double r20 = -ay * s,         r21 = ax * s,          r22 = c;
    return mk(r00 * e.x + r01 * e.y + r02 * e.z,
              r10 * e.x + r11 * e.y + r12 * e.z,
              r20 * e.x + r21 * e.y + r22 * e.z);
}

// rotation_to_z(n) applied to a local direction a (spatial_geometry.py:24-48
// as used by Lambertian.__call__, optics_callables.py:163-165):
// x' = (n.y, -n.x, 0)/|.| (x-hat if degenerate), y' = n x x', z' = n.
__device__ __forceinline__ V3 rotation_to_z_apply(V3 n, V3 a) {
    double px = n.y, py = -n.x;
    if (px == 0.0 && py == 0.0) { px = 1.0; py = 0.0; }
    double len = sqrt(px * px + py * py);
    px /= len; py /= len;
    // y' = n x (px, py, 0)
    double qx = n.y * 0.0 - n.z * py;
    double qy = n.z * px - n.x * 0.0;
    double qz = n.x * py - n.y * px;
    return mk(px * a.x + qx * a.y + n.x * a.z,
              py * a.x + qy * a.y + n.y * a.z,
              0.0 * a.x + qz * a.y + n.z * a.z);
}

// pillbox_sunshape_directions / Lambertian_directions (sources.py:88-117):
// xi1 in [0, 2pi), xi2 in [0, 1).
__device__ __forceinline__ V3 pillbox_direction(double xi1, double xi2, double ang_range) {
    if (ang_range == 0.0) return mk(0.0, 0.0, 1.0);
    double ss = sin(ang_range) * sqrt(xi2);
    return mk(cos(xi1) * ss, sin(xi1) * ss, sqrt(1.0 - ss * ss));
}

// numpy.interp(x, xp, fp) for increasing xp (clamped at the ends).
__device__ __forceinline__ double interp1(double x, const double *xp, const double *fp, int n) {
    if (x <= xp[0]) return fp[0];
    if (x >= xp[n - 1]) return fp[n - 1];
    int lo = 0, hi = n - 1;
    while (hi - lo > 1) {
        int mid = (lo + hi) >> 1;
        if (xp[mid] <= x) lo = mid; else hi = mid;
    }
    double slope = (fp[lo + 1] - fp[lo]) / (xp[lo + 1] - xp[lo]);
    return slope * (x - xp[lo]) + fp[lo];
}

// Bilinear lookup on a regular (theta, wavelength) grid, the "linear" method of
// scipy.interpolate.RegularGridInterpolator used at optics_callables.py:371.
// Layout at xp: thetas[nth], wavelengths[nwl], values[nth*nwl] (row-major).
__device__ __forceinline__ int grid_cell(double x, const double *g, int n) {
    int lo = 0, hi = n - 1;
    while (hi - lo > 1) {
        int mid = (lo + hi) >> 1;
        if (g[mid] <= x) lo = mid; else hi = mid;
    }
    return lo;
}
__device__ __forceinline__ double interp2(double th, double wl, const double *xp, int nth, int nwl) {
    const double *gt = xp, *gw = xp + nth, *val = xp + nth + nwl;
    int i = grid_cell(th, gt, nth), j = grid_cell(wl, gw, nwl);
    double a = (th - gt[i]) / (gt[i + 1] - gt[i]);
    double b = (wl - gw[j]) / (gw[j + 1] - gw[j]);
    double v00 = val[i * nwl + j], v01 = val[i * nwl + j + 1];
    double v10 = val[(i + 1) * nwl + j], v11 = val[(i + 1) * nwl + j + 1];
    return (1.0 - a) * ((1.0 - b) * v00 + b * v01) + a * ((1.0 - b) * v10 + b * v11);
}

// RefractiveHomogenous / RefractiveTransmissiveHomogenous (optics_callables.py:1186-1296, :1323-1345)
// for one hit: the (possibly perturbed) normal, Snell's law in vector form with the total-internal-
// reflection test (optics.refractions, optics.py:159-192), the unpolarised Fresnel reflectance
// (optics.fresnel, optics.py:13-39), the reflect-or-refract draw of single_ray mode and the
// path-length attenuation of Absorbant.attenuate (:874-889) with imposed coefficients.
// u[0] = normal(scale=sigma), u[1] = uniform(0, 2 pi), u[2] = uniform(0, 1).
struct RefractOut {
    V3 dir_refl, dir_refr;
    double e_refl, e_refr;     // energies of the two children
    double m2;                 // index of the medium a refracted ray enters
    bool has_refl, has_refr;   // which of the two child slots exist
};
__device__ __forceinline__ RefractOut apply_refractive(const TrcSurfaceRec *s, V3 d, V3 n, double e_in, double m1,
                                                        const double *u, V3 v_parent, V3 hit) {
    const double *op = s->op;
    RefractOut r;
    if (op[3] >= 0.0) {        // _refract_dirs :1229-1241: normal = rotation_to_z(normal) . error vector
        double st = sin(u[0]);
        n = rotation_to_z_apply(n, mk(st * cos(u[1]), st * sin(u[1]), cos(u[0])));
    }
    double m2 = (m1 == op[0]) ? op[1] : op[0];          // toggle_ref_idx :1204-1219
    r.m2 = m2;
    double nn = m2 / m1;
    double cos1 = n.x * d.x + n.y * d.y + n.z * d.z;
    bool refr = cos1 * cos1 >= 1.0 - nn * nn;
    double R = 1.0;
    r.dir_refr = mk(0.0, 0.0, 0.0);
    if (refr) {
        double cos2 = sqrt(1.0 - 1.0 / (nn * nn) * (1.0 - cos1 * cos1));
        double sg = cos1 < 0.0 ? -1.0 : 1.0;
        r.dir_refr = mk((d.x - cos1 * n.x) / nn + n.x * cos2 * sg,
                        (d.y - cos1 * n.y) / nn + n.y * cos2 * sg,
                        (d.z - cos1 * n.z) / nn + n.z * cos2 * sg);
        double th = acos(fabs(cos1));
        double foo = cos(th);
        double q = m1 / m2 * sin(th);
        double bar = sqrt(1.0 - q * q);
        double rs = (m1 * foo - m2 * bar) / (m1 * foo + m2 * bar);
        double rp = (m1 * bar - m2 * foo) / (m1 * bar + m2 * foo);
        R = (rs * rs + rp * rp) / 2.0;
    }
    r.dir_refl = reflect(d, n);
    if (op[2] != 0.0) {        // single_ray :1252-1281
        bool refl = u[2] <= R;
        r.has_refl = refl; r.has_refr = !refl;
        r.e_refl = e_in; r.e_refr = e_in;
    } else {                   // :1284-1294
        r.has_refl = true; r.has_refr = refr;
        r.e_refl = e_in * R; r.e_refr = e_in * (1.0 - R);
    }
    if (op[4] != 0.0) {        // Absorbant.attenuate :874-889 on every outgoing ray
        double dx = hit.x - v_parent.x, dy = hit.y - v_parent.y, dz = hit.z - v_parent.z;
        double path = sqrt(dx * dx + dy * dy + dz * dz);
        if (op[7] != 1.0) path *= op[7];
        double a_c = (op[4] > 1.0 && m1 == op[1]) ? op[6] : op[5];
        double f = exp(-a_c * path);
        r.e_refl *= f; r.e_refr *= f;
    }
    return r;
}

// ---- refraction between media given by material tables (TRC_F_TABULATED): complex indices
struct Cx { double re, im; };
__device__ __forceinline__ Cx cx(double re, double im) { Cx z; z.re = re; z.im = im; return z; }
__device__ __forceinline__ Cx cadd(Cx a, Cx b) { return cx(a.re + b.re, a.im + b.im); }
__device__ __forceinline__ Cx csub(Cx a, Cx b) { return cx(a.re - b.re, a.im - b.im); }
__device__ __forceinline__ Cx cmul(Cx a, Cx b) { return cx(a.re * b.re - a.im * b.im, a.re * b.im + a.im * b.re); }
__device__ __forceinline__ Cx cscale(Cx a, double k) { return cx(a.re * k, a.im * k); }
__device__ __forceinline__ Cx cdiv(Cx a, Cx b) {
    double den = b.re * b.re + b.im * b.im;
    return cx((a.re * b.re + a.im * b.im) / den, (a.im * b.re - a.re * b.im) / den);
}
__device__ __forceinline__ Cx csqrt(Cx z) {          // principal branch, as numpy.sqrt of a complex array
    double r = hypot(z.re, z.im);
    if (r == 0.0) return cx(0.0, 0.0);
    double a = sqrt(0.5 * (r + fabs(z.re)));
    double b = 0.5 * z.im / a;
    if (z.re >= 0.0) return cx(a, b);
    return cx(fabs(b), z.im < 0.0 ? -a : a);
}
// m(lambda) of a material: knots at `t` = lambdas[n], then n[n], then k[n]; linear, clamped at the ends
__device__ __forceinline__ Cx material_m(const double *t, int n, double wl) {
    return cx(interp1(wl, t, t + n, n), interp1(wl, t, t + 2 * n, n));
}
struct RefractTabOut {
    V3 dir_refl, dir_refr;
    double e_refl, e_refr;
    Cx m2;
    bool has_refl, has_refr;
};
// Refractive._make_refraction_bundle (:836-842) for one hit: m1 the ray's complex index, wl its wavelength.
__device__ __forceinline__ RefractTabOut apply_refractive_tab(const TrcSurfaceRec *s, const double *tables, V3 d, V3 n,
                                                                double e_in, Cx m1, double wl, const double *u) {
    const double *op = s->op;
    RefractTabOut r;
    if (op[3] >= 0.0) {        // _refract_dirs :765-777
        double st = sin(u[0]);
        n = rotation_to_z_apply(n, mk(st * cos(u[1]), st * sin(u[1]), cos(u[0])));
    }
    // toggle_ref_idx :744-756.  The reference compares m1 == material_1.m(wavelength) exactly; the ray's index was
    // produced by the same linear interpolation on the host or on the device, which may differ in the last bits
    // (fused multiply-add), so "equal" is taken to 1e-9 relative here.
    const Cx ma = material_m(tables + s->table_off, s->table_len, wl);
    const Cx mb = material_m(tables + (int)op[0], (int)op[1], wl);
    const bool in_a = fabs(m1.re - ma.re) <= 1e-9 * fabs(ma.re) && fabs(m1.im - ma.im) <= 1e-9 * fabs(ma.im) + 1e-300;
    const Cx m2 = in_a ? mb : ma;
    r.m2 = m2;
    const double nn = m2.re / m1.re;                     // optics.refractions(m1.real, m2.real, ...) :785
    const double cos1 = n.x * d.x + n.y * d.y + n.z * d.z;
    const bool refr = cos1 * cos1 >= 1.0 - nn * nn;
    double R = 1.0;
    r.dir_refr = mk(0.0, 0.0, 0.0);
    if (refr) {
        double cos2 = sqrt(1.0 - 1.0 / (nn * nn) * (1.0 - cos1 * cos1));
        double sg = cos1 < 0.0 ? -1.0 : 1.0;
        r.dir_refr = mk((d.x - cos1 * n.x) / nn + n.x * cos2 * sg,
                        (d.y - cos1 * n.y) / nn + n.y * cos2 * sg,
                        (d.z - cos1 * n.z) / nn + n.z * cos2 * sg);
        // optics.fresnel (optics.py:13-39) in complex arithmetic; the float array R keeps the real part (:839)
        const double th = acos(fabs(cos1));
        const double foo = cos(th), sn = sin(th);
        const Cx q = cscale(cdiv(m1, m2), sn);
        const Cx bar = csqrt(csub(cx(1.0, 0.0), cmul(q, q)));
        const Cx rs = cdiv(csub(cscale(m1, foo), cmul(m2, bar)), cadd(cscale(m1, foo), cmul(m2, bar)));
        const Cx rp = cdiv(csub(cmul(m1, bar), cscale(m2, foo)), cadd(cmul(m1, bar), cscale(m2, foo)));
        R = 0.5 * (cmul(rs, rs).re + cmul(rp, rp).re);
    }
    r.dir_refl = reflect(d, n);
    if (op[2] != 0.0) {        // single_ray :792-823
        bool refl = u[2] <= R;
        r.has_refl = refl; r.has_refr = !refl;
        r.e_refl = e_in; r.e_refr = e_in;
    } else {                   // :826-834
        r.has_refl = true; r.has_refr = refr;
        r.e_refl = e_in * R; r.e_refr = e_in * (1.0 - R);
    }
    return r;
}

struct OpticsOut { V3 dir; double energy; double attenuated; };

// |(d.n) n|: the cosine of the incidence angle as the reference forms it
__device__ __forceinline__ double cos_incidence(V3 d, V3 n) {
    double k = dot(d, n);
    V3 vert = mk(k * n.x, k * n.y, k * n.z);
    return sqrt(vert.x * vert.x + vert.y * vert.y + vert.z * vert.z);
}

// u[0..2]: draws already in the reference's units: normal(scale=sigma) values,
// uniform(0, 2pi) angles, uniform(0,1) variates (see TrcOpticsArgs).
__device__ __forceinline__ OpticsOut apply_optics(const TrcSurfaceRec *s, const double *tables,
                                                  V3 d, V3 n, double e_in, double wavelength,
                                                  const double *u, V3 v_parent, V3 hit) {
    OpticsOut o;
    o.attenuated = 0.0;
    const double *op = s->op;
    switch (s->okind) {
    case TRC_O_TRANSPARENT: case TRC_O_PERIODIC:     // periodic: the energy leaves with the second child
        o.dir = d; o.energy = e_in; break;
    case TRC_O_REFLECTIVE:
        o.dir = reflect(d, n); o.energy = e_in * (1.0 - op[0]); break;
    case TRC_O_REAL_REFLECTIVE: {
        V3 rn = n;
        if (op[1] > 0.0) {
            V3 e;
            if (op[2] != 0.0) {  // bi-variate slope error (optics_callables.py:235-243)
                double tx = tan(u[0]), ty = tan(u[1]);
                double ez = rsqrt(1.0 + tx * tx + ty * ty);
                e = mk(tx * ez, ty * ez, ez);
            } else {             // conical (:245-250)
                double st = sin(u[0]);
                e = mk(st * cos(u[1]), st * sin(u[1]), cos(u[0]));
            }
            rn = rotate_z_to_normal(e, n);
            double inv_len = rsqrt(rn.x * rn.x + rn.y * rn.y + rn.z * rn.z);
            rn = mk(rn.x * inv_len, rn.y * inv_len, rn.z * inv_len);
        }
        o.dir = reflect(d, rn); o.energy = e_in * (1.0 - op[0]); break;
    }
    case TRC_O_LAMBERTIAN: {
        V3 a = pillbox_direction(u[0], u[1], op[1]);
        o.dir = rotation_to_z_apply(n, a); o.energy = e_in * (1.0 - op[0]); break;
    }
    case TRC_O_LAMBERTIAN_SPECULAR: {
        if (u[0] < op[1]) o.dir = reflect(d, n);
        else o.dir = rotation_to_z_apply(n, pillbox_direction(u[1], u[2], M_PI / 2.0));
        o.energy = e_in * (1.0 - op[0]); break;
    }
    case TRC_O_LAMB_DIRECTIONAL: case TRC_O_LAMBSPEC_DIRECTIONAL: case TRC_O_LAMB_DIRECTIONAL_SPECTRAL: {
        // optics_callables.py:340-358, :370-388, :437-455: absorptance looked up
        // by the incidence angle acos(|(d.n) n|).
        const double *xp = tables + s->table_off;
        int nth = s->table_len;
        double k = dot(d, n);
        V3 vert = mk(k * n.x, k * n.y, k * n.z);
        double th = acos(sqrt(vert.x * vert.x + vert.y * vert.y + vert.z * vert.z));
        double ab;
        if (s->okind == TRC_O_LAMB_DIRECTIONAL_SPECTRAL) ab = interp2(th, wavelength, xp, nth, (int)op[1]);
        else ab = interp1(th, xp, xp + nth, nth);
        if (s->okind == TRC_O_LAMBSPEC_DIRECTIONAL) {
            if (u[0] < op[0]) o.dir = reflect(d, n);
            else o.dir = rotation_to_z_apply(n, pillbox_direction(u[1], u[2], M_PI / 2.0));
        } else {
            o.dir = rotation_to_z_apply(n, pillbox_direction(u[0], u[1], M_PI / 2.0));
        }
        o.energy = e_in * (1.0 - ab); break;
    }
    case TRC_O_LAMB_PIECEWISE_SPEC: {
        // optics_callables.py:466-487: absorptance and specularity both follow the incidence angle
        const double *xp = tables + s->table_off;
        int nth = s->table_len;
        double th = acos(cos_incidence(d, n));
        double ab = interp1(th, xp, xp + nth, nth);
        double spec = interp1(th, xp, xp + 2 * nth, nth);
        if (u[0] < spec) o.dir = reflect(d, n);
        else o.dir = rotation_to_z_apply(n, pillbox_direction(u[1], u[2], M_PI / 2.0));
        o.energy = e_in * (1.0 - ab); break;
    }
    case TRC_O_LAMBERTIAN_ABSORBANT: {
        // LambertianAbsorbant.__call__ (:901-906): Lambertian with zero absorptivity, then
        // Absorbant.attenuate (:874-889) over the path from the parent vertex, then absorption
        o.dir = rotation_to_z_apply(n, pillbox_direction(u[0], u[1], op[1]));
        double e0 = e_in * (1.0 - 0.0);
        double dx = hit.x - v_parent.x, dy = hit.y - v_parent.y, dz = hit.z - v_parent.z;
        double path = sqrt(dx * dx + dy * dy + dz * dz);
        if (op[3] != 1.0) path *= op[3];
        double e_att = e0 * exp(-op[2] * path);
        o.attenuated = e_in - e_att;
        o.energy = e_att * (1.0 - op[0]); break;
    }
    case TRC_O_FRESNEL_CONDUCTOR: {
        // optics.py:41-81 (Modest ch. 2: perfect dielectric n1 -> absorbing medium m2 = n2 + i k2)
        const double *xp = tables + s->table_off;
        int nw = s->table_len;
        o.dir = reflect(d, n);
        if (!(wavelength >= xp[0]) || !(wavelength <= xp[nw - 1])) {     // optical_material.check_valid: NaN
            o.energy = e_in * nan(""); break;
        }
        double n2 = interp1(wavelength, xp, xp + nw, nw), k2 = interp1(wavelength, xp, xp + 2 * nw, nw);
        double n1 = op[0];
        double th1 = acos(fabs(n.x * d.x + n.y * d.y + n.z * d.z));
        double s1 = n1 * sin(th1), c1 = n1 * cos(th1);
        double b = n2 * n2 - k2 * k2 - s1 * s1;
        double nk = n2 * k2;
        double a = sqrt(b * b + 4.0 * (nk * nk));
        double pp = sqrt(0.5 * (a + b)), qq = sqrt(0.5 * (a - b));
        double rs = ((c1 - pp) * (c1 - pp) + qq * qq) / ((c1 + pp) * (c1 + pp) + qq * qq);
        double st = s1 * tan(th1);
        double rp = ((pp - st) * (pp - st) + qq * qq) / ((pp + st) * (pp + st) + qq * qq) * rs;
        o.energy = e_in * ((rp + rs) / 2.0); break;
    }
    case TRC_O_REFLECTIVE_SPECTRAL: {
        const double *xp = tables + s->table_off;
        const double *fp = xp + s->table_len;
        o.dir = reflect(d, n);
        o.energy = e_in * (1.0 - interp1(wavelength, xp, fp, s->table_len)); break;
    }
    default:
        o.dir = d; o.energy = e_in; break;
    }
    if (s->flags & TRC_F_IAM) {
        // IAM.__call__ (:277-282) overwrites the energy of the base optics
        double c = cos_incidence(d, n);
        o.energy = e_in * (1.0 - op[0]) * (1.0 - exp(-(op[5] == 1.0 ? c : pow(c, op[5])) / op[4])) / (1.0 - exp(-1.0 / op[4]));
    }
    return o;
}

}  // namespace trc
