// Ray/surface intersection and surface normals for every geometry-manager kind
// on the hot path.  One function per family instead of the reference's
// find_intersections / select_rays / get_normals protocol (surface.py:55-112):
// a thread owns one ray, evaluates the parametric hit `t` (+inf = miss) and, in
// the optics kernel, the normal at the hit.  All arithmetic is float64 so the
// absolute thresholds of the reference (1e-7 flat, 1e-6 quadric, 1e-9 cone)
// keep their meaning and hit indices are bit-exact against the NumPy oracle.
#pragma once
#include <math.h>
#include "../../include/tracer_b200.h"

namespace trc {

struct V3 { double x, y, z; };

__device__ __forceinline__ V3 mk(double x, double y, double z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ __forceinline__ double dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ V3 axpy(V3 v, double t, V3 d) { return mk(v.x + t * d.x, v.y + t * d.y, v.z + t * d.z); }

// rows 0..2 of a 4x4 homogeneous matrix times [p; 1]
__device__ __forceinline__ V3 xform_point(const double *m, V3 p) {
    return mk(m[0] * p.x + m[1] * p.y + m[2] * p.z + m[3],
              m[4] * p.x + m[5] * p.y + m[6] * p.z + m[7],
              m[8] * p.x + m[9] * p.y + m[10] * p.z + m[11]);
}
// R^T d  (global -> local direction), R = frame[:3,:3]
__device__ __forceinline__ V3 rot_to_local(const double *f, V3 d) {
    return mk(f[0] * d.x + f[4] * d.y + f[8] * d.z,
              f[1] * d.x + f[5] * d.y + f[9] * d.z,
              f[2] * d.x + f[6] * d.y + f[10] * d.z);
}
// R n  (local -> global direction)
__device__ __forceinline__ V3 rot_to_global(const double *f, V3 n) {
    return mk(f[0] * n.x + f[1] * n.y + f[2] * n.z,
              f[4] * n.x + f[5] * n.y + f[6] * n.z,
              f[8] * n.x + f[9] * n.y + f[10] * n.z);
}

#define TRC_INF (__longlong_as_double(0x7ff0000000000000LL))

// ---------------------------------------------------------------- flat family
// FlatGeometryManager.find_intersections (flat_surface.py:16-62) followed by
// the aperture trims of FiniteFlatGM subclasses (:202-211, :266-274, :484-492,
// :553-560) and TriangularFace (triangular_face.py:36-74).
__device__ __forceinline__ double isect_flat(const TrcSurfaceRec *s, int kind, V3 v, V3 d) {
    const double *f = s->frame;
    V3 n = mk(f[2], f[6], f[10]);
    V3 c = mk(f[3], f[7], f[11]);
    double dt = dot(d, n);
    if (!(fabs(dt) > 1e-7)) return TRC_INF;
    double vt = dot(n, mk(v.x - c.x, v.y - c.y, v.z - c.z));
    double t = -vt / dt;
    if (t < 1e-7) return TRC_INF;
    if (kind == TRC_G_FLAT_INFINITE) return t;
    V3 g = axpy(v, t, d);
    const double *gp = s->gp;
    if (kind == TRC_G_TRIANGLE) {
        // rel_glob = R @ verts ; w = g - c ; barycentric test.
        V3 e0 = rot_to_global(f, mk(gp[0], gp[1], gp[2]));
        V3 e1 = rot_to_global(f, mk(gp[3], gp[4], gp[5]));
        V3 w = mk(g.x - c.x, g.y - c.y, g.z - c.z);
        double uv = gp[0] * gp[3] + gp[1] * gp[4] + gp[2] * gp[5];
        double n0 = gp[0] * gp[0] + gp[1] * gp[1] + gp[2] * gp[2];
        double n1 = gp[3] * gp[3] + gp[4] * gp[4] + gp[5] * gp[5];
        double r0 = dot(w, e0), r1 = dot(w, e1);
        double den = uv * uv - n0 * n1;
        double b0 = (uv * r1 - n1 * r0) / den;
        double b1 = (uv * r0 - n0 * r1) / den;
        if (b0 < 0.0 || b1 < 0.0 || (b0 + b1) > 1.0) return TRC_INF;
        return t;
    }
    const double *iv = s->inv;
    double lx = iv[0] * g.x + iv[1] * g.y + iv[2] * g.z + iv[3];
    double ly = iv[4] * g.x + iv[5] * g.y + iv[6] * g.z + iv[7];
    if (kind == TRC_G_RECT_PLATE) {
        if (fabs(lx) > gp[0] || fabs(ly) > gp[1]) return TRC_INF;
        return t;
    }
    if (kind == TRC_G_RECT_EXTRUDED) {
        if (fabs(lx) > gp[0] || fabs(ly) > gp[1]) return TRC_INF;
        if (fabs(lx - gp[2]) < gp[4] && fabs(ly - gp[3]) < gp[5]) return TRC_INF;
        return t;
    }
    // TRC_G_ROUND_PLATE
    double r2 = lx * lx + ly * ly;
    if (r2 > gp[0] * gp[0]) return TRC_INF;
    if (gp[1] >= 0.0 && r2 < gp[1] * gp[1]) return TRC_INF;
    if (lx > gp[2]) return TRC_INF;
    return t;
}

// The terms of the triangle test that depend on the face only, written exactly as isect_flat's triangle branch
// writes them (same expressions, hence the same fused multiply-adds): the tree walk of a mesh scene evaluates
// them once per scene into a compact table (128 bytes per face instead of the 384-byte surface record).
struct TriTerms { V3 e0, e1; double uv, n0, n1, den; };
__device__ __forceinline__ TriTerms tri_terms(const double *f, const double *gp) {
    TriTerms r;
    r.e0 = rot_to_global(f, mk(gp[0], gp[1], gp[2]));
    r.e1 = rot_to_global(f, mk(gp[3], gp[4], gp[5]));
    r.uv = gp[0] * gp[3] + gp[1] * gp[4] + gp[2] * gp[5];
    r.n0 = gp[0] * gp[0] + gp[1] * gp[1] + gp[2] * gp[2];
    r.n1 = gp[3] * gp[3] + gp[4] * gp[4] + gp[5] * gp[5];
    r.den = r.uv * r.uv - r.n0 * r.n1;
    return r;
}
// One record of the compact table: [c(3) n(3) e0(3) e1(3) uv n0 n1 den]; den == 0 marks "not a triangle"
// (the walk then reads the full surface record).
#define TRC_TRI_REC_DOUBLES 16
// isect_flat for a triangle given by its compact record: c = frame origin, n = frame z axis.
__device__ __forceinline__ double isect_triangle(V3 c, V3 n, const TriTerms &tt, V3 v, V3 d) {
    double dt = dot(d, n);
    if (!(fabs(dt) > 1e-7)) return TRC_INF;
    double vt = dot(n, mk(v.x - c.x, v.y - c.y, v.z - c.z));
    double t = -vt / dt;
    if (t < 1e-7) return TRC_INF;
    V3 g = axpy(v, t, d);
    V3 w = mk(g.x - c.x, g.y - c.y, g.z - c.z);
    double r0 = dot(w, tt.e0), r1 = dot(w, tt.e1);
    double b0 = (tt.uv * r1 - tt.n1 * r0) / tt.den;
    double b1 = (tt.uv * r0 - tt.n0 * r1) / tt.den;
    if (b0 < 0.0 || b1 < 0.0 || (b0 + b1) > 1.0) return TRC_INF;
    return t;
}

// Circular perforations: the hit is void when strictly inside any hole
// (flat_surface.py:372-373, polygon.py:155-158).  h = (cx, cy, r) triples.
__device__ __forceinline__ bool in_a_hole(const double *h, int n, double lx, double ly) {
    for (int k = 0; k < n; ++k) {
        double dx = lx - h[3 * k], dy = ly - h[3 * k + 1];
        if (sqrt(dx * dx + dy * dy) < h[3 * k + 2]) return true;
    }
    return false;
}

// FlatSimplePolygonGM.in_poly (polygon.py:31-53): crossing number of the ray from
// the point towards +x against the closed profile xs[0..nv], ys[0..nv].  An edge
// wholly at x >= px counts when its ends straddle py; an edge straddling both px and
// py counts when the crossing abscissa (:55-63) is >= px.
__device__ __forceinline__ bool in_polygon(const double *xs, const double *ys, int nv, double px, double py) {
    int crossings = 0;
    bool xp0 = px <= xs[0], yp0 = py <= ys[0];
    for (int k = 0; k < nv; ++k) {
        bool xp1 = px <= xs[k + 1], yp1 = py <= ys[k + 1];
        if (yp0 != yp1) {
            if (xp0 && xp1) {
                ++crossings;
            } else if (xp0 != xp1) {
                double a = (ys[k + 1] - ys[k]) / (xs[k + 1] - xs[k]);
                double y0 = ys[k] - a * xs[k];
                if ((py - y0) / a >= px) ++crossings;
            }
        }
        xp0 = xp1; yp0 = yp1;
    }
    return (crossings & 1) != 0;
}

// Flat kinds whose aperture lives in the table pool (TRC_G_RECT_PERFORATED, TRC_G_POLYGON).
// Kept out of isect_flat so that the flat-only tree walk stays small.
static __device__ __noinline__ double isect_flat_pool(const TrcSurfaceRec *s, int kind, V3 v, V3 d, const double *pool) {
    double t = isect_flat(s, TRC_G_FLAT_INFINITE, v, d);
    if (t == TRC_INF) return t;
    V3 g = axpy(v, t, d);
    const double *iv = s->inv;
    const double *gp = s->gp;
    double lx = iv[0] * g.x + iv[1] * g.y + iv[2] * g.z + iv[3];
    double ly = iv[4] * g.x + iv[5] * g.y + iv[6] * g.z + iv[7];
    if (kind == TRC_G_RECT_PERFORATED) {
        if (fabs(lx) > gp[0] || fabs(ly) > gp[1]) return TRC_INF;
        if (in_a_hole(pool + (int)gp[2], (int)gp[3], lx, ly)) return TRC_INF;
        return t;
    }
    // TRC_G_POLYGON
    const double *xs = pool + (int)gp[0];
    const int nv = (int)gp[1];
    if (!in_polygon(xs, xs + nv + 1, nv, lx, ly)) return TRC_INF;
    if (in_a_hole(xs + 2 * (nv + 1), (int)gp[2], lx, ly)) return TRC_INF;
    return t;
}

// Flat normal always opposes the incoming ray (flat_surface.py:58-60, 84-91).
__device__ __forceinline__ V3 normal_flat(const TrcSurfaceRec *s, V3 d) {
    const double *f = s->frame;
    V3 n = mk(f[2], f[6], f[10]);
    if (dot(d, n) > 0.0) n = mk(-n.x, -n.y, -n.z);
    return n;
}

// ------------------------------------------------------------- quadric family
// Aperture / validity of one candidate root, i.e. the per-class
// _select_coords predicates (quadric.py:115-144, paraboloid.py:89-119,
// :193-223, :257-294, :411-443, sphere_surface.py:122-139, :211-228,
// cone.py:94-120, :285-320, :365-395, cylinder.py:71-110, :171-199,
// quadratic_surface.py:75-105, ellipsoid.py:77-116).  `dflt` reports whether the class falls back
// to QuadricGM's default choice when both roots pass.
__device__ __forceinline__ bool quadric_root_ok(const TrcSurfaceRec *s, int kind, V3 p, double t) {
    const double *gp = s->gp;
    if (!(t == t)) return false;
    switch (kind) {
    case TRC_G_PARABOLOID: case TRC_G_SPHERE: case TRC_G_CONE_INFINITE:
    case TRC_G_CYL_INFINITE: case TRC_G_QUADRIC_FLAT: case TRC_G_PARAB_CYL:
        return t >= 1e-6;
    default: break;
    }
    if (kind == TRC_G_SPHERE_CUT) {
        if (gp[5] != 0.0)      // BoundaryPlane.in_bounds (boundary_shape.py:151-165): gp[1..4] = row 2 of the inverse frame
            return t > 1e-6 && (gp[1] * p.x + gp[2] * p.y + gp[3] * p.z + gp[4]) >= 0.0;
        // BoundarySphere.in_bounds (boundary_shape.py:104-111)
        double dx = p.x - gp[1], dy = p.y - gp[2], dz = p.z - gp[3];
        return t > 1e-6 && gp[4] * gp[4] >= dx * dx + dy * dy + dz * dz;
    }
    V3 l = xform_point(s->inv, p);
    switch (kind) {
    case TRC_G_PARAB_DISH:
        return t > 1e-6 && l.z <= gp[2] && l.z >= 0.0;
    case TRC_G_PARAB_HEX: {
        double ax = fabs(l.x), ay = fabs(l.y);
        bool outside = ax > sqrt(3.0) * gp[2] / 2.0;
        outside |= ay > gp[2] - tan(M_PI / 6.0) * ax;
        return !outside && t > 0.0;
    }
    case TRC_G_PARAB_RECT:
        return !(fabs(l.x) > gp[2] || fabs(l.y) > gp[3]) && t > 1e-6;
    case TRC_G_PARAB_TROUGH:
        return fabs(l.y) <= gp[1] && l.z <= gp[2] && l.z >= 0.0 && t > 1e-6;
    case TRC_G_HEMISPHERE:
        return l.z <= 0.0 && t > 1e-6;
    case TRC_G_SPHERE_RECT:
        return l.z <= 0.0 && t > 1e-6 && fabs(l.x) <= gp[1] && fabs(l.y) <= gp[2];
    case TRC_G_CONE_FINITE:
        return l.z >= 0.0 && l.z <= gp[2] && t > 1e-9;
    case TRC_G_FRUSTUM:
        return gp[2] <= l.z && l.z <= gp[3] && t > 1e-6;
    case TRC_G_FRUSTUM_RECT:
        return gp[2] <= l.z && l.z <= gp[3] && fabs(l.x) <= gp[4] && fabs(l.y) <= gp[5] && t > 1e-6;
    case TRC_G_CYL_FINITE: {
        double ang = atan2(l.y, l.x);
        if (ang < 0.0) ang = 2.0 * M_PI + ang;
        return fabs(l.z) <= gp[1] && ang >= gp[2] && ang <= gp[3] && t > 1e-6;
    }
    case TRC_G_CYL_RECT:
        return -gp[1] <= l.z && l.z <= gp[1] && fabs(l.x) <= gp[2] && fabs(l.y) <= gp[3] && t > 1e-6;
    case TRC_G_QUADRIC_RECT:
        return !(fabs(l.x) > gp[6] || fabs(l.y) > gp[7]) && t > 1e-6;
    case TRC_G_ELLIPSOID:         // ellipsoid.py:93-110 (1e-7, limits inclusive)
        return t > 1e-7 && l.x >= gp[3] && l.x <= gp[4] && l.y >= gp[5] && l.y <= gp[6] &&
               l.z >= gp[7] && l.z <= gp[8];
    }
    return false;
}

// QuadricGM.find_intersections (quadric.py:28-113) with the per-class
// get_ABC: paraboloid.py:28-42, :343-356, sphere_surface.py:52-68,
// cone.py:59-72, cylinder.py:32-57, quadratic_surface.py:46-61.
__device__ __forceinline__ double isect_quadric(const TrcSurfaceRec *s, int kind, V3 v, V3 d) {
    const double *gp = s->gp;
    double A, B, C;
    if (kind == TRC_G_SPHERE || kind == TRC_G_HEMISPHERE || kind == TRC_G_SPHERE_RECT ||
        kind == TRC_G_SPHERE_CUT) {
        const double *f = s->frame;
        V3 w = mk(v.x - f[3], v.y - f[7], v.z - f[11]);
        A = d.x * d.x + d.y * d.y + d.z * d.z;
        B = 2.0 * (d.x * w.x + d.y * w.y + d.z * w.z);
        C = (w.x * w.x + w.y * w.y + w.z * w.z) - gp[0] * gp[0];
    } else {
        V3 dl = rot_to_local(s->frame, d);
        V3 vl = xform_point(s->inv, v);
        switch (kind) {
        case TRC_G_PARABOLOID: case TRC_G_PARAB_DISH: case TRC_G_PARAB_HEX: case TRC_G_PARAB_RECT: {
            double a = gp[0], b = gp[1];
            A = a * (dl.x * dl.x) + b * (dl.y * dl.y);
            B = 2.0 * a * dl.x * vl.x + 2.0 * b * dl.y * vl.y - dl.z;
            C = a * (vl.x * vl.x) + b * (vl.y * vl.y) - vl.z;
            break;
        }
        case TRC_G_PARAB_CYL: case TRC_G_PARAB_TROUGH: {
            double a = gp[0];
            A = a * (dl.x * dl.x);
            B = 2.0 * a * dl.x * vl.x - dl.z;
            C = a * (vl.x * vl.x) - vl.z;
            break;
        }
        case TRC_G_CONE_INFINITE: case TRC_G_CONE_FINITE: case TRC_G_FRUSTUM: case TRC_G_FRUSTUM_RECT: {
            double c = gp[0], a = gp[1];
            double cdz = c * dl.z, cz = c * (vl.z - a);
            A = dl.x * dl.x + dl.y * dl.y - cdz * cdz;
            B = 2.0 * (vl.x * dl.x + vl.y * dl.y - (c * c) * (vl.z - a) * dl.z);
            C = vl.x * vl.x + vl.y * vl.y - cz * cz;
            break;
        }
        case TRC_G_CYL_INFINITE: case TRC_G_CYL_FINITE: case TRC_G_CYL_RECT: {
            A = dl.x * dl.x + dl.y * dl.y;
            B = 2.0 * (dl.x * vl.x + dl.y * vl.y);
            C = (vl.x * vl.x + vl.y * vl.y) - gp[0] * gp[0];
            break;
        }
        case TRC_G_ELLIPSOID: {   // ellipsoid.py:20-34
            double a = gp[0], b = gp[1], c = gp[2];
            A = a * (dl.x * dl.x) + b * (dl.y * dl.y) + c * (dl.z * dl.z);
            B = 2.0 * a * dl.x * vl.x + 2.0 * b * dl.y * vl.y + 2.0 * c * dl.z * vl.z;
            C = a * (vl.x * vl.x) + b * (vl.y * vl.y) + c * (vl.z * vl.z) - 1.0;
            break;
        }
        default: {  // TRC_G_QUADRIC_FLAT / TRC_G_QUADRIC_RECT
            double a = gp[0], b = gp[1], c = gp[2], dd = gp[3], e = gp[4], ff = gp[5];
            A = a * (dl.x * dl.x) + b * (dl.y * dl.y) + c * dl.x * dl.y;
            B = 2.0 * a * dl.x * vl.x + 2.0 * b * dl.y * vl.y + c * (vl.x * dl.y + vl.y * dl.x) +
                dd * dl.x + e * dl.y - dl.z;
            C = a * (vl.x * vl.x) + b * (vl.y * vl.y) + c * vl.x * vl.y + dd * vl.x + e * vl.y + ff - vl.z;
            break;
        }
        }
    }
    double delta = B * B - 4.0 * A * C;
    if (!(delta >= 1e-6)) return TRC_INF;
    double sq = sqrt(delta);
    double t0, t1;
    if (A == 0.0) {
        if (B == 0.0) return TRC_INF;
        t0 = t1 = -C / B;
    } else if (B == 0.0) {
        double r = sqrt(-C / A);
        t0 = -r; t1 = r;
    } else {
        double sgn = (B > 0.0) ? 1.0 : -1.0;
        double q = -0.5 * (B + sgn * sq);
        t0 = q / A;
        t1 = C / q;
    }
    bool ok0 = quadric_root_ok(s, kind, axpy(v, t0, d), t0);
    bool ok1 = quadric_root_ok(s, kind, axpy(v, t1, d), t1);
    if (kind == TRC_G_PARAB_HEX && ok0 && ok1) {
        // both inside the hexagon: QuadricGM's default choice decides (paraboloid.py:208)
        ok0 = t0 >= 1e-6; ok1 = t1 >= 1e-6;
    }
    // both valid -> index 1 (quadric.py:137-138); exactly one -> that one
    if (ok1) return t1;
    if (ok0) return t0;
    return TRC_INF;
}

// Normals of the quadric family (paraboloid.py:44-69, :358-384,
// sphere_surface.py:35-50, cone.py:28-57, cylinder.py:20-30,
// quadratic_surface.py:22-44).  p = global hit point, d = incoming direction.
__device__ __forceinline__ V3 normal_quadric(const TrcSurfaceRec *s, int kind, V3 p, V3 d) {
    const double *gp = s->gp;
    const double *f = s->frame;
    if (kind == TRC_G_SPHERE || kind == TRC_G_HEMISPHERE || kind == TRC_G_SPHERE_RECT ||
        kind == TRC_G_SPHERE_CUT) {
        V3 c = mk(f[3], f[7], f[11]);
        double sides = (c.x - p.x) * d.x + (c.y - p.y) * d.y + (c.z - p.z) * d.z;
        V3 n = mk(p.x - c.x, p.y - c.y, p.z - c.z);
        if (sides < 0.0) n = mk(-n.x, -n.y, -n.z);
        double len = sqrt(n.x * n.x + n.y * n.y + n.z * n.z);
        return mk(n.x / len, n.y / len, n.z / len);
    }
    V3 h = xform_point(s->inv, p);
    V3 dl = rot_to_local(f, d);
    V3 n;
    double flip_eps = 0.0;
    bool apex = false;
    switch (kind) {
    case TRC_G_PARABOLOID: case TRC_G_PARAB_DISH: case TRC_G_PARAB_HEX: case TRC_G_PARAB_RECT:
        n = mk(2.0 * h.x * gp[0], 2.0 * h.y * gp[1], -1.0); break;
    case TRC_G_PARAB_CYL: case TRC_G_PARAB_TROUGH:
        n = mk(2.0 * h.x * gp[0], 0.0, -1.0); break;
    case TRC_G_CONE_INFINITE: case TRC_G_CONE_FINITE: case TRC_G_FRUSTUM: case TRC_G_FRUSTUM_RECT:
        n = mk(2.0 * h.x, 2.0 * h.y, -2.0 * (h.z - gp[1]) * (gp[0] * gp[0]));
        flip_eps = 1e-9;
        apex = (h.z == gp[1]);
        break;
    case TRC_G_CYL_INFINITE: case TRC_G_CYL_FINITE: case TRC_G_CYL_RECT: {
        V3 ln = mk(h.x / gp[0], h.y / gp[0], 0.0);
        if (ln.x * dl.x + ln.y * dl.y > 0.0) ln = mk(-ln.x, -ln.y, -0.0);
        return rot_to_global(f, ln);
    }
    case TRC_G_ELLIPSOID:         // ellipsoid.py:36-61
        n = mk(2.0 * h.x * gp[0], 2.0 * h.y * gp[1], 2.0 * h.z * gp[2]); break;
    default:
        n = mk(2.0 * h.x * gp[0] + gp[2] * h.y + gp[3], 2.0 * h.y * gp[1] + gp[2] * h.x + gp[4], -1.0); break;
    }
    double len = sqrt(n.x * n.x + n.y * n.y + n.z * n.z);
    n = mk(n.x / len, n.y / len, n.z / len);
    if (dot(dl, n) > flip_eps) n = mk(-n.x, -n.y, -n.z);
    if (apex) n = mk(0.0, 0.0, -1.0);
    return rot_to_global(f, n);
}

__device__ __forceinline__ bool kind_is_flat_pool(int kind) { return kind == TRC_G_RECT_PERFORATED || kind == TRC_G_POLYGON; }
__device__ __forceinline__ bool kind_is_flat(int kind) { return kind <= TRC_G_TRIANGLE || kind_is_flat_pool(kind); }

// GeometryManager.find_intersections for any kind, plus the engine's
// "t == 0 -> inf" rule (tracer_engine.py:58).
// `pool` = the scene's table pool (apertures of the TRC_G_RECT_PERFORATED / TRC_G_POLYGON kinds).
__device__ __forceinline__ double intersect(const TrcSurfaceRec *s, V3 v, V3 d, const double *pool) {
    int kind = s->gkind;
    double t = kind <= TRC_G_TRIANGLE ? isect_flat(s, kind, v, d)
             : kind_is_flat_pool(kind) ? isect_flat_pool(s, kind, v, d, pool)
                                       : isect_quadric(s, kind, v, d);
    if (t == 0.0) t = TRC_INF;
    return t;
}

__device__ __forceinline__ V3 surface_normal(const TrcSurfaceRec *s, V3 p, V3 d) {
    int kind = s->gkind;
    return kind_is_flat(kind) ? normal_flat(s, d) : normal_quadric(s, kind, p, d);
}

}  // namespace trc
