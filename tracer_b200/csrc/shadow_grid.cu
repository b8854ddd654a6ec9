// Shadow grid of a directionally coherent bundle: host-side builder (C++, no kernel here).
//
// The first bounce of a solar trace is a bundle whose directions lie in a narrow cone (half angle of the
// sunshape, 43.6 mrad for Buie) around the sun vector.  For such rays the point where a ray crosses a fixed
// plane x_n = h0 tells which surfaces it can touch: a ray through a point p of a surface box crosses the plane
// at key = p_ab - (p_n - h0) * (d_ab / d_n), so the keys of all rays of the cone that touch the box fill the
// box's "shadow" -- its (a, b) rectangle swept by (p_n - h0) times the cone's slope box.  The grid stores, per
// cell of the plane, the surfaces whose shadow covers it; the tree-walk kernel (nearest.cu, stage A) looks the
// cell up instead of walking the tree, and every candidate still takes the float64 test of the reference, so
// the hit indices are the walk's bit for bit.  Rays outside the slope box take the walk.
//
// Replaces, for such bundles, KdTree.traversal (accel_tree.py:213-312) as the pruning structure; the reference
// has no counterpart of the grid itself.
#include <algorithm>
#include <cmath>
#include <vector>

#include "common.cuh"

namespace {

struct AxisChoice {
    int n, a, b;
    double sgn, t0a, t0b, da, db, h0, smax, cost;
    bool ok;
};

// bounding box of the slopes (d_a / d_n, d_b / d_n) over the cone of half angle `theta` around d0: the cone maps
// to a conic region of the slope plane whose extremes lie on the image of the cone's rim, sampled densely
AxisChoice slope_box(const double d0[3], double theta, int n) {
    AxisChoice c = {};
    c.n = n; c.a = (n + 1) % 3; c.b = (n + 2) % 3; c.ok = false;
    const double alpha = std::acos(std::min(1.0, std::fabs(d0[n])));
    if (alpha + theta > 1.40) return c;                 // the cone comes within 10 degrees of grazing this plane
    c.sgn = d0[n] < 0.0 ? -1.0 : 1.0;
    // orthonormal pair perpendicular to d0
    double u[3] = {0, 0, 0}, w[3];
    int small = 0;
    for (int k = 1; k < 3; ++k) if (std::fabs(d0[k]) < std::fabs(d0[small])) small = k;
    u[small] = 1.0;
    double dotu = d0[small];
    double len = 0.0;
    for (int k = 0; k < 3; ++k) { u[k] -= dotu * d0[k]; len += u[k] * u[k]; }
    len = std::sqrt(len);
    for (int k = 0; k < 3; ++k) u[k] /= len;
    w[0] = d0[1] * u[2] - d0[2] * u[1]; w[1] = d0[2] * u[0] - d0[0] * u[2]; w[2] = d0[0] * u[1] - d0[1] * u[0];
    double lo_a = d0[c.a] / d0[n], hi_a = lo_a, lo_b = d0[c.b] / d0[n], hi_b = lo_b;
    const int kSamples = 4096;
    const double ct = std::cos(theta), st = std::sin(theta);
    for (int k = 0; k < kSamples; ++k) {
        const double phi = 2.0 * M_PI * k / kSamples;
        double d[3];
        for (int j = 0; j < 3; ++j) d[j] = ct * d0[j] + st * (std::cos(phi) * u[j] + std::sin(phi) * w[j]);
        const double ta = d[c.a] / d[n], tb = d[c.b] / d[n];
        lo_a = std::min(lo_a, ta); hi_a = std::max(hi_a, ta);
        lo_b = std::min(lo_b, tb); hi_b = std::max(hi_b, tb);
    }
    // the rim is sampled every 1.5 mrad of phi: an extreme between two samples is missed by < 1.2e-6 of the range
    const double pad_a = 1e-3 * (hi_a - lo_a) + 1e-9, pad_b = 1e-3 * (hi_b - lo_b) + 1e-9;
    c.t0a = 0.5 * (lo_a + hi_a); c.da = 0.5 * (hi_a - lo_a) + pad_a;
    c.t0b = 0.5 * (lo_b + hi_b); c.db = 0.5 * (hi_b - lo_b) + pad_b;
    c.smax = std::max(std::fabs(c.t0a) + c.da, std::fabs(c.t0b) + c.db);
    c.ok = true;
    return c;
}

struct Rect { double a0, a1, b0, b1; };

// shadow of the box [lo, hi] on the plane x_n = h0 for slopes in the (widened) box of `c`
Rect shadow(const AxisChoice &c, const double *lo, const double *hi, double pad_slope, double margin) {
    const double h_lo = lo[c.n] - c.h0, h_hi = hi[c.n] - c.h0;
    Rect r;
    double sweep_lo[2], sweep_hi[2];
    const double t0[2] = {c.t0a, c.t0b}, dt[2] = {c.da + pad_slope, c.db + pad_slope};
    for (int k = 0; k < 2; ++k) {
        const double ta = t0[k] - dt[k], tb = t0[k] + dt[k];
        const double p[4] = {h_lo * ta, h_lo * tb, h_hi * ta, h_hi * tb};
        sweep_lo[k] = *std::min_element(p, p + 4);
        sweep_hi[k] = *std::max_element(p, p + 4);
    }
    r.a0 = lo[c.a] - sweep_hi[0] - margin; r.a1 = hi[c.a] - sweep_lo[0] + margin;
    r.b0 = lo[c.b] - sweep_hi[1] - margin; r.b1 = hi[c.b] - sweep_lo[1] + margin;
    return r;
}

struct P2 { double x, y; };

// convex hull (counter-clockwise, monotone chain) of n points; returns the number of hull points written to `out`
int convex_hull(P2 *pts, int n, P2 *out) {
    std::sort(pts, pts + n, [](const P2 &a, const P2 &b) { return a.x < b.x || (a.x == b.x && a.y < b.y); });
    auto cross = [](const P2 &o, const P2 &a, const P2 &b) { return (a.x - o.x) * (b.y - o.y) - (a.y - o.y) * (b.x - o.x); };
    int k = 0;
    for (int i = 0; i < n; ++i) { while (k >= 2 && cross(out[k - 2], out[k - 1], pts[i]) <= 0) --k; out[k++] = pts[i]; }
    for (int i = n - 2, t = k + 1; i >= 0; --i) { while (k >= t && cross(out[k - 2], out[k - 1], pts[i]) <= 0) --k; out[k++] = pts[i]; }
    return k > 1 ? k - 1 : k;
}

// Shadow of a planar quad (4 corners, global coordinates): the hull of every corner's key under the four corners of
// the slope box -- the key is affine in the point for a fixed slope and affine in the slope for a fixed point, so the
// keys of all (point of the quad, slope of the box) pairs lie in the hull of these 16.
int quad_shadow(const AxisChoice &c, const double *quad, double pad_slope, P2 *hull) {
    P2 pts[16];
    int n = 0;
    const double ta[2] = {c.t0a - c.da - pad_slope, c.t0a + c.da + pad_slope};
    const double tb[2] = {c.t0b - c.db - pad_slope, c.t0b + c.db + pad_slope};
    for (int k = 0; k < 4; ++k) {
        const double *p = quad + 3 * k;
        const double h = p[c.n] - c.h0;
        for (int i = 0; i < 2; ++i)
            for (int j = 0; j < 2; ++j) { pts[n].x = p[c.a] - h * ta[i]; pts[n].y = p[c.b] - h * tb[j]; ++n; }
    }
    return convex_hull(pts, n, hull);
}

void free_grid(TrcScene *sc) {
    cudaFree(sc->d_grid_bits); cudaFree(sc->d_grid_cells); cudaFree(sc->d_grid_list);
    sc->d_grid_bits = sc->d_grid_cells = nullptr;
    sc->d_grid_list = nullptr;
    sc->has_grid = false;
    sc->grid = TrcShadowGrid();
    for (int k = 0; k < 3; ++k) sc->grid_occ[k] = 0;       // occupancy depends on the grid's shared-memory share
}

}  // namespace

void trc_scene_free_shadow_grid(TrcScene *sc) { free_grid(sc); }

extern "C" int trc_scene_set_shadow_grid(TrcScene *sc, int32_t n_boxes, const double *box_lo, const double *box_hi,
                                         const double *quads, const int32_t *surface_ids, const double *direction,
                                         double half_angle, int32_t *enabled) {
    if (!sc || n_boxes < 0 || (n_boxes > 0 && (!box_lo || !box_hi || !surface_ids)) || !enabled)
        return fail(TRC_ERR_BAD_ARGUMENT, "trc_scene_set_shadow_grid: bad argument%s");
    *enabled = 0;
    TRC_CUDA(cudaSetDevice(sc->device));
    TRC_CUDA(cudaDeviceSynchronize());           // a trace in flight may still read the grid being replaced
    free_grid(sc);
    if (n_boxes == 0 || !direction || !(half_angle >= 0.0) || half_angle > 0.5) return TRC_OK;
    double d0[3], len = 0.0;
    for (int k = 0; k < 3; ++k) len += direction[k] * direction[k];
    len = std::sqrt(len);
    if (!(len > 0.0) || !std::isfinite(len)) return TRC_OK;
    for (int k = 0; k < 3; ++k) d0[k] = direction[k] / len;
    const int max_words = trc_walk_grid_word_budget(sc);
    if (max_words < 64) return TRC_OK;           // no shared memory to spare next to this scene's tree cache
    for (int i = 0; i < n_boxes; ++i)
        if (surface_ids[i] < 0 || surface_ids[i] >= (1 << 22))
            return fail(TRC_ERR_BAD_ARGUMENT, "trc_scene_set_shadow_grid: surface index out of range%s");

    // scene scale: the kernel's float32 keys are good to rmax * (1 + 2 smax) * 2^-19 for |v| <= rmax
    double extent = 0.0;
    for (int i = 0; i < 3 * n_boxes; ++i) {
        if (!std::isfinite(box_lo[i]) || !std::isfinite(box_hi[i])) return TRC_OK;
        extent = std::max(extent, std::max(std::fabs(box_lo[i]), std::fabs(box_hi[i])));
    }
    const double rmax = 8.0 * std::max(extent, 1e-3);

    AxisChoice best = {};
    best.ok = false;
    std::vector<double> heights(n_boxes);
    for (int n = 0; n < 3; ++n) {
        AxisChoice c = slope_box(d0, half_angle, n);
        if (!c.ok) continue;
        for (int i = 0; i < n_boxes; ++i) heights[i] = 0.5 * (box_lo[3 * i + n] + box_hi[3 * i + n]);
        std::nth_element(heights.begin(), heights.begin() + n_boxes / 2, heights.end());
        c.h0 = heights[n_boxes / 2];           // median height: most boxes sweep a short shadow
        c.cost = 0.0;
        for (int i = 0; i < n_boxes; ++i) {
            Rect r = shadow(c, box_lo + 3 * i, box_hi + 3 * i, 0.0, 0.0);
            c.cost += (r.a1 - r.a0) * (r.b1 - r.b0);
        }
        if (!best.ok || c.cost < best.cost) best = c;
    }
    if (!best.ok) return TRC_OK;
    const AxisChoice &c = best;
    const double pad_slope = 1e-5 * (1.0 + c.smax);                 // float32 slope of a ray at the edge of the box
    const double key_err = rmax * (1.0 + 2.0 * c.smax) * std::ldexp(1.0, -18);

    std::vector<Rect> rects(n_boxes);
    double ga0 = 1e300, ga1 = -1e300, gb0 = 1e300, gb1 = -1e300;
    for (int i = 0; i < n_boxes; ++i) {
        rects[i] = shadow(c, box_lo + 3 * i, box_hi + 3 * i, pad_slope, key_err);
        ga0 = std::min(ga0, rects[i].a0); ga1 = std::max(ga1, rects[i].a1);
        gb0 = std::min(gb0, rects[i].b0); gb1 = std::max(gb1, rects[i].b1);
    }
    double wa = std::max(ga1 - ga0, 1e-6), wb = std::max(gb1 - gb0, 1e-6);
    const double n_cells_max = 32.0 * max_words;
    double cell = std::sqrt(wa * wb / n_cells_max);
    int nx, ny;
    for (;;) {
        nx = (int)std::ceil(wa / cell) + 1; ny = (int)std::ceil(wb / cell) + 1;
        if ((double)nx * ny <= n_cells_max) break;
        cell *= 1.02;
    }
    if (nx > 32000 || ny > 32000) return TRC_OK;
    const double cell_pad = 2e-3 * cell;                            // rounding of (key - origin) / cell in float32
    const int64_t n_cells = (int64_t)nx * ny;
    std::vector<uint32_t> count(n_cells, 0u);
    const double m = key_err + cell_pad;                            // outward margin of a hull shadow
    // Cells of surface i's shadow: the rectangle of its box, or -- a planar surface whose corners are known -- the hull
    // of its corners' keys, rasterised row by row (the x extent of the hull over the widened row band).
    auto for_each_cell = [&](int i, auto &&fn) {
        const double *quad = quads ? quads + 12 * (size_t)i : nullptr;
        if (!quad || !std::isfinite(quad[0])) {
            int a0 = (int)std::floor((rects[i].a0 - cell_pad - ga0) / cell), a1 = (int)std::floor((rects[i].a1 + cell_pad - ga0) / cell);
            int b0 = (int)std::floor((rects[i].b0 - cell_pad - gb0) / cell), b1 = (int)std::floor((rects[i].b1 + cell_pad - gb0) / cell);
            a0 = std::max(a0, 0); a1 = std::min(a1, nx - 1); b0 = std::max(b0, 0); b1 = std::min(b1, ny - 1);
            for (int b = b0; b <= b1; ++b)
                for (int a = a0; a <= a1; ++a) fn((int64_t)b * nx + a);
            return;
        }
        P2 hull[17];
        const int nh = quad_shadow(c, quad, pad_slope, hull);
        double ylo = 1e300, yhi = -1e300;
        for (int k = 0; k < nh; ++k) { ylo = std::min(ylo, hull[k].y); yhi = std::max(yhi, hull[k].y); }
        int b0 = std::max(0, (int)std::floor((ylo - m - gb0) / cell)), b1 = std::min(ny - 1, (int)std::floor((yhi + m - gb0) / cell));
        for (int b = b0; b <= b1; ++b) {
            const double ya = gb0 + b * cell - m, yb = gb0 + (b + 1) * cell + m;
            double xlo = 1e300, xhi = -1e300;
            for (int k = 0; k < nh; ++k) {
                const P2 p = hull[k], q = hull[(k + 1) % nh];
                // the part of edge p-q inside the band ya <= y <= yb
                double t0 = 0.0, t1 = 1.0;
                const double dy = q.y - p.y;
                if (dy == 0.0) { if (p.y < ya || p.y > yb) continue; }
                else {
                    double ta_ = (ya - p.y) / dy, tb_ = (yb - p.y) / dy;
                    if (ta_ > tb_) std::swap(ta_, tb_);
                    t0 = std::max(t0, ta_); t1 = std::min(t1, tb_);
                    if (t0 > t1) continue;
                }
                const double x0 = p.x + t0 * (q.x - p.x), x1 = p.x + t1 * (q.x - p.x);
                xlo = std::min(xlo, std::min(x0, x1)); xhi = std::max(xhi, std::max(x0, x1));
            }
            if (xlo > xhi) continue;
            int a0 = std::max(0, (int)std::floor((xlo - m - ga0) / cell)), a1 = std::min(nx - 1, (int)std::floor((xhi + m - ga0) / cell));
            for (int a = a0; a <= a1; ++a) fn((int64_t)b * nx + a);
        }
    };
    int64_t total = 0;
    for (int i = 0; i < n_boxes; ++i) for_each_cell(i, [&](int64_t k) { ++count[k]; ++total; });
    if (total >= (1 << 24)) return TRC_OK;                          // list offsets are 24 bits
    std::vector<uint32_t> cells(n_cells, 0u), bits((n_cells + 31) / 32, 0u);
    std::vector<uint32_t> first(n_cells, 0u);
    uint32_t run = 0;
    for (int64_t k = 0; k < n_cells; ++k) {
        if (count[k] > 255u) return TRC_OK;                         // too many surfaces per cell for this layout
        if (count[k]) bits[k >> 5] |= 1u << (k & 31);
        if (count[k] > 1u) { first[k] = run; run += count[k]; }
    }
    std::vector<int32_t> list(std::max<uint32_t>(run, 1u), 0);
    std::vector<uint32_t> fill(n_cells, 0u);
    for (int i = 0; i < n_boxes; ++i)
        for_each_cell(i, [&](int64_t k) {
            if (count[k] == 1u) cells[k] = ((uint32_t)surface_ids[i] << 8) | 1u;
            else {
                list[first[k] + fill[k]++] = surface_ids[i];
                cells[k] = (first[k] << 8) | count[k];
            }
        });
    TRC_CUDA(cudaMalloc(&sc->d_grid_bits, bits.size() * sizeof(uint32_t)));
    TRC_CUDA(cudaMalloc(&sc->d_grid_cells, cells.size() * sizeof(uint32_t)));
    TRC_CUDA(cudaMalloc(&sc->d_grid_list, list.size() * sizeof(int32_t)));
    TRC_CUDA(cudaMemcpy(sc->d_grid_bits, bits.data(), bits.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
    TRC_CUDA(cudaMemcpy(sc->d_grid_cells, cells.data(), cells.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
    TRC_CUDA(cudaMemcpy(sc->d_grid_list, list.data(), list.size() * sizeof(int32_t), cudaMemcpyHostToDevice));
    TrcShadowGrid &g = sc->grid;
    g.bits = sc->d_grid_bits; g.cells = sc->d_grid_cells; g.list = sc->d_grid_list;
    g.n_words = (int)bits.size();
    g.axis_n = c.n; g.axis_a = c.a; g.axis_b = c.b;
    g.nx = nx; g.ny = ny;
    g.h0 = (float)c.h0; g.sgn = (float)c.sgn;
    // the kernel tests float32 copies of the slope box: shrink nothing, the shadows were swept with pad_slope more
    g.t0a = (float)c.t0a; g.t0b = (float)c.t0b; g.da = (float)c.da; g.db = (float)c.db;
    g.ga0 = (float)ga0; g.gb0 = (float)gb0; g.inv_ca = (float)(1.0 / cell); g.inv_cb = (float)(1.0 / cell);
    g.rmax = (float)rmax;
    sc->has_grid = true;
    *enabled = 1;
    return TRC_OK;
}
