// Ray sources on device: pillbox (disk_bundle, sources.py:175-239) and Buie
// sunshape (buie_sunshape / Buie_directions_sampling, sources.py:330-464).
// One thread per ray, Philox keyed by (seed, global ray index) so a bundle is
// reproducible from its seed alone and independent of how it is sharded.
#include "common.cuh"
#include "geom.cuh"
#include "optics.cuh"
#include "philox.cuh"

using namespace trc;

namespace {

struct Mat3 { double m[9]; };
struct Vec3h { double v[3]; };

__device__ __forceinline__ V3 mat_apply(const Mat3 &r, V3 a) {
    return mk(r.m[0] * a.x + r.m[1] * a.y + r.m[2] * a.z,
              r.m[3] * a.x + r.m[4] * a.y + r.m[5] * a.z,
              r.m[6] * a.x + r.m[7] * a.y + r.m[8] * a.z);
}

__device__ __forceinline__ void store_ray(double *vert, double *dir, double *energy, int64_t ld, int64_t i,
                                          V3 p, V3 d, double e) {
    vert[i] = p.x; vert[ld + i] = p.y; vert[2 * ld + i] = p.z;
    dir[i] = d.x; dir[ld + i] = d.y; dir[2 * ld + i] = d.z;
    energy[i] = e;
}

__global__ void __launch_bounds__(256)
pillbox_disk_kernel(int64_t n, int64_t first, uint64_t seed, Vec3h center, Mat3 rot_dir, Mat3 rot_pos,
                    double r_in, double r_out, double ang0, double ang1, double ang_range, double ray_energy,
                    double rect_hx, double rect_hy,
                    double *__restrict__ vert, double *__restrict__ dir, double *__restrict__ energy, int64_t ld) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint64_t id = (uint64_t)(first + i);
    double u0, u1, u2, u3;
    uniform2(seed, id, 0x50494c4cu, 0, u0, u1);   // direction draws
    uniform2(seed, id, 0x50494c4cu, 1, u2, u3);   // position draws
    V3 a = pillbox_direction(2.0 * M_PI * u0, u1, ang_range);
    V3 d = mat_apply(rot_dir, a);
    V3 pl;
    if (rect_hx > 0.0) {          // rectangular source (rect_bundle, sources.py:241-264): uniform in +-hx, +-hy
        pl = mk((2.0 * u2 - 1.0) * rect_hx, (2.0 * u3 - 1.0) * rect_hy, 0.0);
    } else {
        double th = ang0 + (ang1 - ang0) * u3;
        double rs = sqrt(r_in * r_in + u2 * (r_out * r_out - r_in * r_in));
        pl = mk(rs * cos(th), rs * sin(th), 0.0);
    }
    V3 p = mat_apply(rot_pos, pl);
    store_ray(vert, dir, energy, ld, i, mk(p.x + center.v[0], p.y + center.v[1], p.z + center.v[2]), d, ray_energy);
}

__global__ void __launch_bounds__(256)
buie_disk_kernel(int64_t n, int64_t first, uint64_t seed, const TrcBuieTables *__restrict__ tb,
                 Vec3h center, Mat3 rot_dir, Mat3 rot_pos, double radius, double ray_energy,
                 double rect_hx, double rect_hy,
                 double *__restrict__ vert, double *__restrict__ dir, double *__restrict__ energy, int64_t ld) {
    __shared__ double s_cdf[212], s_th[212], s_w[212];
    const int nseg = tb->n_seg;
    for (int k = threadIdx.x; k <= nseg; k += blockDim.x) { s_cdf[k] = tb->cdf[k]; s_th[k] = tb->theta[k]; s_w[k] = tb->w[k]; }
    __syncthreads();
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint64_t id = (uint64_t)(first + i);
    double xv1, uphi, rth, uxi;
    uniform2(seed, id, 0x42554945u, 0, xv1, uphi);   // start position
    uniform2(seed, id, 0x42554945u, 1, rth, uxi);    // polar / azimuth of the direction
    V3 p;
    if (rect_hx > 0.0) {          // rect_buie_sunshape (sources.py:487-488): width * (u - 0.5), height * (u' - 0.5)
        p = mat_apply(rot_pos, mk((2.0 * xv1 - 1.0) * rect_hx, (2.0 * uphi - 1.0) * rect_hy, 0.0));
    } else {
        double phiv = 2.0 * M_PI * uphi;
        double rs = radius * sqrt(xv1);
        p = mat_apply(rot_pos, mk(rs * cos(phiv), rs * sin(phiv), 0.0));
    }
    double theta = 0.0;
    if (rth >= s_cdf[nseg]) {
        if (tb->has_aureole) {
            double g2 = tb->gamma + 2.0;
            double k1 = g2 / (pow(10.0, 3.0 * tb->gamma) * exp(tb->kappa)) * tb->sum_integ_dni - pow(tb->theta_dni, g2);
            theta = pow((rth - 1.0) * k1 + rth * pow(tb->theta_tot, g2), 1.0 / g2);
        }
    } else {
        int lo = 0, hi = nseg;            // cdf[lo] <= rth < cdf[hi]
        while (hi - lo > 1) {
            int mid = (lo + hi) >> 1;
            if (s_cdf[mid] <= rth) lo = mid; else hi = mid;
        }
        double A = s_w[lo], B = s_w[lo + 1];
        double t0 = s_th[lo], t1 = s_th[lo + 1];
        double C = 2.0 * tb->sum_integ_dni * (rth - s_cdf[lo]) * (t1 - t0);
        double k = (t0 - t1) * A;
        theta = -(-A * t1 + B * t0 + sqrt(k * k + C * (B - A))) / (A - B);
    }
    double xi1 = 2.0 * M_PI * uxi;
    double st = sin(theta);
    V3 d = mat_apply(rot_dir, mk(cos(xi1) * st, sin(xi1) * st, cos(theta)));
    store_ray(vert, dir, energy, ld, i, mk(p.x + center.v[0], p.y + center.v[1], p.z + center.v[2]), d, ray_energy);
}

// The Buie tables are tiny (5 KB); keep one device copy per call site alive in
// a small cache keyed by content so repeated source calls do not re-upload.
struct BuieCache { TrcBuieTables host; TrcBuieTables *dev; bool valid; };
constexpr int kMaxDevices = 64;
thread_local BuieCache g_buie_dev[kMaxDevices];     // one cached device copy per CUDA device

}  // namespace

static int source_pillbox(int64_t n, int64_t first_ray, uint64_t seed, const double *center3,
                          const double *rot_dir9, const double *rot_pos9, double radius_in,
                          double radius, double ang0, double ang1, double ang_range,
                          double ray_energy, double rect_hx, double rect_hy, double *d_vert, double *d_dir,
                          double *d_energy, int64_t ld, void *stream) {
    if (n < 0 || !center3 || !rot_dir9 || !rot_pos9 || (n > 0 && (!d_vert || !d_dir || !d_energy)) || ld < n)
        return fail(TRC_ERR_BAD_ARGUMENT, "trc_source_pillbox: bad argument%s");
    if (n == 0) return TRC_OK;
    Vec3h c; Mat3 rd, rp;
    for (int k = 0; k < 3; ++k) c.v[k] = center3[k];
    for (int k = 0; k < 9; ++k) { rd.m[k] = rot_dir9[k]; rp.m[k] = rot_pos9[k]; }
    pillbox_disk_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        n, first_ray, seed, c, rd, rp, radius_in, radius, ang0, ang1, ang_range, ray_energy, rect_hx, rect_hy,
        d_vert, d_dir, d_energy, ld);
    return check_launch("source_pillbox");
}

extern "C" int trc_source_pillbox_disk(int64_t n, int64_t first_ray, uint64_t seed, const double *center3,
                                       const double *rot_dir9, const double *rot_pos9, double radius_in,
                                       double radius, double ang0, double ang1, double ang_range,
                                       double ray_energy, double *d_vert, double *d_dir, double *d_energy,
                                       int64_t ld, void *stream) {
    return source_pillbox(n, first_ray, seed, center3, rot_dir9, rot_pos9, radius_in, radius, ang0, ang1, ang_range,
                          ray_energy, 0.0, 0.0, d_vert, d_dir, d_energy, ld, stream);
}

extern "C" int trc_source_pillbox_rect(int64_t n, int64_t first_ray, uint64_t seed, const double *center3,
                                       const double *rot_dir9, const double *rot_pos9, double half_x,
                                       double half_y, double ang_range, double ray_energy, double *d_vert,
                                       double *d_dir, double *d_energy, int64_t ld, void *stream) {
    if (!(half_x > 0.0) || !(half_y > 0.0))
        return fail(TRC_ERR_BAD_ARGUMENT, "trc_source_pillbox_rect: half extents must be positive%s");
    return source_pillbox(n, first_ray, seed, center3, rot_dir9, rot_pos9, 0.0, 0.0, 0.0, 0.0, ang_range, ray_energy,
                          half_x, half_y, d_vert, d_dir, d_energy, ld, stream);
}

static int source_buie(int64_t n, int64_t first_ray, uint64_t seed, const TrcBuieTables *tables,
                       const double *center3, const double *rot_dir9, const double *rot_pos9,
                       double radius, double ray_energy, double rect_hx, double rect_hy, double *d_vert,
                       double *d_dir, double *d_energy, int64_t ld, void *stream) {
    if (n < 0 || !tables || tables->n_seg < 1 || tables->n_seg > 210 || !center3 || !rot_dir9 || !rot_pos9 ||
        (n > 0 && (!d_vert || !d_dir || !d_energy)) || ld < n)
        return fail(TRC_ERR_BAD_ARGUMENT, "trc_source_buie: bad argument%s");
    if (n == 0) return TRC_OK;
    cudaStream_t st = (cudaStream_t)stream;
    int dev = 0;
    TRC_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) return fail(TRC_ERR_BAD_ARGUMENT, "trc_source_buie: device index out of range%s");
    BuieCache &g_buie = g_buie_dev[dev];
    if (!g_buie.valid || memcmp(&g_buie.host, tables, sizeof(TrcBuieTables)) != 0) {
        if (!g_buie.dev) TRC_CUDA(cudaMalloc(&g_buie.dev, sizeof(TrcBuieTables)));
        // synchronous copy: the cached device table may still be read by an
        // earlier launch on another stream
        TRC_CUDA(cudaDeviceSynchronize());
        TRC_CUDA(cudaMemcpy(g_buie.dev, tables, sizeof(TrcBuieTables), cudaMemcpyHostToDevice));
        g_buie.host = *tables;
        g_buie.valid = true;
    }
    Vec3h c; Mat3 rd, rp;
    for (int k = 0; k < 3; ++k) c.v[k] = center3[k];
    for (int k = 0; k < 9; ++k) { rd.m[k] = rot_dir9[k]; rp.m[k] = rot_pos9[k]; }
    buie_disk_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(n, first_ray, seed, g_buie.dev, c, rd, rp, radius,
                                                                  ray_energy, rect_hx, rect_hy, d_vert, d_dir,
                                                                  d_energy, ld);
    return check_launch("source_buie");
}

extern "C" int trc_source_buie_disk(int64_t n, int64_t first_ray, uint64_t seed, const TrcBuieTables *tables,
                                    const double *center3, const double *rot_dir9, const double *rot_pos9,
                                    double radius, double ray_energy, double *d_vert, double *d_dir,
                                    double *d_energy, int64_t ld, void *stream) {
    return source_buie(n, first_ray, seed, tables, center3, rot_dir9, rot_pos9, radius, ray_energy, 0.0, 0.0, d_vert,
                       d_dir, d_energy, ld, stream);
}

extern "C" int trc_source_buie_rect(int64_t n, int64_t first_ray, uint64_t seed, const TrcBuieTables *tables,
                                    const double *center3, const double *rot_dir9, const double *rot_pos9,
                                    double half_x, double half_y, double ray_energy, double *d_vert,
                                    double *d_dir, double *d_energy, int64_t ld, void *stream) {
    if (!(half_x > 0.0) || !(half_y > 0.0))
        return fail(TRC_ERR_BAD_ARGUMENT, "trc_source_buie_rect: half extents must be positive%s");
    return source_buie(n, first_ray, seed, tables, center3, rot_dir9, rot_pos9, 0.0, ray_energy, half_x, half_y,
                       d_vert, d_dir, d_energy, ld, stream);
}
