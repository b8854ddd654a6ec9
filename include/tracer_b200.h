/*
 * tracer_b200.h -- C ABI of libtracer_b200.so (sm_100a).
 *
 * This is the drop-in boundary for the Monte-Carlo bounce loop of
 * casselineau/Tracer.  The reference has no FFI (it is pure NumPy), so each
 * entry point below names the reference *Python* interface it replaces
 * (paths relative to the reference tree).  All pointers named `d_*` are DEVICE
 * pointers owned by the caller (torch CUDA tensors in the Python host layer);
 * the library never allocates, frees or retains them.  Scene handles own small
 * device tables and are freed by trc_scene_destroy().
 *
 * Ray bundles are SoA float64 exactly as the reference's RayBundle
 * (tracer/ray_bundle.py:6-63): vertices (3,n) and directions (3,n) row-major
 * with a row stride `ld` (elements), energy (n,), parents (n,) int64.
 *
 * Every function returns 0 on success or a negative TrcStatus; the message is
 * available from trc_last_error() (thread-local).  No C++ exception crosses
 * this boundary.  All work is enqueued on the given CUDA stream
 * (a cudaStream_t passed as void*; NULL = legacy default stream).
 */
#ifndef TRACER_B200_H
#define TRACER_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TRC_ABI_VERSION 15

typedef enum TrcStatus {
    TRC_OK = 0,
    TRC_ERR_BAD_ARGUMENT = -1,
    TRC_ERR_UNSUPPORTED_KIND = -2,
    TRC_ERR_CUDA = -3,
    TRC_ERR_CAPACITY = -4
} TrcStatus;

/* Geometry-manager kinds: one per reference GeometryManager subclass on the
 * hot path. */
typedef enum TrcGeomKind {
    TRC_G_FLAT_INFINITE = 0,   /* FlatGeometryManager        flat_surface.py:11   */
    TRC_G_RECT_PLATE = 1,      /* RectPlateGM                flat_surface.py:181  gp: hx, hy */
    TRC_G_RECT_EXTRUDED = 2,   /* ExtrudedRectPlateGM        flat_surface.py:253  gp: hx, hy, cx, cy, ehx, ehy */
    TRC_G_ROUND_PLATE = 3,     /* RoundPlateGM / StraightCut flat_surface.py:457,548  gp: Re, Ri(<0 none), x_cut(+inf none) */
    TRC_G_TRIANGLE = 4,        /* TriangularFace             triangular_face.py:12 gp: v0x v0y v0z v1x v1y v1z */
    TRC_G_PARABOLOID = 5,      /* Paraboloid                 paraboloid.py:11     gp: a, b */
    TRC_G_PARAB_DISH = 6,      /* ParabolicDishGM            paraboloid.py:71     gp: a, b, h */
    TRC_G_PARAB_HEX = 7,       /* HexagonalParabolicDishGM   paraboloid.py:174    gp: a, b, R */
    TRC_G_PARAB_RECT = 8,      /* RectangularParabolicDishGM paraboloid.py:225    gp: a, b, hw, hh */
    TRC_G_PARAB_CYL = 9,       /* ParabolicCylinder          paraboloid.py:328    gp: a */
    TRC_G_PARAB_TROUGH = 10,   /* ParabolicTroughGM          paraboloid.py:386    gp: a, hl, h */
    TRC_G_SPHERE = 11,         /* SphericalGM                sphere_surface.py:9  gp: r */
    TRC_G_HEMISPHERE = 12,     /* HemisphereGM               sphere_surface.py:117 gp: r */
    TRC_G_SPHERE_RECT = 13,    /* SphericalRectFacet         sphere_surface.py:206 gp: r, lx/2, ly/2 */
    TRC_G_CONE_INFINITE = 14,  /* InfiniteCone               cone.py:7            gp: c, a */
    TRC_G_CONE_FINITE = 15,    /* FiniteCone                 cone.py:74           gp: c, a(=0), h */
    TRC_G_FRUSTUM = 16,        /* ConicalFrustum             cone.py:261          gp: c, a, zmin, zmax */
    TRC_G_FRUSTUM_RECT = 17,   /* RectCutConicalFrustum      cone.py:356          gp: c, a, zmin, zmax, hw, hh */
    TRC_G_CYL_INFINITE = 18,   /* InfiniteCylinder           cylinder.py:12       gp: R */
    TRC_G_CYL_FINITE = 19,     /* FiniteCylinder             cylinder.py:59       gp: R, half_h, ang0, ang1 */
    TRC_G_CYL_RECT = 20,       /* RectCutCylinder            cylinder.py:161      gp: R, half_h, hw, hh */
    TRC_G_QUADRIC_FLAT = 21,   /* FlatQuadricSurfaceGM       quadratic_surface.py:4   gp: a b c d e f */
    TRC_G_QUADRIC_RECT = 22,   /* RectFlatQuadricSurfaceGM   quadratic_surface.py:63  gp: a b c d e f hw hh */
    TRC_G_SPHERE_CUT = 23,     /* CutSphereGM sphere_surface.py:168 bounded by a BoundarySphere (boundary_shape.py:89-111)
                                  gp: r, cx, cy, cz (bound centre, global), R_bound, 0
                                  or by a BoundaryPlane (:151-165) gp: r, a, b, c, d, 1 with (a, b, c, d) = row 2 of the
                                  inverse of (surface frame . bound transform): valid where a x + b y + c z + d >= 0 */
    TRC_G_ELLIPSOID = 24,      /* EllipsoidGM                ellipsoid.py:5,63    gp: 1/a^2, 1/b^2, 1/c^2, xlo, xhi, ylo,
                                  yhi, zlo, zhi (-inf / +inf where there is no limit) */
    TRC_G_RECT_PERFORATED = 25,/* PerforatedRectPlateGM      flat_surface.py:357  gp: hx, hy, pool offset, n_holes;
                                  pool: (cx, cy, r) per hole */
    TRC_G_POLYGON = 26,        /* FlatSimplePolygonGM / PerforatedPolygonGM  polygon.py:7,139
                                  gp: pool offset, n_vertices, n_holes; pool: xs[nv+1], ys[nv+1] (closed
                                  profile), then (cx, cy, r) per hole */
    TRC_G_KIND_COUNT = 27
} TrcGeomKind;
/* "pool" = the scene's float64 table pool (trc_scene_create `tables`), shared with the optics
 * tables; offsets are counted in doubles and stored in gp as exact small integers. */

/* Optics kinds: the base optics class behind the accountant wrapper
 * (optics_callables.py). */
typedef enum TrcOpticsKind {
    TRC_O_TRANSPARENT = 0,        /* Transparent         optics_callables.py:93  */
    TRC_O_REFLECTIVE = 1,         /* Reflective          :116  op: abs */
    TRC_O_REAL_REFLECTIVE = 2,    /* RealReflective      :214  op: abs, sigma, bi_var */
    TRC_O_LAMBERTIAN = 3,         /* Lambertian          :143  op: abs, ang_range */
    TRC_O_LAMBERTIAN_SPECULAR = 4,/* LambertianSpecular  :553  op: abs, specularity */
    TRC_O_LAMB_DIRECTIONAL = 5,   /* Lambertian_directional_axisymmetric_piecewise :331
                                     table = (thetas, abs_th); always diffuse */
    TRC_O_REFLECTIVE_SPECTRAL = 6,/* Reflective_spectral :178  table = (wavelengths, absorptances) */
    TRC_O_LAMBSPEC_DIRECTIONAL = 7,/* LambertianSpecular_directional_axisymmetric_piecewise :427
                                     op: specularity; table = (thetas, abs_th) */
    TRC_O_LAMB_DIRECTIONAL_SPECTRAL = 8, /* Lambertian_directional_axisymmetric_piecewise_spectral :359
                                     op[1] = n_wavelengths; table = thetas[table_len],
                                     wavelengths[n_wl], absorptance[table_len*n_wl] (bilinear) */
    TRC_O_LAMB_PIECEWISE_SPEC = 9,/* Lambertian_piecewise_Specular_directional_axisymmetric_piecewise :457
                                     table = thetas[n], abs_th[n], spec_th[n]: absorptance AND specularity
                                     interpolated at the incidence angle */
    TRC_O_LAMBERTIAN_ABSORBANT = 10,/* LambertianAbsorbant :891 (+ Absorbant.attenuate :874-889, imposed
                                     attenuation coefficient): op: absorptivity, ang_range, a_c, scaling.
                                     E_att = E exp(-a_c |hit - parent vertex| scaling), E' = E_att (1 - abs) */
    TRC_O_PERIODIC = 11,          /* PeriodicBoundary :690-723: op: period.  TWO children per hit: a
                                     zero-energy stub at the hit, then the live ray moved by period * normal
                                     (block A, block B of the surface's segment; see TrcOpticsArgs) */
    TRC_O_FRESNEL_CONDUCTOR = 12, /* FresnelConductorHomogenous :1523-1558 with optics.fresnel_conductor /
                                     fresnel_to_attenuating (optics.py:41-81): op: n1; table = wavelengths[n],
                                     n2[n], k2[n] (complex index of the conductor, linear in wavelength);
                                     specular, E' = E (R_p + R_s) / 2; NaN outside the table's range */
    TRC_O_REFRACTIVE = 13,        /* RefractiveHomogenous :1186-1296 (+ Refractive :726-858, optics.refractions /
                                     optics.fresnel optics.py:13-39,159-192) and RefractiveTransmissiveHomogenous
                                     :1323-1345 (Absorbant.attenuate :874-889 with imposed coefficients).
                                     op: n_a, n_b (the two indices, _ref_idxs), single_ray (1/0), sigma (< 0: None),
                                     n_att (0 none, 1 or 2 imposed attenuation coefficients), a_c[0], a_c[1], scaling.
                                     Needs the parent bundle's ref_index column (d_ref_index).  Per hit, with
                                     m1 = the ray's index and m2 = the other one: TWO child slots, block A = the
                                     reflected ray, block B = the refracted ray (as TRC_O_PERIODIC lays them out).
                                     single_ray: the ray goes to A when draw <= R (R = 1 under total internal
                                     reflection) else to B, and the other slot is VOID; otherwise A gets E R and
                                     B gets E (1 - R), B void under total internal reflection.  Void slots carry
                                     d_cull = 2 and are not children: the caller drops them. */
    TRC_O_KIND_COUNT = 14
} TrcOpticsKind;

/* Surface flag bits. */
#define TRC_F_ONE_SIDED   1u   /* OneSidedReflective / OneSidedRealReflective (:195,:492): E'=0 where d.up>0 */
#define TRC_F_UNBOUNDED   2u   /* no finite AABB: tested for every ray, outside the BVH */
#define TRC_F_POLY        8u   /* Lambertian_directional_axisymmetric_piecewise_Polychromatic (:393-425): the optics
                                  kernel treats the surface as a Lambertian that absorbs nothing; trc_spectra then absorbs
                                  each ray's spectrum sample by sample (table = thetas, then wavelengths and the
                                  absorptance grid; op[6] = number of wavelengths) and integrates the energy */
#define TRC_F_TABULATED   16u  /* TRC_O_REFRACTIVE between media given by material tables (Refractive :726-858,
                                  RefractiveAbsorbant :908-944): the indices are complex, m(lambda) = n + i k interpolated
                                  linearly from knots in the table pool -- material 1 at table_off (table_len knots:
                                  lambdas, then n, then k), material 2 at op[0] (op[1] knots).  The ray's index (columns
                                  d_ref_index + d_ref_index_im) equal to material 1's at its wavelength enters material 2
                                  and vice versa; Snell with the real parts, Fresnel with the complex ones (real part of
                                  R kept, :839); op[4] != 0: every child is attenuated by exp(-4 pi k path / lambda)
                                  with its own k and wavelength (optics.attenuations, optics.py:205-212), path scaled by
                                  op[7].  Needs d_wavelengths. */
#define TRC_F_IAM         4u   /* Martin-Ruiz incidence-angle modifier (IAM.__call__ :277-282) replaces the
                                  energy of Lambertian_IAM :303 / RealReflective_IAM :320:
                                  E' = E (1 - abs) (1 - exp(-cos^c / a_r)) / (1 - exp(-1 / a_r)),
                                  cos = |(d.n) n|; op[4] = a_r, op[5] = c */

/* One record per surface in Assembly.get_surfaces() order (assembly.py:60-77).
 * frame = rows 0..2 of Surface._temp_frame (has_frame.py:75), inv = rows 0..2
 * of numpy.linalg.inv(frame) computed on the host with the same routine the
 * reference uses (flat_surface.py:160).  Row-major, 4 columns. */
typedef struct TrcSurfaceRec {
    double frame[12];
    double inv[12];
    double gp[12];       /* geometry parameters, see TrcGeomKind */
    double op[8];        /* optics parameters, see TrcOpticsKind (op[4..5]: TRC_F_IAM) */
    int32_t gkind;
    int32_t okind;
    uint32_t flags;
    int32_t table_off;   /* offset (in doubles) into the scene's optics table pool, -1 = none */
    int32_t table_len;   /* number of (x, y) knots */
    int32_t pad_[3];
} TrcSurfaceRec;         /* 384 bytes */

/* Binary BVH node over per-surface global AABBs.  Replaces the relevance
 * mask produced by KdTree.traversal (accel_tree.py:213-312): instead of
 * materialising an (S, N) boolean matrix, the nearest-hit kernel walks this
 * tree per ray.  Leaf: count > 0, `first` indexes the leaf-surface list.
 * Inner: count == 0, children at `first` and `first + 1`. */
typedef struct TrcBvhNode {
    double lo[3];
    double hi[3];
    int32_t first;
    int32_t count;
    int32_t pad_[2];
} TrcBvhNode;            /* 64 bytes */

/* Largest leaf the tree walk accepts (bounds its candidate queue). */
#define TRC_BVH_MAX_LEAF 4

typedef struct TrcScene TrcScene;

int trc_abi_version(void);
const char *trc_last_error(void);
/* Number of CUDA kernels this library has launched in the process so far. */
uint64_t trc_launch_count(void);

/* Replaces the per-trace Python walk over Assembly.get_surfaces()
 * (tracer_engine.py:165-166): uploads the compiled surface table.
 * `tables` is a pool of float64 (x..., y...) interpolation tables for the
 * directional / spectral optics (may be NULL when n_table == 0). */
int trc_scene_create(const TrcSurfaceRec *surfaces, int32_t n_surfaces,
                     const double *tables, int32_t n_table,
                     TrcScene **out);
/* After Assembly.transform_children / HeliostatField.track_sun
 * (assembly.py:137-147, models/heliostat_field.py:89): refresh frames only. */
int trc_scene_update_frames(TrcScene *scene, const double *frames12,
                            const double *inv12, int32_t n_surfaces);
/* Acceleration structure.  Replaces KdTree.__init__/build_tree output
 * (accel_tree.py:20-146).  nodes[0] is the root.  leaf_surfaces lists the
 * surface indices referenced by leaves; `unbounded` lists the always-relevant
 * surfaces (accel_tree.py:59-73). */
/* Host-only binned-SAH builder over n global AABBs (lo3/hi3: n x 3 float64).
 * Replaces KdTree.build_tree + determine_split (accel_tree.py:42-204); the
 * layout is free because results do not depend on it.  Writes at
 * most 2n nodes (node_capacity >= 2n), the node count and the leaf ordering
 * (leaf k of a node covers leaf_order[first .. first+count)).  1 <= leaf_size
 * <= TRC_BVH_MAX_LEAF.  Needs no GPU. */
int trc_bvh_build(const double *lo3, const double *hi3, int32_t n, int32_t leaf_size,
                  TrcBvhNode *nodes, int32_t node_capacity, int32_t *n_nodes_out,
                  int32_t *leaf_order);
int trc_scene_set_bvh(TrcScene *scene, const TrcBvhNode *nodes, int32_t n_nodes,
                      const int32_t *leaf_surfaces, int32_t n_leaf_surfaces,
                      const int32_t *unbounded, int32_t n_unbounded);
/* Shadow grid for a directionally coherent bundle (every direction within `half_angle` radians of `direction`:
 * the first bounce of a solar source, sources.py:88-117, 330-464).  The nearest-hit entry points called with
 * use_bvh == 2 then find the candidates of such rays with one lookup in a plane grid of the surfaces' shadows
 * instead of the tree walk (KdTree.traversal, accel_tree.py:213-312, is what both replace); rays outside the cone
 * take the walk, and every candidate takes the same float64 test, so results do not change.  box_lo / box_hi:
 * global AABBs (n_boxes x 3 float64) of the bounded surfaces, surface_ids their indices.  *enabled = 0 when no
 * grid was made (cone too wide or grazing every axis, too many surfaces per cell, no shared memory to spare):
 * use_bvh == 2 is then an ordinary walk.  A NULL direction or negative half_angle removes the grid.  Host-built;
 * synchronises the device (call it between traces). */
int trc_scene_set_shadow_grid(TrcScene *scene, int32_t n_boxes, const double *box_lo, const double *box_hi,
                              const double *quads /* nullable: n_boxes x 4 corners x 3, global coordinates, of the
                                                     surfaces that lie inside a planar quadrilateral (plates,
                                                     triangles: last corner repeated); NaN in the first entry = none.
                                                     Their shadow is then the hull of the corners' shadows instead of
                                                     the box's, which halves the candidates of a tilted plate */,
                              const int32_t *surface_ids, const double *direction, double half_angle,
                              int32_t *enabled);
int trc_scene_destroy(TrcScene *scene);
int32_t trc_scene_num_surfaces(const TrcScene *scene);

/* Nearest hit across all surfaces for every ray.  Replaces
 * TracerEngine.intersect_ray (tracer_engine.py:27-64) together with every
 * GeometryManager.find_intersections it calls and,
 * when use_bvh != 0, KdTree.traversal (accel_tree.py:213-312).  use_bvh: 0 = every surface for every ray,
 * 1 = tree walk, 2 = tree walk with the scene's shadow grid for the rays it covers (trc_scene_set_shadow_grid;
 * the same in trc_trace_hits, trc_trace_hits_slice and TrcTraceArgs, where 2 applies to the bundle passed in).
 * d_hit_surf[i] = index of the first surface hit by ray i (lowest index wins
 * ties, strict '<' at tracer_engine.py:59), or -1. */
int trc_nearest_hit(const TrcScene *scene, int64_t n_rays,
                    const double *d_vertices, const double *d_directions, int64_t ld,
                    int32_t use_bvh, int32_t *d_hit_surf, void *stream);

/* Stable counting/radix sort of (key, value) pairs by key; keys < 0 are
 * dropped.  With keys = d_hit_surf and values = NULL (meaning 0..n-1) this is
 * the surface-major, ray-minor child ordering of the optics loop +
 * concatenate_rays (tracer_engine.py:218-266, ray_bundle.py:197-223).
 * With keys = cull flags it is the "survivors first, weak rays at the back"
 * order of tracer_engine.py:268-274.
 * d_work must hold trc_sort_work_bytes(n, n_keys) bytes.  The number of kept
 * pairs is written to d_count[0] (int64, device). */
int64_t trc_sort_work_bytes(int64_t n, int32_t n_keys);
int trc_sort_pairs(int64_t n, int32_t n_keys,
                   const int32_t *d_keys_in, const int32_t *d_vals_in,
                   int32_t *d_keys_out, int32_t *d_vals_out,
                   int64_t *d_count, void *d_work, void *stream);

/* trc_sort_pairs without a caller-provided work buffer and without any allocation: the work area is a
 * file-scope device array of the library (about 2 MB, held by every GPU that loads the library until exit).
 * No counterpart in the reference; added for callers that cannot call
 * cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  Same results as
 * trc_sort_pairs.  Limits: n <= TRC_SORT_STATIC_MAX_PAIRS, n_keys <= TRC_SORT_STATIC_MAX_KEYS, otherwise
 * TRC_ERR_CAPACITY.  All calls share the array: calls that overlap on the GPU race; the array keeps the last
 * call's contents (harmless: every cell is written before it is read). */
#define TRC_SORT_STATIC_MAX_PAIRS 4194304
#define TRC_SORT_STATIC_MAX_KEYS 256
int trc_sort_pairs_static(int64_t n, int32_t n_keys,
                          const int32_t *d_keys_in, const int32_t *d_vals_in,
                          int32_t *d_keys_out, int32_t *d_vals_out,
                          int64_t *d_count, void *stream);

/* Segment offsets of a sorted key array holding *d_count (<= n_max) entries:
 * d_offsets[k] = first position whose key >= k, for k in [0, n_keys]; i.e.
 * surface k owns slots [off[k], off[k+1]).  This is what the optics loop's
 * per-surface `intersections` masks encode (tracer_engine.py:218-228). */
int trc_segment_offsets_dyn(const int64_t *d_count, int64_t n_max, int32_t n_keys,
                            const int32_t *d_sorted_keys, int64_t *d_offsets, void *stream);

/* One hit of a bounce: the parent ray (copied while it is on chip in the
 * nearest-hit kernel, so the optics kernel never gathers from the multi-GB
 * parent bundle), its index in the parent bundle and the surface it hit. */
typedef struct TrcHitRecord {
    double v[3];
    double d[3];
    double energy;
    int32_t ray;
    int32_t surf;
} TrcHitRecord;          /* 64 bytes */

#define TRC_HIT_CHUNK 256 /* rays per chunk of trc_trace_hits */

/* trc_nearest_hit fused with the emission of the hit list.  Rays are processed
 * in chunks of TRC_HIT_CHUNK consecutive rays; chunk c appends its
 * d_chunk_cnt[c] records (ascending ray index) at d_records[d_chunk_off[c] ...].
 * Chunks land in arbitrary order; trc_order_hits restores the reference order.
 * *d_total (device, zeroed by this call) receives the number of hits; records
 * beyond `capacity` are dropped (the caller checks total <= capacity and
 * retries with a larger buffer).  d_hit_surf (n int32, as trc_nearest_hit) is
 * optional (NULL to skip). */
int trc_trace_hits(const TrcScene *scene, int64_t n_rays,
                   const double *d_vertices, const double *d_directions, const double *d_energy, int64_t ld,
                   int32_t use_bvh, int32_t *d_hit_surf,
                   TrcHitRecord *d_records, int64_t capacity,
                   uint32_t *d_chunk_off, uint32_t *d_chunk_cnt, int64_t *d_total, void *stream);

/* One slice of a bundle that is still arriving from the host (the engine copies
 * a host RayBundle in slices on a second stream and traces slice k while slice
 * k+1 is in flight).  Same as trc_trace_hits for the rays [ray_base, ray_base +
 * n_rays) of a larger bundle: every pointer addresses the slice's FIRST ray /
 * first chunk (ray_base must be a multiple of TRC_HIT_CHUNK, `ld` stays the row
 * stride of the whole bundle), records carry ray = ray_base + local index, and
 * *d_total is NOT zeroed: the caller zeroes it before the first slice and reads
 * the total after the last one.  The chunk tables of all slices together are
 * what trc_order_hits expects. */
int trc_trace_hits_slice(const TrcScene *scene, int64_t n_rays,
                         const double *d_vertices, const double *d_directions, const double *d_energy, int64_t ld,
                         int32_t use_bvh, int32_t *d_hit_surf,
                         TrcHitRecord *d_records, int32_t *d_keys /* nullable: surface index per record */,
                         int64_t capacity,
                         uint32_t *d_chunk_off, uint32_t *d_chunk_cnt, int64_t *d_total,
                         int64_t ray_base, void *stream);

/* Child order of the optics loop + concatenate_rays (tracer_engine.py:218-266,
 * ray_bundle.py:197-223): surface-major, parent-ray-minor.  From the chunk
 * table of trc_trace_hits builds, for every slot q < h of that order,
 * d_slot_surf[q] and d_slot_rec[q] (index of its record in d_records), plus the
 * per-surface segment offsets d_offsets[0..n_keys] (as trc_segment_offsets_dyn).
 * d_work must hold trc_order_work_bytes(n_chunks, h, n_keys) bytes. */
int64_t trc_order_work_bytes(int64_t n_chunks, int64_t h, int32_t n_keys);
int trc_order_hits(int64_t n_chunks, const uint32_t *d_chunk_off, const uint32_t *d_chunk_cnt,
                   const TrcHitRecord *d_records, int64_t h, int32_t n_keys,
                   int32_t *d_slot_surf, int32_t *d_slot_rec, int64_t *d_offsets,
                   void *d_work, void *stream);

/* Optics + accounting for the h ordered hits of one bounce.  Replaces the body
 * of the optics loop: Surface.select_rays / get_outgoing (surface.py:72-94),
 * GM.get_normals / get_intersection_points_global, every optics __call__
 * (optics_callables.py:93-627) and the Accountant.count calls
 * (optics_callables.py:1949-1961).
 *   slot q reads record d_records[d_slot_rec[q]] (parent ray + surface).
 *   child q: vertex = hit point, direction per optics, energy, parent index.
 *   d_draws[k] (k < 3) are injected random draws in slot order (NULL => the
 *   kernel draws from Philox4x32-10 keyed by (seed, source-ray id, bounce, draw slot)).
 *   d_wavelengths: per-PARENT-ray wavelengths (indexed by record.ray), nullable.
 *   Optional tallies (NULL to skip): d_absorbed[q] = E_in - E_out
 *   (AbsorptionAccountant :1638-1643), d_normals (3,h) (NormalAccountant),
 *   d_e_in[q], d_dirs_in (3,h) (Reception / Direction accountants),
 *   d_attenuated[q] = E_in - E_att (AttenuationAccountant :1670; 0 for optics without attenuation),
 *   d_out_wavelengths[q] (child wavelength = parent's).
 *   d_cull[q] = 1 where child energy <= min_energy (tracer_engine.py:242). */
typedef struct TrcOpticsArgs {
    int64_t h;
    const TrcHitRecord *d_records;
    const int32_t *d_slot_rec;
    const double *d_wavelengths;  /* nullable: per-parent-ray wavelength (spectral optics) */
    double *d_out_vertices;       /* child bundle, slot order */
    double *d_out_directions;
    double *d_out_energy;
    int64_t *d_out_parents;
    int64_t ld_out;
    const double *d_draws[3];
    uint64_t seed;
    uint32_t bounce;
    uint32_t pad_;
    double min_energy;
    double *d_absorbed;           /* nullable */
    double *d_normals;            /* nullable, (3,h) with stride ld_out */
    double *d_e_in;               /* nullable */
    double *d_dirs_in;            /* nullable, (3,h) with stride ld_out */
    double *d_out_wavelengths;    /* nullable */
    int32_t *d_cull;              /* nullable; int32 so it can key trc_sort_pairs */
    int64_t *d_n_cull;            /* nullable: device counter incremented by the number of culled children */
    double *d_attenuated;         /* nullable */
    /* Scenes with TRC_O_PERIODIC surfaces (NULL otherwise => child index == slot).  The child
     * bundle then has more entries than hits: with c_s = hits on surface s and m_s = 1 or 2
     * children per hit, surface s owns children [d_child_off[s], d_child_off[s+1]) and the j-th
     * hit of the surface writes child d_child_off[s] + j (and + c_s + j for its second child).
     * d_slot_surf / d_seg are the outputs of trc_order_hits; ld_out >= d_child_off[n_surfaces];
     * the child arrays, d_cull and d_out_wavelengths are indexed by child, every other tally by slot. */
    const int32_t *d_slot_surf;
    const int64_t *d_seg;
    const int64_t *d_child_off;
    /* RayBundle.ref_index (ray_bundle.py:33): the refractive index each parent ray travels in, and the
     * same column of the children (inherited, or the medium entered by a refracted ray).  Both NULL
     * unless the bundle carries the property; TRC_O_REFRACTIVE surfaces need them.  d_n_void counts the
     * void child slots (d_cull = 2) left by TRC_O_REFRACTIVE surfaces. */
    const double *d_ref_index;
    double *d_out_ref_index;
    int64_t *d_n_void;
    /* Identity of the SOURCE ray each ray descends from: the Philox draws of a hit are keyed by
     * (seed, source-ray id, bounce, draw slot), so a bundle sharded over several GPUs (sources.py
     * `.../procs`, tracer_engine_mp.py:19-35) draws exactly what the whole bundle draws.  Parent ray r has
     * id d_ray_ids[r], or ray_id_base + r when d_ray_ids is NULL (a source bundle); children inherit it
     * into d_out_ray_ids (nullable; indexed by child).  The two children of a TRC_O_REFRACTIVE split get
     * (id << 1) and (id << 1) | 1. */
    const int64_t *d_ray_ids;
    int64_t ray_id_base;
    int64_t *d_out_ray_ids;
    /* imaginary part of ref_index, per ray / per child (nullable: zero): scenes with TRC_F_TABULATED surfaces */
    const double *d_ref_index_im;
    double *d_out_ref_index_im;
} TrcOpticsArgs;
int trc_optics(const TrcScene *scene, const TrcOpticsArgs *args, void *stream);

/* Polychromatic rays: bundle properties `wavelengths` and `spectra`, both (W, n) (optics_callables.py:393-425,
 * ray_bundle.py:65-100).  Runs after trc_optics on a scene without two-children optics and finishes each child
 * slot q (child index == slot): copies the parent's W wavelengths, and writes the child's spectrum
 *   - TRC_F_POLY surfaces: spectrum * (1 - absorptance(theta_in, wavelength_w)), bilinear as scipy's
 *     RegularGridInterpolator (:403-412); the slot's energy becomes the trapezoid integral of the new spectrum
 *     over the ray's wavelengths (numpy.trapz, :413), d_absorbed[q] the energy lost, d_cull[q] is re-decided;
 *   - Reflective, RealReflective (and their one-sided forms), Lambertian: spectrum * (1 - absorptivity)
 *     (:137, :173, :266); Lambertian_directional_axisymmetric_piecewise: spectrum * (1 - absorptance(theta_in)) (:358);
 *   - every other kind: the parent's spectrum.
 * d_lost_spectra (nullable, (W, ld_out)) receives parent minus child spectrum: what PolychromaticAccountant.count
 * stores (:1835-1838).  d_normals / d_dirs_in: the accountant columns trc_optics wrote for the same slots.
 * *d_n_cull must be zero on entry; it counts every slot whose final energy is <= min_energy. */
typedef struct TrcSpectraArgs {
    int64_t h;                        /* child slots */
    int32_t n_samples;                /* W */
    const int32_t *d_slot_surf;       /* (h,) surface of each slot */
    const int64_t *d_parents;         /* (h,) parent ray of each slot */
    const double *d_in_wavelengths, *d_in_spectra; int64_t ld_in;   /* (W, ld_in) of the incoming bundle */
    const double *d_normals, *d_dirs_in; int64_t ld_cols;           /* (3, ld_cols) per slot */
    double *d_out_wavelengths, *d_out_spectra; int64_t ld_out;      /* (W, ld_out) of the children */
    double *d_energy;                 /* (h,) child energy: read (== parent energy on TRC_F_POLY surfaces), rewritten there */
    double *d_absorbed;               /* (h,) nullable */
    double *d_lost_spectra;           /* (W, ld_out) nullable */
    int32_t *d_cull;                  /* (h,) */
    int64_t *d_n_cull;
    double min_energy;
} TrcSpectraArgs;
int trc_spectra(const TrcScene *scene, const TrcSpectraArgs *args, void *stream);


/* The reference loop (tracer_engine.py:193-274) for up to TRC_TRACE_MAX_BOUNCES iterations, enqueued back to
 * back on the stream with NO host synchronisation in between: trc_trace_hits -> child order -> trc_optics
 * (Philox draws) -> per-surface tallies -> survivors-first reorder, per bounce, then one asynchronous copy of
 * every count to h_counts.  The caller synchronises the stream once and reads h_counts.  The ray count of
 * every bounce after the first, the hit counts and the cull counts live in device memory; kernels are launched
 * on grids sized from `capacity[j]` (hits the buffers of bounce j can hold; a multiple of 256).  A bounce
 * with more hits than its capacity does nothing (hits is still reported), and nothing after it runs on real
 * data: the caller enlarges the capacity and traces again from that bounce (its input, the children of the
 * bounce before, is intact).  Same kernels and the same results as the step-by-step entry points.  Not for
 * scenes with two-children optics (TRC_O_PERIODIC, TRC_O_REFRACTIVE) nor injected draws.
 *
 * Memory: two caller-owned device arenas sized by trc_trace_plan.  d_scratch is reused by every bounce and may
 * be freed after the synchronisation; d_keep holds the outputs of bounce j at plan.bounce[j] (byte offsets, -1 =
 * absent): rows of `ld` = capacity[j] elements of which the first `hits` are valid.
 *   children in slot order (surface-major, parent-minor): off_vertices (3,ld), off_directions (3,ld), off_energy,
 *     off_parents (int64), off_ray_ids (int64), off_wavelengths;
 *   off_alt_*: the same children with survivors first and weak rays at the back (tracer_engine.py:242-245,
 *     268-274) -- only written when the bounce's `reorder` word is 1 (some but not all children weak); otherwise
 *     the slot order already IS that order.  The next bounce reads whichever applies (decided on the device);
 *   accountant columns in slot order: off_absorbed, off_normals, off_e_in, off_dirs_in, off_attenuated;
 *   off_seg: int64 (n_surfaces + 1), surface s owns slots [seg[s], seg[s+1]).
 * h_counts (page-locked host memory, plan.counts_words int64): per bounce [rays in, hits, survivors, reorder]
 * followed, with TRC_TRACE_SEG, by the n_surfaces + 1 words of seg. */
#define TRC_COL_NORMALS    1u
#define TRC_COL_E_IN       2u
#define TRC_COL_DIRS_IN    4u
#define TRC_COL_ATTENUATED 8u
#define TRC_TRACE_MAX_BOUNCES 8
#define TRC_TRACE_TIME_WALK 1   /* bracket the first bounce's nearest-hit kernel with CUDA events (trc_trace_walk_ms) */
#define TRC_TRACE_WALK_DONE 2   /* the caller ran trc_trace_init and then the first bounce's nearest-hit pass itself
                                   (trc_trace_hits_slice on a bundle still arriving from the host) into the scratch
                                   buffers named by the plan */
#define TRC_TRACE_SEG       4   /* copy seg of every bounce to h_counts (accountants need it on the host) */
#define TRC_TRACE_ZERO_TALLIES 8 /* zero d_absorbed_per_surface / d_hits_per_surface first (trc_trace_init's kernel) */
typedef struct TrcTraceArgs {
    int64_t n;                        /* rays entering the first bounce */
    const double *d_vertices;         /* (3,n) rows of stride ld */
    const double *d_directions;
    const double *d_energy;
    int64_t ld;
    const double *d_wavelengths;      /* nullable, per ray */
    const int64_t *d_ray_ids;         /* nullable: source-ray ids (see TrcOpticsArgs); else ray_id_base + index */
    int64_t ray_id_base;
    int32_t use_bvh;
    uint32_t first_bounce;            /* index of the first bounce in the whole trace (Philox stream) */
    int32_t n_bounces;
    int32_t flags;                    /* TRC_TRACE_* */
    uint64_t seed;
    double min_energy;
    uint32_t columns;                 /* TRC_COL_*: accountant columns to produce besides `absorbed` */
    uint32_t pad_;
    int64_t capacity[TRC_TRACE_MAX_BOUNCES];
    void *d_keep;
    int64_t keep_bytes;
    void *d_scratch;
    int64_t scratch_bytes;
    double *d_absorbed_per_surface;   /* (n_surfaces) running totals, updated */
    int64_t *d_hits_per_surface;      /* (n_surfaces) running totals, updated */
    int64_t *h_counts;                /* page-locked host memory */
} TrcTraceArgs;
typedef struct TrcTraceBounceLayout {
    int64_t ld;
    int64_t off_vertices, off_directions, off_energy, off_parents, off_ray_ids, off_wavelengths;
    int64_t off_alt_vertices, off_alt_directions, off_alt_energy, off_alt_parents, off_alt_ray_ids, off_alt_wavelengths;
    int64_t off_absorbed, off_normals, off_e_in, off_dirs_in, off_attenuated;
    int64_t off_seg;
} TrcTraceBounceLayout;
typedef struct TrcTracePlan {
    int64_t keep_bytes, scratch_bytes, counts_words;
    TrcTraceBounceLayout bounce[TRC_TRACE_MAX_BOUNCES];
    /* byte offsets into d_scratch of the first bounce's hit list, for TRC_TRACE_WALK_DONE callers:
     * records, keys, chunk tables and the int64 hit total (trc_trace_hits_slice arguments) */
    int64_t off_records, off_keys, off_chunk_off, off_chunk_cnt, off_total;
} TrcTracePlan;
/* Needs n, ld, n_bounces, capacity, columns, flags and whether d_wavelengths is set; no GPU work. */
int trc_trace_plan(const TrcScene *scene, const TrcTraceArgs *args, TrcTracePlan *plan);
int trc_trace_init(const TrcScene *scene, const TrcTraceArgs *args, void *stream);
int trc_trace(const TrcScene *scene, const TrcTraceArgs *args, void *stream);
/* Device time of the first bounce's nearest-hit kernel of the last trc_trace with TRC_TRACE_TIME_WALK on this
 * scene (waits for it).  A TrcScene serves one stream / one host thread at a time. */
int trc_trace_walk_ms(const TrcScene *scene, float *ms);

/* Row gather: out[r, j] = in[r, idx[j]] for r < n_rows (float64 rows) and
 * optionally an int64 row.  Applies the cull permutation to a child bundle
 * (RayBundle.inherit with an index selector, ray_bundle.py:117-143). */
int trc_gather_rows(int64_t n_out, const int32_t *d_idx,
                    int32_t n_rows, const double *d_in, int64_t ld_in,
                    double *d_out, int64_t ld_out,
                    const int64_t *d_in_i64, int64_t *d_out_i64, void *stream);

/* Sources (tracer/sources.py).  All draw from Philox4x32-10 keyed by
 * (seed, first_ray + i); rot_dir / rot_pos are the 3x3 row-major matrices
 * returned by rotation_to_z (spatial_geometry.py:24-48) for the ray direction
 * and the source-plane normal.
 *   trc_source_pillbox_disk: disk_bundle (sources.py:175-239)
 *   trc_source_buie_disk:    buie_sunshape (sources.py:412-464) with the
 *                            Buie_directions_sampling tables (:330-384)
 *                            precomputed on the host (n_seg = 210). */
int trc_source_pillbox_disk(int64_t n, int64_t first_ray, uint64_t seed,
                            const double *center3, const double *rot_dir9, const double *rot_pos9,
                            double radius_in, double radius, double ang0, double ang1,
                            double ang_range, double ray_energy,
                            double *d_vertices, double *d_directions, double *d_energy,
                            int64_t ld, void *stream);
typedef struct TrcBuieTables {
    int32_t n_seg;            /* 210 */
    int32_t has_aureole;      /* CSR > 0 */
    double cdf[212];          /* CDF_DNI[0..n_seg] */
    double theta[212];        /* theta_int[0..n_seg] */
    double w[212];            /* phi_dni*cos*sin at theta_int[i] */
    double sum_integ_dni;
    double gamma, kappa, theta_dni, theta_tot;
} TrcBuieTables;
int trc_source_buie_disk(int64_t n, int64_t first_ray, uint64_t seed,
                         const TrcBuieTables *tables,
                         const double *center3, const double *rot_dir9, const double *rot_pos9,
                         double radius, double ray_energy,
                         double *d_vertices, double *d_directions, double *d_energy,
                         int64_t ld, void *stream);
/* Rectangular source planes: rect_bundle / oblique_solar_rect_bundle (sources.py:241-302) and
 * rect_buie_sunshape (:466-515).  Start points uniform in [-half_x, half_x] x [-half_y, half_y]
 * of the source plane's local frame (rot_pos), directions as the disk variants. */
int trc_source_pillbox_rect(int64_t n, int64_t first_ray, uint64_t seed,
                            const double *center3, const double *rot_dir9, const double *rot_pos9,
                            double half_x, double half_y, double ang_range, double ray_energy,
                            double *d_vertices, double *d_directions, double *d_energy,
                            int64_t ld, void *stream);
int trc_source_buie_rect(int64_t n, int64_t first_ray, uint64_t seed,
                         const TrcBuieTables *tables,
                         const double *center3, const double *rot_dir9, const double *rot_pos9,
                         double half_x, double half_y, double ray_energy,
                         double *d_vertices, double *d_directions, double *d_energy,
                         int64_t ld, void *stream);

/* Flux-map binning on device.  Replaces Surface.global_to_local
 * (surface.py:114-126) + RectPlateGM.get_fluxmap's histogram2d
 * (flat_surface.py:237-251): d_bins (nx*ny, row-major [ix][iy]) += energy. */
int trc_fluxmap_rect(int64_t n, const double *d_hits, int64_t ld, const double *d_energy,
                     const double *proj12 /* host, rows 0..2 of rounded inverse frame */,
                     double hx, double hy, int32_t nx, int32_t ny,
                     double *d_bins, void *stream);

/* Flux-map binning in the mesh coordinates of the other geometry managers
 * (get_fluxmap of RoundPlateGM flat_surface.py:524, ParabolicDishGM paraboloid.py:151,
 * FiniteCylinder cylinder.py:139, SphericalGM sphere_surface.py:100, and RectPlateGM with
 * explicit edges): local = proj12 . [hit; 1] (Surface.global_to_local, surface.py:114-126),
 * (u, v) by `coord`, weights d_energy summed into d_bins[iu * nv + iv] (caller zeroes it)
 * with numpy.histogram2d's rule on the explicit edge arrays d_edges_u[0..nu], d_edges_v[0..nv]
 * (device memory; [e_i, e_i+1), last bin closed).  The host divides by the bin areas. */
typedef enum TrcFluxCoord {
    TRC_FLUX_XY = 0,         /* u = x, v = y */
    TRC_FLUX_R_ATAN_XY = 1,  /* u = sqrt(x^2+y^2), v = atan2(x, y) wrapped to [0, 2pi)  (RoundPlateGM) */
    TRC_FLUX_R_ATAN_YX = 2,  /* u = sqrt(x^2+y^2), v = atan2(y, x) wrapped             (dishes) */
    TRC_FLUX_Z_PHI = 3,      /* u = z, v = atan2(y, x) wrapped                         (cylinders) */
    TRC_FLUX_THETA_PHI = 4,  /* u = acos(z / param), v = atan2(y, x) wrapped; param = radius (spheres) */
    TRC_FLUX_COORD_COUNT = 5
} TrcFluxCoord;
int trc_fluxmap_binned(int32_t coord, int64_t n, const double *d_hits, int64_t ld, const double *d_energy,
                       const double *proj12, double param,
                       const double *d_edges_u, int32_t nu, const double *d_edges_v, int32_t nv,
                       double *d_bins, void *stream);

/* Per-surface energy totals: d_totals[k] += sum of d_values[q] over the h slots
 * whose (sorted) key d_sorted_keys[q] == k.  Feeds the absorbed-energy tally that
 * is all-reduced across GPUs (the merge of tracer_engine_mp.py:64-119 reduced to
 * its sums). */
int trc_segment_sums(int64_t h, const int32_t *d_sorted_keys, const double *d_values,
                     double *d_totals, void *stream);

/* Measurement aid, not part of the path: sustained FMA rate of the float32 (fp64 = 0) or float64 (fp64 = 1)
 * pipe of the current device in TFLOP/s (2 flop per FMA) -- the roofline denominator SURVEY.md 8(d) asks the
 * builder to measure for the brute-force nearest-hit kernel.  Synchronous. */
int trc_fma_peak(int32_t fp64, double *tflops);
/* Instrumentation (no counterpart in the reference): time, in ms, to stream n rays of a bundle (vertices and
 * directions, six float64 rows of stride ld) through persistent warps that own chunks of 256 rays, as the shadow-grid
 * kernel reads them.  variant 1, 2, 4: that many 32-ray groups in flight per warp by register prefetch (16 warps per
 * SM), chunk ids from one atomic counter; 11, 12, 14: the same with the chunks dealt round robin; 108: one
 * cp.async.bulk per row and chunk into a two-chunk shared-memory ring, 8 warps per SM.  Allocates a few bytes,
 * synchronises; scripts/stream_probe.py. */
int trc_stream_probe(int32_t variant, const double *d_vertices, const double *d_directions, int64_t ld, int64_t n,
                     float *ms_out);

#ifdef __cplusplus
}
#endif
#endif /* TRACER_B200_H */
