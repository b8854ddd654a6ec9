"""TracerEngine: the reference's driver API (tracer/tracer_engine.py:11-295) on
top of the sm_100a kernels.

``TracerEngine(assembly).ray_tracer(bundle, reps, min_energy, tree, accel,
Kd_Tree)`` keeps the reference signature, return value and side effects
(``self.tree``, accountant data behind ``get_all_hits()``, log messages) but the
per-bounce work -- every find_intersections, the nearest-hit reduction, the
optics callables, accountants, culling and child-bundle concatenation -- runs on
the GPU through libtracer_b200.so.  With Philox draws the whole loop is device
driven (``trc_trace``): Python plans buffer capacities from the previous trace,
enqueues up to eight bounces in one foreign call, synchronises ONCE and reads
every count back together; no count visits the host in between.

Random draws: ``rng='philox'`` (default) draws on device, keyed by
(seed, source-ray id, bounce, draw slot), so a bundle sharded over GPUs draws
what the whole bundle draws.  ``rng='numpy'`` draws on the host from
``numpy.random`` in exactly the reference's order (bounce -> surface ->
draw kind -> hit) and injects them, so a trace with the
same ``numpy.random.seed`` reproduces the reference bundle for bundle.
"""
import ctypes as C
import logging

import numpy as np

from . import _lib as L
from .ray_bundle import RayBundle, _DeviceColumns
from .scene import CompiledScene
from .trace_tree import RayTree


class _LazyRows(object):
    """Device-resident chunk of accountant data; looks like an ndarray to
    numpy.hstack (shape, len, __array__) and is copied to the host on demand."""

    def __init__(self, tensor):
        self.t = tensor
        self.shape = tuple(tensor.shape)
        self.ndim = tensor.dim()
        self._host = None

    def __len__(self):
        return self.shape[0]

    def __array__(self, dtype=None, copy=None):
        if self._host is None:
            self._host = self.t.cpu().numpy()
            self.t = None
        return self._host if dtype is None else self._host.astype(dtype)


class _DeviceSeg(object):
    """Per-surface slot offsets of a bounce that stay in HBM (scenes with very many surfaces, e.g. a 1e6-face
    mesh: 8 MB per bounce): the few entries the engine needs arrive with one small gather, the whole array is
    only copied when somebody asks for it (``TriangleMeshObject.face_of_hits``)."""

    def __init__(self, tensor, known=None):
        self.t = tensor
        self.known = known or {}
        self._host = None

    def __len__(self):
        return int(self.t.shape[0]) if self._host is None else len(self._host)

    def __array__(self, dtype=None, copy=None):
        if self._host is None:
            self._host = self.t.cpu().numpy()
            self.t = None
        return self._host if dtype is None else self._host.astype(dtype)

    def __getitem__(self, i):
        if isinstance(i, slice):
            if self._host is not None:
                return self._host[i]
            return _DeviceSeg(self.t[i].clone())
        if self._host is None and i in self.known:
            return self.known[i]
        return self.__array__()[i]


INLINE_SEG_MAX = 4096          # scenes with more surfaces keep the slot offsets of a bounce on the device


class _DevBundle(object):
    """Engine-internal view of a bundle living in HBM.  ``slices`` is set while a
    host bundle is still arriving: [(first ray, end ray, CUDA event of its copy)]."""
    __slots__ = ('V', 'D', 'E', 'W', 'R', 'RI', 'I', 'SP', 'id_base', 'n', 'slices', 'keep')

    def __init__(self, V, D, E, W=None, slices=None, keep=None, R=None, I=None, id_base=0, SP=None, RI=None):
        self.V, self.D, self.E, self.W = V, D, E, W
        self.RI = RI              # imaginary part of ref_index (media given by material tables), else None
        self.SP = SP              # spectra (W, n) of polychromatic rays; W is then (W, n) too
        self.R = R                # ref_index column (refractive scenes), else None
        self.I = I                # source-ray id of each ray (int64), None: id_base + index
        self.id_base = int(id_base)
        self.n = int(V.shape[1])
        self.slices = slices
        self.keep = keep          # host arrays of the copies in flight


STREAM_MIN_RAYS = 1 << 22     # host bundles at least this large are uploaded in slices ...
STREAM_SLICES = 8             # ... this many, each traced while the next one is copied
FULL_CAP_RAYS = 1 << 18       # bundles up to this size get buffers for "every ray hits" (no overflow possible)
BLIND_BOUNCES = 4             # bounces enqueued per call while the depth of the trace is unknown


GRID_MIN_RAYS = 1 << 16       # bundles at least this large with a narrow direction cone use the scene's shadow grid
CONE_SAMPLE = 1 << 16         # rays of a host bundle looked at to estimate its cone
MAX_CONE = 0.25               # rad


def _direction_cone(bundle):
    """(axis, half angle) of a cone holding the bundle's directions, or None.  Device sources state it
    (sources._wrap); for a large host bundle it is estimated from the first CONE_SAMPLE rays with a margin.
    It is a pruning hint only: the kernel checks every ray against the cone's slope box and walks the tree
    for those outside."""
    cone = getattr(bundle, '_dir_cone', None)
    if cone is not None:
        return cone
    props = getattr(bundle, '_props', None)
    d = props.get('directions') if props is not None else None
    if not isinstance(d, np.ndarray) or d.ndim != 2 or d.shape[0] != 3 or d.shape[1] < (1 << 20):
        return None
    sample = np.asarray(d[:, :CONE_SAMPLE], dtype=float)
    norm = np.sqrt(np.sum(sample ** 2, axis=0))
    if not np.all(norm > 0.):
        return None
    unit = sample / norm
    axis = unit.mean(axis=1)
    length = np.sqrt(np.sum(axis ** 2))
    if not length > 0.5:
        return None
    axis = axis / length
    half = float(np.arccos(np.clip(np.min(np.dot(axis, unit)), -1., 1.))) * 1.05 + 1e-4
    return (axis, half) if half <= MAX_CONE else None


def _round_cap(x):
    return max(256, (int(x) + 255) // 256 * 256)


class TracerEngine(object):
    def __init__(self, parent_assembly, loglevel=logging.DEBUG, device=None, rng='philox', seed=0,
                 device_accel='auto', leaf_size=1):
        """
        parent_assembly - the top-level Assembly (reference's or this package's).
        loglevel        - level of the per-bounce log lines (as the reference).
        device          - torch device (default: current CUDA device).
        rng             - 'philox' (device draws) or 'numpy' (host draws in the
                          reference's order, injected into the kernels).
        seed            - Philox seed.
        device_accel    - 'auto' | 'brute' | 'bvh': how the nearest-hit kernel
                          finds candidates.  Results are identical.
        """
        import torch
        L.load()  # fail loudly when the CUDA library is missing
        if not torch.cuda.is_available():
            raise RuntimeError("tracer_b200 needs a CUDA device: the bounce loop has no CPU path")
        self._asm = parent_assembly
        self.loglevel = loglevel
        self.device = torch.device('cuda', torch.cuda.current_device()) if device is None else torch.device(device)
        if rng not in ('philox', 'numpy'):
            raise ValueError("rng must be 'philox' or 'numpy'")
        self.rng = rng
        self.seed = int(seed)
        if device_accel not in ('auto', 'brute', 'bvh'):
            raise ValueError("device_accel must be 'auto', 'brute' or 'bvh'")
        self.device_accel = device_accel
        self.leaf_size = leaf_size
        self._scene = None
        self._frames_key = None
        self._hit_frac = {}       # bounce -> hit fraction of the previous trace (sizes the hit list)
        self._hist = {}           # bounce -> (rays in, hits, survivors) of the previous trace (plans trc_trace)
        self._cap_floor = {}      # bounce -> smallest capacity that is known to fit (set after an overflow)
        self._counts_host = None  # page-locked int64 words for the one read-back of a trace
        self._tallies_dirty = False
        self._copy_stream = None
        self._keep_tree = False
        self.shadow_grid = True    # first bounce of a narrow-cone bundle through the scene's shadow grid (results equal)
        self.fused_bounce = True   # sequence a bounce inside the library (trc_bounce) when the configuration allows
        self.kernel_events = None  # set to a list to collect (bounce, n, H, start, end) CUDA events of trc_trace_hits
        self.stats = {}

    @staticmethod
    def _to_host(t):
        """Device tensor -> NumPy array.  Large results land in page-locked memory (torch's
        caching host allocator: ~50 GB/s instead of the ~5 GB/s of a pageable copy, and the
        block is recycled once the caller drops the array)."""
        import torch
        if t.numel() == 0:
            return np.zeros(tuple(t.shape))
        t = t.contiguous()
        if t.numel() < (1 << 20):
            return t.cpu().numpy()
        host = torch.empty(t.shape, dtype=t.dtype, pin_memory=True)
        host.copy_(t, non_blocking=True)
        torch.cuda.current_stream(t.device).synchronize()
        return host.numpy()

    # ------------------------------------------------------------ scene
    def compile(self, force=False):
        """(Re)compile the assembly into device tables when it changed."""
        surfaces = self._asm.get_surfaces()
        # Every transform of the scene graph assigns a NEW frame array to each surface
        # (has_frame.py / reference has_frame.py:75), so object identity is a cheap change
        # detector; the arrays are kept referenced so that their ids cannot be recycled.
        if hasattr(surfaces, 'parts'):        # SurfaceSeq: a mesh object is one entry (its frame array is re-made
            surfaces = [s for p in surfaces.parts for s in (p if isinstance(p, list) else [p._mesh])]   # on transform)
            frames = [s._global_frames if hasattr(s, '_global_frames') else s._temp_frame for s in surfaces]
        else:
            frames = [s._temp_frame for s in surfaces]
        # Optics and geometry managers are part of the key by identity as well: replacing one (surf._opt = ...)
        # recompiles.  Editing a parameter IN PLACE (opt._abs = 0.5, gm._half_dims[0] = 2.) is not seen --
        # call compile(force=True) after such an edit (the reference re-reads the live objects every trace).
        try:
            managers = [s._opt for s in surfaces], [s._geom for s in surfaces]
        except AttributeError:              # a mesh object stands for its faces: no managers of its own
            managers = ([getattr(s, '_opt', None) for s in surfaces], [getattr(s, '_geom', None) for s in surfaces])
        key = (tuple(map(id, surfaces)), tuple(map(id, frames)), tuple(map(id, managers[0])), tuple(map(id, managers[1])))
        if force or self._scene is None or key != self._frames_key:
            if self._scene is not None:
                self._scene.close()
            self._scene = CompiledScene(self._asm, leaf_size=self.leaf_size)
            self._hit_frac = {}
            self._hist, self._cap_floor = {}, {}
            self._frames_key = key
            self._frames_ref = (surfaces, frames, managers)
        return self._scene

    # ----------------------------------------------------------- bundles
    def _to_device(self, bundle):
        import torch

        def col(name, required=True):
            if not bundle.has_property(name) or not getattr(bundle, 'is_set', lambda n: True)(name):
                if required:
                    raise ValueError(f"bundle has no {name}")
                return None
            raw = bundle._props.get(name) if hasattr(bundle, '_props') else None
            if isinstance(raw, _DeviceColumns):
                t = raw.pieces[0] if len(raw.pieces) == 1 else torch.cat(raw.pieces, dim=-1)
                return t.to(self.device)
            try:
                arr = getattr(bundle, 'get_' + name)()
            except (KeyError, AttributeError):
                if required:
                    raise ValueError(f"bundle has no {name}")
                return None
            arr = np.ascontiguousarray(arr, dtype=np.float64)
            return torch.from_numpy(arr).to(self.device, non_blocking=True)

        id_base = int(getattr(bundle, '_ray_id_base', 0) or 0)
        streamed = self._upload_in_slices(bundle)
        if streamed is not None:
            streamed.id_base = id_base
            return streamed
        V, D, E = col('vertices'), col('directions'), col('energy')
        W = col('wavelengths', required=False) if bundle.has_property('wavelengths') else None
        SP = col('spectra', required=False) if bundle.has_property('spectra') else None
        if SP is not None:
            if W is None or SP.dim() != 2 or W.shape != SP.shape or SP.shape[1] != E.shape[0]:
                raise ValueError("polychromatic bundle: 'wavelengths' and 'spectra' must both be (W, n)")
        elif W is not None and W.shape != E.shape:
            raise ValueError("bundle wavelengths must be (n,) (or (W, n) next to 'spectra')")
        R, RI = self._ref_index_columns(bundle)
        if R is not None and R.shape != E.shape:
            raise ValueError("bundle ref_index must be (n,)")
        if V.dim() != 2 or V.shape[0] != 3 or D.shape != V.shape or E.shape[0] != V.shape[1]:
            raise ValueError("bundle arrays must be vertices (3,n), directions (3,n), energy (n,)")
        return _DevBundle(V.contiguous(), D.contiguous(), E.contiguous(), None if W is None else W.contiguous(),
                          R=None if R is None else R.contiguous(), id_base=id_base,
                          SP=None if SP is None else SP.contiguous(), RI=None if RI is None else RI.contiguous())

    def _ref_index_columns(self, bundle):
        """(real, imaginary) device columns of the bundle's ref_index; the imaginary one is None for a real index."""
        import torch
        if not bundle.has_property('ref_index') or not getattr(bundle, 'is_set', lambda n: True)('ref_index'):
            return None, None
        raw = bundle._props.get('ref_index') if hasattr(bundle, '_props') else None
        if isinstance(raw, _DeviceColumns):
            t = raw.pieces[0] if len(raw.pieces) == 1 else torch.cat(raw.pieces, dim=-1)
            t = t.to(self.device)
            if t.is_complex():
                return t.real.contiguous(), t.imag.contiguous()
            return t, None
        try:
            arr = np.asarray(bundle.get_ref_index())
        except (KeyError, AttributeError):
            return None, None
        to_dev = (lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64)).to(self.device, non_blocking=True))
        if np.iscomplexobj(arr):
            return to_dev(arr.real), to_dev(arr.imag)
        return to_dev(arr), None

    def _upload_in_slices(self, bundle):
        """Large host bundle: allocate it in HBM and enqueue the host->device copy in
        STREAM_SLICES slices on a second stream, one event per slice; bounce 1 traces
        slice k (trc_trace_hits_slice) while slice k+1 is on the wire.  Returns None
        when the bundle is device-backed, small, or carries wavelengths."""
        import torch
        props = getattr(bundle, '_props', None)
        if props is None or bundle.has_property('wavelengths') and bundle.is_set('wavelengths'):
            return None
        if props.get('ref_index') is not None:
            return None
        raw = [props.get(k) for k in ('vertices', 'directions', 'energy')]
        if any(not isinstance(a, np.ndarray) for a in raw):
            return None
        hV, hD, hE = (np.ascontiguousarray(a, dtype=np.float64) for a in raw)
        if hV.ndim != 2 or hV.shape[0] != 3 or hD.shape != hV.shape or hE.shape != (hV.shape[1],):
            return None          # the regular path raises the reference's error
        n = hV.shape[1]
        if n < STREAM_MIN_RAYS:
            return None
        dev = self.device
        V = torch.empty((3, n), dtype=torch.float64, device=dev)
        D = torch.empty((3, n), dtype=torch.float64, device=dev)
        E = torch.empty(n, dtype=torch.float64, device=dev)
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(device=dev)
        cs = self._copy_stream
        cs.wait_stream(torch.cuda.current_stream(dev))       # the blocks may have just been freed by compute work
        step = -(-n // STREAM_SLICES)
        step = -(-step // L.HIT_CHUNK) * L.HIT_CHUNK
        tV, tD, tE = torch.from_numpy(hV), torch.from_numpy(hD), torch.from_numpy(hE)
        slices = []
        with torch.cuda.stream(cs):
            for a in range(0, n, step):
                b = min(n, a + step)
                for k in range(3):
                    V[k, a:b].copy_(tV[k, a:b], non_blocking=True)
                    D[k, a:b].copy_(tD[k, a:b], non_blocking=True)
                E[a:b].copy_(tE[a:b], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(cs)
                slices.append((a, b, ev))
        return _DevBundle(V, D, E, None, slices, keep=(tV, tD, tE))

    # ------------------------------------------------------------ draws
    @staticmethod
    def _draws_for(rec, table, n, theta):
        """numpy.random draws of one optics callable on n hits, in the order its __call__ consumes them."""
        k = rec['okind']
        vals = [None, None, None]
        if k == L.O_REAL_REFLECTIVE:
            sig = rec['op'][1]
            if sig > 0.:
                if rec['op'][2] != 0.:
                    vals[0] = np.random.normal(scale=sig, size=(n,))
                    vals[1] = np.random.normal(scale=sig, size=(n,))
                else:
                    vals[0] = np.random.normal(scale=sig, size=(n,))
                    vals[1] = np.random.uniform(low=0., high=2. * np.pi, size=(n,))
        elif k == L.O_LAMB_PIECEWISE_SPEC:
            th_tab, rest = table
            spec = np.interp(theta, th_tab, rest[len(th_tab):])
            r = np.random.rand(n)
            diffuse = ~(r < spec)
            nd = int(diffuse.sum())
            vals[0] = r
            vals[1], vals[2] = np.zeros(n), np.zeros(n)
            vals[1][diffuse] = np.random.uniform(low=0., high=2. * np.pi, size=nd)
            vals[2][diffuse] = np.random.uniform(size=nd)
        elif k == L.O_REFRACTIVE:
            # _refract_dirs (:1229-1231) then _make_output_refraction_bundles (:1255)
            if rec['op'][3] >= 0.:
                vals[0] = np.random.normal(scale=rec['op'][3], size=n)
                vals[1] = np.random.uniform(low=0., high=2. * np.pi, size=n)
            if rec['op'][2] != 0.:
                vals[2] = np.random.uniform(size=n)
        elif k in (L.O_LAMBERTIAN, L.O_LAMB_DIRECTIONAL, L.O_LAMB_DIRECTIONAL_SPECTRAL,
                   L.O_LAMBERTIAN_ABSORBANT):
            ang = rec['op'][1] if k in (L.O_LAMBERTIAN, L.O_LAMBERTIAN_ABSORBANT) else np.pi / 2.
            vals[0] = np.random.uniform(low=0., high=2. * np.pi, size=n)
            if ang != 0.:
                vals[1] = np.random.uniform(size=n)
        elif k in (L.O_LAMBERTIAN_SPECULAR, L.O_LAMBSPEC_DIRECTIONAL):
            spec = rec['op'][1] if k == L.O_LAMBERTIAN_SPECULAR else rec['op'][0]
            r = np.random.rand(n)
            diffuse = ~(r < spec)
            nd = int(diffuse.sum())
            vals[0] = r
            vals[1], vals[2] = np.zeros(n), np.zeros(n)
            vals[1][diffuse] = np.random.uniform(low=0., high=2. * np.pi, size=nd)
            vals[2][diffuse] = np.random.uniform(size=nd)
        return vals

    def _numpy_draws(self, scene, seg, H, theta_in=None, back=None):
        """Host draws in the reference's consumption order; returns the (3, H) device array of the surfaces'
        (front) optics and, for scenes with BiFacial managers, a second one for the back sides (else None).
        ``theta_in`` (host, slot order) is needed by optics whose specular/diffuse split depends on the incidence
        angle; ``back`` (host bool, slot order) marks the hits on the back of a BiFacial surface: a side draws
        (for ALL the surface's hits) only when some hit is on it, the back side first (optics_callables.py:1891-1894)."""
        import torch
        draws = np.zeros((3, H))
        draws_b = np.zeros((3, H)) if scene.has_bifacial else None
        counts = np.diff(seg)
        for s in np.nonzero(counts)[0]:
            n = int(counts[s])
            a, b = int(seg[s]), int(seg[s + 1])
            theta = None if theta_in is None else theta_in[a:b]
            passes = [(scene.records[s], scene.optics[s]['table'], draws)]
            if scene.has_bifacial and scene.bifacial_mask[s]:
                passes = []
                if back[a:b].any():
                    passes.append((scene.records_back[s], scene.optics[s]['back']['table'], draws_b))
                if not back[a:b].all():
                    passes.append((scene.records[s], scene.optics[s]['table'], draws))
            for rec, table, dest in passes:
                vals = self._draws_for(rec, table, n, theta)
                for j in range(3):
                    if vals[j] is not None:
                        dest[j, a:b] = vals[j]
        to_dev = (lambda x: torch.from_numpy(x).to(self.device))
        return to_dev(draws), (to_dev(draws_b) if draws_b is not None else None)

    # ------------------------------------------------------------ bounce
    def _bounce(self, scene, cur, bounce, min_energy, use_bvh):
        """One iteration of the reference loop body (tracer_engine.py:193-274).
        Returns (record tensors dict, n_survivors, H).

        Launch sequence: trc_trace_hits (nearest hit fused with the emission of
        64-byte hit records) -> trc_order_hits (surface-major / ray-minor slots)
        -> trc_optics -> trc_segment_sums [-> cull sort + gathers].  Two host
        syncs: the hit count (sizes the child bundle) and the cull count."""
        import torch
        lib = L.load()
        dev = self.device
        st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        n, S = cur.n, scene.num_surfaces
        i32, i64, f64 = torch.int32, torch.int64, torch.float64
        n_chunks = (n + L.HIT_CHUNK - 1) // L.HIT_CHUNK
        chunk_off = torch.empty(n_chunks, dtype=i32, device=dev)
        chunk_cnt = torch.empty(n_chunks, dtype=i32, device=dev)
        counters = torch.zeros(3, dtype=i64, device=dev)          # [hits, culled, void child slots]
        # capacity of the hit list: the hit fraction seen at this bounce of the previous
        # trace (+30 %), else every ray; an overflow re-runs the kernel with the exact size
        frac = self._hit_frac.get(bounce)
        cap = n if frac is None or n <= (1 << 20) else min(n, int(1.3 * frac * n) + 65536)
        while True:
            records = torch.empty((cap, L.HIT_RECORD_BYTES // 8), dtype=f64, device=dev)
            if cur.slices is not None:
                # the bundle is still arriving: trace each slice as soon as its copy has landed
                stream = torch.cuda.current_stream(dev)
                ld = cur.V.stride(0)
                for a, b, ev in cur.slices:
                    stream.wait_event(ev)
                    c0 = a // L.HIT_CHUNK
                    L.check(lib.trc_trace_hits_slice(
                        scene.handle, b - a, cur.V.data_ptr() + 8 * a, cur.D.data_ptr() + 8 * a,
                        cur.E.data_ptr() + 8 * a, ld, int(use_bvh), None, L.dptr(records), None, cap,
                        chunk_off.data_ptr() + 4 * c0, chunk_cnt.data_ptr() + 4 * c0, L.dptr(counters), a, st))
                cur.slices = cur.keep = None   # resident from here on (an overflow re-runs the plain call)
                H = int(counters[0].item())
                if H <= cap:
                    break
                cap = H
                counters.zero_()
                continue
            if self.kernel_events is not None:
                ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                ev0.record()
            L.check(lib.trc_trace_hits(scene.handle, n, L.dptr(cur.V), L.dptr(cur.D), L.dptr(cur.E),
                                       cur.V.stride(0), int(use_bvh), None, L.dptr(records), cap,
                                       L.dptr(chunk_off), L.dptr(chunk_cnt), L.dptr(counters), st))
            if self.kernel_events is not None:
                ev1.record()
            H = int(counters[0].item())
            if self.kernel_events is not None:
                self.kernel_events.append((bounce, n, H, lambda a=ev0, b=ev1: a.elapsed_time(b)))
            if H <= cap:
                break
            cap = H
        self._hit_frac[bounce] = H / float(n) if n else 0.
        if H == 0:
            return None, 0, 0
        slot_surf = torch.empty(H, dtype=i32, device=dev)
        slot_rec = torch.empty(H, dtype=i32, device=dev)
        seg = torch.empty(S + 1, dtype=i64, device=dev)
        work = torch.empty(int(lib.trc_order_work_bytes(n_chunks, H, S)), dtype=torch.uint8, device=dev)
        L.check(lib.trc_order_hits(n_chunks, L.dptr(chunk_off), L.dptr(chunk_cnt), L.dptr(records), H, S,
                                   L.dptr(slot_surf), L.dptr(slot_rec), L.dptr(seg), L.dptr(work), st))
        need_seg_host = scene.has_accountants or self.rng == 'numpy' or scene.has_split
        seg_host = seg.cpu().numpy() if need_seg_host else None
        # children: one per hit, or (PeriodicBoundary) two -- then child index != slot
        Hc, child_off, child_off_host = H, None, None
        if scene.has_split:
            counts = np.diff(seg_host)
            child_off_host = np.concatenate([[0], np.cumsum(counts * scene.children_per_hit)]).astype(np.int64)
            Hc = int(child_off_host[-1])
            child_off = torch.from_numpy(child_off_host).to(dev)

        want = set(scene.accountant_sources)
        if scene.has_refractive and scene.has_accountants:
            want.add('e_in')                             # accountants of refractive surfaces re-derive their columns
        poly = cur.SP is not None
        if poly:
            want.update(('normals', 'dirs_in'))          # trc_spectra reads the incidence angle from them
        if 'wavelengths' in want and cur.W is None:
            raise ValueError("Spectral accountant needs a 'wavelengths' bundle property")
        n_w = int(cur.SP.shape[0]) if poly else 0

        def outputs():
            """One set of child columns + accountant columns (slot order) for an optics pass."""
            o = dict(V=torch.empty((3, Hc), dtype=f64, device=dev), D=torch.empty((3, Hc), dtype=f64, device=dev),
                     E=torch.empty(Hc, dtype=f64, device=dev), P=torch.empty(Hc, dtype=i64, device=dev),
                     cull=torch.empty(Hc, dtype=i32, device=dev),
                     absorbed=torch.empty(H, dtype=f64, device=dev),   # always: feeds the per-surface totals
                     normals=torch.empty((3, Hc), dtype=f64, device=dev) if 'normals' in want else None,   # row stride = ld_out
                     e_in=torch.empty(H, dtype=f64, device=dev) if 'e_in' in want else None,
                     dirs_in=torch.empty((3, Hc), dtype=f64, device=dev) if 'dirs_in' in want else None,
                     attenuated=torch.empty(H, dtype=f64, device=dev) if 'attenuated' in want else None,
                     W=None, SP=None, lost_spectra=None,
                     R=torch.empty(Hc, dtype=f64, device=dev) if cur.R is not None else None,
                     RI=torch.empty(Hc, dtype=f64, device=dev) if (cur.R is not None and scene.has_tabulated) else None,
                     I=torch.empty(Hc, dtype=i64, device=dev))
            if poly:
                o['W'] = torch.empty((n_w, Hc), dtype=f64, device=dev)
                o['SP'] = torch.empty((n_w, Hc), dtype=f64, device=dev)
                if 'lost_spectra' in want:
                    o['lost_spectra'] = torch.empty((n_w, Hc), dtype=f64, device=dev)
            elif cur.W is not None:
                o['W'] = torch.empty(Hc, dtype=f64, device=dev)
            return o

        def optics_args(o, draws):
            a = L.OpticsArgs()
            a.h = H
            a.d_records, a.d_slot_rec = records.data_ptr(), slot_rec.data_ptr()
            a.d_wavelengths = cur.W.data_ptr() if (cur.W is not None and not poly) else None
            a.d_out_vertices, a.d_out_directions = o['V'].data_ptr(), o['D'].data_ptr()
            a.d_out_energy, a.d_out_parents, a.ld_out = o['E'].data_ptr(), o['P'].data_ptr(), Hc
            if child_off is not None:
                a.d_slot_surf, a.d_seg, a.d_child_off = slot_surf.data_ptr(), seg.data_ptr(), child_off.data_ptr()
            a.seed, a.bounce, a.min_energy = self.seed, bounce, float(min_energy)
            if cur.R is not None:
                a.d_ref_index, a.d_out_ref_index = cur.R.data_ptr(), o['R'].data_ptr()
                a.d_ref_index_im = cur.RI.data_ptr() if cur.RI is not None else None
                a.d_out_ref_index_im = o['RI'].data_ptr() if o['RI'] is not None else None
            a.d_ray_ids = cur.I.data_ptr() if cur.I is not None else None
            a.ray_id_base, a.d_out_ray_ids = cur.id_base, o['I'].data_ptr()
            for j in range(3):
                a.d_draws[j] = draws[j].data_ptr() if draws is not None else None
            a.d_absorbed = o['absorbed'].data_ptr()
            for name, field in (('normals', 'd_normals'), ('e_in', 'd_e_in'), ('dirs_in', 'd_dirs_in'),
                                ('attenuated', 'd_attenuated')):
                setattr(a, field, o[name].data_ptr() if o[name] is not None else None)
            a.d_out_wavelengths = o['W'].data_ptr() if (o['W'] is not None and not poly) else None
            a.d_cull = o['cull'].data_ptr()
            a.d_n_cull = counters[1:].data_ptr()
            a.d_n_void = counters[2:].data_ptr()
            return a

        def optics_pass(handle, o, draws):
            """trc_optics (+ trc_spectra for polychromatic rays) with the surface table behind ``handle``."""
            L.check(lib.trc_optics(handle, C.byref(optics_args(o, draws)), st))
            if poly:
                counters[1:2].zero_()                       # trc_spectra re-decides every slot's cull flag
                sa = L.SpectraArgs()
                sa.h, sa.n_samples = H, n_w
                sa.d_slot_surf, sa.d_parents = slot_surf.data_ptr(), o['P'].data_ptr()
                sa.d_in_wavelengths, sa.d_in_spectra, sa.ld_in = cur.W.data_ptr(), cur.SP.data_ptr(), cur.W.stride(0)
                sa.d_normals, sa.d_dirs_in, sa.ld_cols = o['normals'].data_ptr(), o['dirs_in'].data_ptr(), Hc
                sa.d_out_wavelengths, sa.d_out_spectra, sa.ld_out = o['W'].data_ptr(), o['SP'].data_ptr(), Hc
                sa.d_energy, sa.d_absorbed = o['E'].data_ptr(), o['absorbed'].data_ptr()
                sa.d_lost_spectra = o['lost_spectra'].data_ptr() if o['lost_spectra'] is not None else None
                sa.d_cull, sa.d_n_cull, sa.min_energy = o['cull'].data_ptr(), counters[1:].data_ptr(), float(min_energy)
                L.check(lib.trc_spectra(handle, C.byref(sa), st))

        # BiFacial managers (optics_callables.py:1885-1901): which hits are on the back of their surface
        back = back_host = back_counts = None
        if scene.has_bifacial:
            up, bif = scene.bifacial_tensors(dev)
            d_in = records.index_select(0, slot_rec.long())[:, 3:6]
            proj = (d_in * up.index_select(0, slot_surf.long())).sum(dim=1)
            back = (torch.round(proj, decimals=6) > 0.) & bif.index_select(0, slot_surf.long())
            back_counts = torch.bincount(slot_surf[back].long(), minlength=S).cpu().numpy()

        draws = draws_back = None
        if self.rng == 'numpy':
            theta_in = None
            if scene.needs_theta_for_draws:
                # the specular/diffuse split of these optics follows the incidence angle: a first
                # pass of the kernel (draws unused for normals) provides it to the host
                tmp = outputs()
                tn = tmp['normals'] if tmp['normals'] is not None else torch.empty((3, Hc), dtype=f64, device=dev)
                td = tmp['dirs_in'] if tmp['dirs_in'] is not None else torch.empty((3, Hc), dtype=f64, device=dev)
                tmp['normals'], tmp['dirs_in'] = tn, td
                L.check(lib.trc_optics(scene.handle, C.byref(optics_args(tmp, None)), st))
                counters[1:].zero_()
                hn, hd = tn[:, :H].cpu().numpy(), td[:, :H].cpu().numpy()
                vertical = np.sum(hd * hn, axis=0) * hn          # optics_callables.py:469-472
                theta_in = np.arccos(np.sqrt(np.sum(vertical ** 2, axis=0)))
                del tmp
            back_host = back.cpu().numpy() if back is not None else None
            draws, draws_back = self._numpy_draws(scene, seg_host, H, theta_in, back_host)

        out = outputs()
        optics_pass(scene.handle, out, draws)
        src_back = None
        if scene.has_bifacial:
            # the back sides' optics on every slot (same geometry, second surface table), then slot by slot the
            # side the ray is on.  Each side's accountants still see ALL the hits of their surface (below).
            out_b = outputs()
            optics_pass(scene.handle_back, out_b, draws_back)
            src_back = out_b
            merged = {}
            for k, t in out.items():
                merged[k] = None if t is None else torch.where(back, out_b[k], t)
            src_front, out = out, merged
            counters[1] = (out['cull'] == 1).sum()

        oV, oD, oE, oP, cull, oW, oR, oI, oS = (out[k] for k in ('V', 'D', 'E', 'P', 'cull', 'W', 'R', 'I', 'SP'))
        oRI = out['RI']
        absorbed = out['absorbed']
        # per-surface absorbed energy and hit counts, kept on device: the payload of
        # the multi-GPU all-reduce (tracer_engine_mp.py) and a tally for 'fast' optics
        L.check(lib.trc_segment_sums(H, L.dptr(slot_surf), L.dptr(absorbed), L.dptr(self.absorbed_per_surface), st))
        self.hits_per_surface += seg[1:] - seg[:-1]
        if scene.has_accountants:
            def columns(o):
                return dict(hits=o['V'], absorbed=o['absorbed'], e_out=o['E'], normals=o['normals'], e_in=o['e_in'],
                            dirs_in=o['dirs_in'], wavelengths=o['W'], attenuated=o['attenuated'],
                            lost_spectra=o['lost_spectra'], cull=o['cull'], parents=o['P'], w_in=cur.W)
            if scene.has_bifacial:
                self._account(scene, seg_host, columns(src_front), child_off_host, back=(columns(src_back), back_counts))
            else:
                self._account(scene, seg_host, columns(out), child_off_host)

        n_cull, n_void = (int(x) for x in counters[1:].tolist())
        del records, work

        def reorder(keys, n_keys):
            # stable counting sort of the children by key through the library's radix passes
            k2 = torch.empty(Hc, dtype=i32, device=dev)
            perm = torch.empty(Hc, dtype=i32, device=dev)
            count = torch.zeros(1, dtype=i64, device=dev)
            w = torch.empty(int(lib.trc_sort_work_bytes(Hc, n_keys)), dtype=torch.uint8, device=dev)
            L.check(lib.trc_sort_pairs(Hc, n_keys, L.dptr(keys), None, L.dptr(k2), L.dptr(perm), L.dptr(count),
                                       L.dptr(w), st))
            pV, pD = torch.empty_like(oV), torch.empty_like(oD)
            pE, pP = torch.empty_like(oE), torch.empty_like(oP)
            L.check(lib.trc_gather_rows(Hc, L.dptr(perm), 3, L.dptr(oV), Hc, L.dptr(pV), Hc, None, None, st))
            L.check(lib.trc_gather_rows(Hc, L.dptr(perm), 3, L.dptr(oD), Hc, L.dptr(pD), Hc, None, None, st))
            L.check(lib.trc_gather_rows(Hc, L.dptr(perm), 1, L.dptr(oE), Hc, L.dptr(pE), Hc, L.dptr(oP), L.dptr(pP), st))
            pl = perm.long()
            pW = oW.index_select(-1, pl) if oW is not None else None
            pS = oS.index_select(-1, pl) if oS is not None else None
            pR = oR.index_select(0, pl) if oR is not None else None
            pI = oI.index_select(0, pl)
            if oRI is not None:
                reorder.RI = oRI.index_select(0, pl)
            return pV, pD, pE, pP, pW, pR, pI, pS, perm
        reorder.RI = None

        if scene.has_bifacial:
            # children of a BiFacial surface: the back side's first, each side in hit order (:1891-1899)
            key = 2 * slot_surf + (~back).to(i32) * scene.bifacial_tensors(dev)[1].index_select(0, slot_surf.long()).to(i32)
            oV, oD, oE, oP, oW, oR, oI, oS, perm = reorder(key.contiguous(), 2 * S)
            oRI = reorder.RI
            cull = cull.index_select(0, perm.long()).contiguous()
        if n_void:
            # refractive surfaces leave void child slots (cull code 2): drop them, keeping the order
            is_void = (cull == 2).to(i32)
            pV, pD, pE, pP, pW, pR, pI, pS, perm = reorder(is_void, 2)
            Hc -= n_void
            cull = cull.index_select(0, perm[:Hc].long()).contiguous()
            oV, oD = pV[:, :Hc].contiguous(), pD[:, :Hc].contiguous()
            oE, oP = pE[:Hc].contiguous(), pP[:Hc].contiguous()
            oW = pW[..., :Hc].contiguous() if pW is not None else None
            oS = pS[..., :Hc].contiguous() if pS is not None else None
            oR = pR[:Hc].contiguous() if pR is not None else None
            oRI = reorder.RI[:Hc].contiguous() if reorder.RI is not None else None
            oI = pI[:Hc].contiguous()
            if Hc == 0:
                return None, 0, H
        rec = dict(V=oV, D=oD, E=oE, P=oP, W=oW, R=oR, I=oI, SP=oS, RI=oRI)
        if n_cull and n_cull < Hc:
            # survivors first, weak rays at the back (tracer_engine.py:242-245, 268-274)
            pV, pD, pE, pP, pW, pR, pI, pS, _ = reorder(cull, 2)
            rec = dict(V=pV, D=pD, E=pE, P=pP, W=pW, R=pR, I=pI, SP=pS, RI=reorder.RI)
        rec['pre_cull'] = dict(V=oV, D=oD, E=oE, P=oP, W=oW, R=oR, SP=oS, RI=oRI)
        return rec, Hc - n_cull, H

    # ------------------------------------------------- device-driven trace
    def _plan_block(self, scene, cur, bounce, max_bounces):
        """How many bounces to enqueue from ``bounce`` on and the hit capacity of each, from what the previous
        trace of this scene did (``self._hist``).  Without two-children optics a bounce cannot hit more rays
        than entered it, so the capacity of the bounce before is a hard bound."""
        n = cur.n
        hist = self._hist
        full = n <= FULL_CAP_RAYS
        caps, n_in_est, bound = [], float(n), n
        depth_known = bounce in hist
        for j in range(min(max_bounces, L.TRACE_MAX_BOUNCES)):
            h = hist.get(bounce + j)
            if h is None or h[0] == 0:
                if depth_known and j > 0:
                    break                       # the last trace ended here; a surprise survivor starts a new block
                if j >= BLIND_BOUNCES or (j > 0 and not full):
                    break                       # unknown territory: worst-case buffers, so few bounces at a time
                est_h = float(bound)            # nothing known: every ray may hit
                surv = 1.
            else:
                est_h = n_in_est * h[1] / h[0]
                surv = h[2] / float(h[1]) if h[1] else 0.
            cap = bound if full else min(bound, int(1.3 * est_h) + 4096)
            cap = _round_cap(max(cap, min(bound, self._cap_floor.get(bounce + j, 0))))
            caps.append(cap)
            n_in_est = est_h * surv
            bound = cap
            if depth_known and h is not None and h[2] == 0:
                break                           # nothing survived this bounce last time
        return caps

    def _count_hits(self, scene, cur, use_bvh):
        """Hit count of the bundle's first bounce alone (a nearest-hit pass that emits no records): sizes the
        buffers of a large first trace exactly instead of for the worst case."""
        import torch
        lib = L.load()
        dev = self.device
        st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        n = cur.n
        n_chunks = (n + L.HIT_CHUNK - 1) // L.HIT_CHUNK
        chunk = torch.empty(2 * n_chunks, dtype=torch.int32, device=dev)
        total = torch.zeros(1, dtype=torch.int64, device=dev)
        L.check(lib.trc_trace_hits(scene.handle, n, L.dptr(cur.V), L.dptr(cur.D), L.dptr(cur.E), cur.V.stride(0),
                                   int(use_bvh), None, None, 0, L.dptr(chunk), chunk.data_ptr() + 4 * n_chunks,
                                   L.dptr(total), st))
        return int(total.item())

    def _trace_block(self, scene, cur, bounce, max_bounces, min_energy, use_bvh):
        """Up to ``max_bounces`` iterations of the reference loop body through trc_trace: one foreign call
        enqueues every kernel, one stream synchronisation, one read-back of all counts.  Returns a list of
        (record dict or None, survivors, hits, rays in), one entry per bounce that ran on valid data; an
        overflowing bounce (more hits than planned) ends the list early and raises its capacity floor, so the
        caller's next block starts with it."""
        import torch
        lib = L.load()
        dev = self.device
        stream = torch.cuda.current_stream(dev)
        st = C.c_void_p(stream.cuda_stream)
        n, S = cur.n, scene.num_surfaces
        want = scene.accountant_sources
        if 'wavelengths' in want and cur.W is None:
            raise ValueError("Spectral accountant needs a 'wavelengths' bundle property")
        cols = ((L.COL_NORMALS if 'normals' in want else 0) | (L.COL_E_IN if 'e_in' in want else 0) |
                (L.COL_DIRS_IN if 'dirs_in' in want else 0) | (L.COL_ATTENUATED if 'attenuated' in want else 0))
        if bounce not in self._hist and bounce not in self._cap_floor and n > FULL_CAP_RAYS and cur.slices is None:
            self._cap_floor[bounce] = self._count_hits(scene, cur, use_bvh)
            self._hist[bounce] = (n, self._cap_floor[bounce], self._cap_floor[bounce])
        caps = self._plan_block(scene, cur, bounce, max_bounces)
        nb = len(caps)
        a, plan = L.TraceArgs(), L.TracePlan()
        a.n, a.ld = n, cur.V.stride(0)
        a.d_vertices, a.d_directions, a.d_energy = cur.V.data_ptr(), cur.D.data_ptr(), cur.E.data_ptr()
        a.d_wavelengths = cur.W.data_ptr() if cur.W is not None else None
        a.d_ray_ids = cur.I.data_ptr() if cur.I is not None else None
        a.ray_id_base = cur.id_base
        a.use_bvh, a.first_bounce, a.n_bounces = int(use_bvh), bounce, nb
        timed = self.kernel_events is not None and bounce == 0 and cur.slices is None
        seg_on_host = scene.has_accountants and S + 1 <= INLINE_SEG_MAX
        a.flags = (L.TRACE_SEG if seg_on_host else 0) | (L.TRACE_TIME_WALK if timed else 0)
        if self._tallies_dirty:
            a.flags |= L.TRACE_ZERO_TALLIES     # the trace's first launch zeroes the per-surface tallies
            self._tallies_dirty = False
        a.seed, a.min_energy, a.columns = self.seed, float(min_energy), cols
        for j, c in enumerate(caps):
            a.capacity[j] = c
        L.check(lib.trc_trace_plan(scene.handle, C.byref(a), C.byref(plan)))
        keep = torch.empty(int(plan.keep_bytes), dtype=torch.uint8, device=dev)
        scratch = torch.empty(int(plan.scratch_bytes), dtype=torch.uint8, device=dev)
        words = int(plan.counts_words)
        if self._counts_host is None or self._counts_host.numel() < words:
            self._counts_host = torch.empty(max(words, 64), dtype=torch.int64, pin_memory=True)
        counts = self._counts_host
        a.d_keep, a.keep_bytes = keep.data_ptr(), keep.numel()
        a.d_scratch, a.scratch_bytes = scratch.data_ptr(), scratch.numel()
        a.d_absorbed_per_surface = self.absorbed_per_surface.data_ptr()
        a.d_hits_per_surface = self.hits_per_surface.data_ptr()
        a.h_counts = counts.data_ptr()
        if cur.slices is not None:
            # the bundle is still arriving: run the first nearest-hit pass slice by slice, as each copy lands
            a.flags |= L.TRACE_WALK_DONE
            L.check(lib.trc_trace_init(scene.handle, C.byref(a), st))
            sp = scratch.data_ptr()
            ld = cur.V.stride(0)
            for lo, hi, ev in cur.slices:
                stream.wait_event(ev)
                c0 = lo // L.HIT_CHUNK
                L.check(lib.trc_trace_hits_slice(
                    scene.handle, hi - lo, cur.V.data_ptr() + 8 * lo, cur.D.data_ptr() + 8 * lo,
                    cur.E.data_ptr() + 8 * lo, ld, int(use_bvh), None, sp + plan.off_records,
                    sp + plan.off_keys, caps[0], sp + plan.off_chunk_off + 4 * c0, sp + plan.off_chunk_cnt + 4 * c0,
                    sp + plan.off_total, lo, st))
            cur.slices = cur.keep = None       # resident from here on
        L.check(lib.trc_trace(scene.handle, C.byref(a), st))
        stream.synchronize()                    # the only synchronisation of the block
        per = 4 + (S + 1 if seg_on_host else 0)
        host = counts[:nb * per].numpy()
        k64, ki64 = keep.view(torch.float64), keep.view(torch.int64)
        f64, i64 = k64, ki64
        out = []
        for j in range(nb):
            n_in, H, n_surv, reorder = (int(x) for x in host[per * j:per * j + 4])
            if timed and j == 0:
                ms = C.c_float(0.)
                L.check(lib.trc_trace_walk_ms(scene.handle, C.byref(ms)))
                self.kernel_events.append((0, n_in, H, lambda v=ms.value: v))
            if H > caps[j]:
                # more hits than planned: nothing of this bounce is valid; the next block starts here, larger
                self._cap_floor[bounce + j] = H
                self._hist[bounce + j] = (n_in, H, H)
                self.stats_overflows = getattr(self, 'stats_overflows', 0) + 1
                break
            self._hist[bounce + j] = (n_in, H, n_surv)
            self._hit_frac[bounce + j] = H / float(n_in) if n_in else 0.
            if H == 0:
                out.append((None, 0, 0, n_in))
                break
            lay = plan.bounce[j]
            ld = int(lay.ld)

            def view(off, base, rows, lay_ld=ld, h=H):
                if off < 0:
                    return None
                if rows:
                    return torch.as_strided(base, (rows, h), (lay_ld, 1), off >> 3)
                return torch.as_strided(base, (h,), (1,), off >> 3)
            pre = dict(V=view(lay.off_vertices, f64, 3), D=view(lay.off_directions, f64, 3),
                       E=view(lay.off_energy, f64, 0), P=view(lay.off_parents, i64, 0),
                       W=view(lay.off_wavelengths, f64, 0), I=view(lay.off_ray_ids, i64, 0))
            first = pre
            if reorder:        # some but not all children are weak: the survivors-first copy was made
                first = dict(V=view(lay.off_alt_vertices, f64, 3), D=view(lay.off_alt_directions, f64, 3),
                             E=view(lay.off_alt_energy, f64, 0), P=view(lay.off_alt_parents, i64, 0),
                             W=view(lay.off_alt_wavelengths, f64, 0), I=view(lay.off_alt_ray_ids, i64, 0))
            if scene.has_accountants:
                if seg_on_host:
                    seg = host[per * j + 4:per * (j + 1)]
                else:
                    seg_dev = torch.as_strided(ki64, (S + 1,), (1,), lay.off_seg >> 3)
                    need = scene.seg_entries_needed(dev)
                    vals = seg_dev.index_select(0, need[1]).cpu().tolist()
                    seg = _DeviceSeg(seg_dev, dict(zip(need[0], vals)))
                src = dict(hits=pre['V'], e_out=pre['E'], wavelengths=pre['W'])
                for name, off, rows in (('absorbed', lay.off_absorbed, 0), ('normals', lay.off_normals, 3),
                                        ('e_in', lay.off_e_in, 0), ('dirs_in', lay.off_dirs_in, 3),
                                        ('attenuated', lay.off_attenuated, 0)):
                    if name in want:
                        src[name] = view(off, f64, rows)
                self._account(scene, seg, src, copy=not self._keep_tree)
            rec = dict(first)
            rec['pre_cull'] = dict(V=pre['V'], D=pre['D'], E=pre['E'], P=pre['P'], W=pre['W'])
            out.append((rec, n_surv, H, n_in))
            if n_surv == 0:
                break
        return out

    @staticmethod
    def _account(scene, seg, src, child_off=None, copy=True, back=None):
        """What Accountant.count() stores, one chunk per surface per bounce
        (optics_callables.py:1597-1790), kept in HBM until read.  ``src`` maps a
        source name to the tensor the optics kernel filled: slot order, except the
        child columns (hit points, outgoing energy) which follow ``child_off`` when
        some surface of the scene emits two children per hit.  ``copy=False`` stores views of the
        bounce's arena instead of copies (used when the ray tree keeps that arena alive anyway)."""
        take = (lambda t: t.clone()) if copy else (lambda t: t)
        for i0, i1, mesh in scene.mesh_segments:           # one chunk per bounce for a whole mesh
            a, b = int(seg[i0]), int(seg[i1])
            names = set(name for _, _, name in scene.optics[i0]['accountants'])
            if b == a or not names:
                continue
            if child_off is not None:
                raise NotImplementedError("TriangleMeshObject in a scene with two-children optics")
            part = seg[i0:i1 + 1]
            if not isinstance(part, _DeviceSeg):
                part = np.array(part, dtype=np.int64)
            mesh._add_chunk(part, {k: take(src[k][..., a:b]) for k in names})
        for s in scene.account_list:
            orec = scene.optics[s]
            a, b = int(seg[s]), int(seg[s + 1])
            if b == a:
                continue
            feeds = [(orec['accountants'], src)]
            if orec.get('back') is not None:
                # BiFacial: a side's manager ran (on every hit of the surface) iff some hit was on that side
                n_back = int(back[1][s])
                feeds = ([(orec['back']['accountants'], back[0])] if n_back > 0 else []) + \
                        ([(orec['accountants'], src)] if n_back < b - a else [])
            if orec['okind'] == L.O_REFRACTIVE:
                # single_ray refraction (one child per hit): the child bundle as the callable returns it is
                # [reflected in hit order][refracted in hit order]; count() subtracts it AS IS from the hit-ordered
                # incoming energies (optics_callables.py:1638-1643)
                lo, h = int(child_off[s]), b - a
                blk = src['e_out'][lo:lo + 2 * h]
                e_child = blk[src['cull'][lo:lo + 2 * h] != 2]
                for acc, field, name in orec['accountants']:
                    if name == 'absorbed':
                        t = src['e_in'][a:b] - e_child
                    elif name == 'e_out':
                        t = e_child
                    elif name == 'hits':
                        t = src['hits'][:, lo:lo + h]
                    elif name == 'wavelengths':
                        # SpectralAccountant.count reads the INCOMING rays' wavelengths in hit order (:1811); the
                        # children's differ in single_ray mode (inherited by rank, :797-815)
                        t = src['w_in'].index_select(0, src['parents'][lo:lo + h])
                    else:                                   # e_in, dirs_in, normals: slot order
                        t = src[name][..., a:b]
                    getattr(acc, field).append(_LazyRows(t.clone()))
                continue
            for accs, cols in feeds:
                for acc, field, name in accs:
                    t = cols[name]
                    if child_off is not None and name in ('hits', 'e_out', 'wavelengths'):
                        lo = int(child_off[s])
                        getattr(acc, field).append(_LazyRows(take(t[..., lo:lo + (b - a)])))
                    else:
                        getattr(acc, field).append(_LazyRows(take(t[..., a:b])))

    @staticmethod
    def _as_bundle(rec, n=None):
        def cols(t):
            if t is None:
                return None
            return _DeviceColumns([t if n is None else t[..., :n]])
        kw = {}
        if rec.get('W') is not None:
            kw['wavelengths'] = cols(rec['W'])
        if rec.get('R') is not None:
            if rec.get('RI') is not None:       # complex index (media given by material tables)
                import torch
                kw['ref_index'] = cols(torch.complex(rec['R'], rec['RI']))
            else:
                kw['ref_index'] = cols(rec['R'])
        if rec.get('SP') is not None:
            kw['spectra'] = cols(rec['SP'])
        return RayBundle(vertices=cols(rec['V']), directions=cols(rec['D']), energy=cols(rec['E']),
                         parents=cols(rec['P']), **kw)

    # ------------------------------------------------------------- trace
    def ray_tracer(self, bundle, reps=100, min_energy=1e-10, tree=True, accel=False, Kd_Tree=None, **kwargs):
        """Trace ``bundle`` through the assembly; same contract as
        tracer_engine.py:124-295.  Returns (vertices, directions) of the last
        live bundle as NumPy arrays."""
        import torch
        self.reps = reps
        self.minener = min_energy
        self._keep_tree = bool(tree)      # the tree's bundles view the per-bounce arenas: accountant rows may too
        self.tree = RayTree()
        if tree is True:
            self.tree.append(bundle)
        with torch.cuda.device(self.device):      # scene tables, BVH and every launch on the engine's device
            return self._ray_tracer(bundle, reps, min_energy, tree, accel, Kd_Tree)

    def _ray_tracer(self, bundle, reps, min_energy, tree, accel, Kd_Tree):
        import torch
        scene = self.compile()
        self.Kd_Tree = Kd_Tree if Kd_Tree is not None else (scene if accel else None)
        if self.device_accel == 'bvh' or accel:
            use_bvh = True
        elif self.device_accel == 'brute':
            use_bvh = False
        else:
            use_bvh = scene.num_surfaces > 16
        if scene.needs_wavelengths and not bundle.has_property('wavelengths'):
            raise ValueError("scene has spectral optics but the bundle carries no 'wavelengths'")

        cur = self._to_device(bundle)
        # a bundle of nearly parallel rays (a solar source) finds its candidates in the scene's shadow grid
        grid = False
        if use_bvh and self.shadow_grid and cur.n >= GRID_MIN_RAYS:
            cone = _direction_cone(bundle)
            grid = cone is not None and scene.set_shadow_grid(cone[0], cone[1])
        if scene.has_refractive and cur.R is None:
            raise ValueError("scene has refractive optics but the bundle carries no 'ref_index'")
        if scene.has_tabulated and cur.RI is None:
            import torch as _t
            cur.RI = _t.zeros_like(cur.R)
        source_E = cur.E
        # device-driven (trc_trace) unless draws are injected or some optics emits two children per hit
        if scene.needs_spectra and cur.SP is None:
            raise ValueError("scene has polychromatic optics but the bundle carries no 'spectra' (with 'wavelengths', both (W, n))")
        if cur.SP is not None and (scene.needs_wavelengths or scene.has_split or
                                   bool((scene.records['flags'] & L.F_IAM).any())):
            raise NotImplementedError("polychromatic bundle (a 'spectra' property) in a scene with spectral, IAM, periodic "
                                      "or refractive optics")
        device_driven = (self.fused_bounce and self.rng == 'philox' and not scene.has_split and cur.R is None
                         and cur.SP is None and not scene.has_bifacial)
        if device_driven and cur.n > 0 and reps > 0:
            tall = torch.empty(2 * scene.num_surfaces, dtype=torch.float64, device=self.device)
            self.absorbed_per_surface = tall[:scene.num_surfaces]
            self.hits_per_surface = tall[scene.num_surfaces:].view(torch.int64)
            self._tallies_dirty = True
        else:
            self.absorbed_per_surface = torch.zeros(scene.num_surfaces, dtype=torch.float64, device=self.device)
            self.hits_per_surface = torch.zeros(scene.num_surfaces, dtype=torch.int64, device=self.device)
            self._tallies_dirty = False
        ray_bounces = 0
        per_bounce = []
        last_rec = None
        i = 0
        depleted = False
        while i < reps and not depleted:
            how = (2 if grid and i == 0 else 1) if use_bvh else 0      # the use_bvh argument of the C entry points
            if device_driven:
                results = self._trace_block(scene, cur, i, reps - i, min_energy, how)
            else:
                rec, n_surv, H = self._bounce(scene, cur, i, min_energy, how)
                results = [(rec, n_surv, H, cur.n)]
            for rec, n_surv, H, n_in in results:
                i += 1
                ray_bounces += n_in
                per_bounce.append((n_in, H, n_surv))
                if rec is None:
                    last_rec = None
                    cur = _DevBundle(cur.V[:, :0], cur.D[:, :0], cur.E[:0], None if cur.W is None else cur.W[..., :0])
                    logging.log(self.loglevel, 'Ray bundle depleted')
                    depleted = True
                    break
                last_rec = rec
                if tree:
                    full = dict(rec)
                    full.pop('pre_cull')
                    self.tree.append(self._as_bundle(full))
                W = rec['W'][..., :n_surv] if rec['W'] is not None else None
                SP = rec['SP'][..., :n_surv] if rec.get('SP') is not None else None
                R = rec['R'][:n_surv] if rec.get('R') is not None else None
                RI = rec['RI'][:n_surv] if rec.get('RI') is not None else None
                I = rec['I'][:n_surv] if rec.get('I') is not None else None
                cur = _DevBundle(rec['V'][:, :n_surv], rec['D'][:, :n_surv], rec['E'][:n_surv], W, R=R, I=I, SP=SP, RI=RI)
                if cur.n == 0:
                    logging.log(self.loglevel, 'Ray bundle depleted')
                    depleted = True
                    break
        if not tree:
            # only the last record, in pre-cull order (tracer_engine.py:288-291)
            self.tree.append(self._as_bundle(last_rec['pre_cull']) if last_rec is not None
                             else RayBundle.empty_bund())
        self.stats = dict(ray_bounces=ray_bounces, per_bounce=per_bounce, surfaces=scene.num_surfaces,
                          use_bvh=use_bvh, shadow_grid=bool(grid))
        if cur.n != 0:
            logging.warning(f'{cur.n} rays left at the end of the simulation')
            frac = float((cur.E.sum() / source_E.sum()).item()) * 100.
            logging.warning(f'Remaining energy in last bundle: {frac}%')
        return self._to_host(cur.V), self._to_host(cur.D)
