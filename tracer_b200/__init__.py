"""tracer_b200 -- B200-native Monte-Carlo bounce loop behind Tracer's API.

Scene description (Assembly, AssembledObject, Surface, geometry managers,
optics_callables, RayBundle, sources) mirrors casselineau/Tracer's ``tracer``
package; ``TracerEngine.ray_tracer`` runs on sm_100a kernels through
libtracer_b200.so (include/tracer_b200.h).  Importing this package does not
need a GPU; constructing a TracerEngine does.
"""
__all__ = ['assembly', 'object', 'surface', 'optics_callables', 'ray_bundle', 'sources', 'spatial_geometry',
           'tracer_engine', 'trace_tree']
