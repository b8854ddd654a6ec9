"""Stub of the `shapely` names the reference imports at module top
(ray_trace_utils/sampling.py:3, tracer/polygon.py:4).  They are only used by
mesh()/polygon_sampling, never on the trace path."""


class Polygon(object):
    def __init__(self, *a, **k):
        raise NotImplementedError("shapely stub")


class MultiPolygon(Polygon):
    pass


def constrained_delaunay_triangles(*a, **k):
    raise NotImplementedError("shapely stub")
