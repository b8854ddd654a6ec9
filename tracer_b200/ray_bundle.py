"""RayBundle with the reference's API (tracer/ray_bundle.py:6-223), host- or
device-backed.

Properties are named arrays whose last axis is the ray index: ``vertices``
(3,n), ``directions`` (3,n), ``energy`` (n,), ``parents`` (n,), ``ref_index``
(n,) plus arbitrary extras (``wavelengths`` ...).  Accessors ``get_<prop>
(selector=None)`` / ``set_<prop>(value, selector=None)`` are resolved
dynamically instead of being installed per instance.

A property value may be a NumPy array or a *lazy device segment list*: the GPU
engine records each bounce as [survivors | culled] pieces that live in HBM and
are only copied to the host (and concatenated) the first time somebody reads
them -- ``engine.tree[k].get_vertices()`` looks exactly like the reference's.
"""
import numpy as np


class _DeviceColumns(object):
    """Lazy value of one property: pieces of torch CUDA tensors, concatenated
    along the ray axis on first host access."""

    def __init__(self, pieces):
        self.pieces = [p for p in pieces if p is not None]

    def num(self):
        return int(sum(p.shape[-1] for p in self.pieces))

    def to_host(self):
        import torch
        if len(self.pieces) == 1:
            return self.pieces[0].cpu().numpy()
        return torch.cat(self.pieces, dim=-1).cpu().numpy()


class RayBundle(object):
    _BASE = ('vertices', 'directions', 'energy', 'parents', 'ref_index')

    def __init__(self, vertices=None, directions=None, energy=None, parents=None, ref_index=None, **kwds):
        self.__dict__['_props'] = {}
        self.__dict__['_check_attr'] = []
        vals = dict(vertices=vertices, directions=directions, energy=energy, parents=parents,
                    ref_index=ref_index)
        vals.update(kwds)
        for name, v in vals.items():
            self._create_property(name, v)

    # -- dynamic accessors --------------------------------------------------
    def _create_property(self, propname, init_val):
        if '_' + propname not in self._check_attr:
            self._check_attr.append('_' + propname)
        if init_val is not None:
            self._props[propname] = init_val

    def _host(self, name):
        v = self._props[name]
        if isinstance(v, _DeviceColumns):
            v = v.to_host()
            self._props[name] = v
        return v

    def __getattr__(self, attr):
        if attr.startswith('get_'):
            name = attr[4:]
            if '_' + name in self.__dict__.get('_check_attr', ()):
                def getter(selector=None, _n=name):
                    if _n not in self._props:
                        raise KeyError('_' + _n)
                    v = self._host(_n)
                    return v if selector is None else v[..., selector]
                return getter
        elif attr.startswith('set_'):
            name = attr[4:]
            if '_' + name in self.__dict__.get('_check_attr', ()):
                def setter(new_val, selector=None, _n=name):
                    if selector is None:
                        self._props[_n] = new_val
                    else:
                        self._host(_n)[..., selector] = new_val
                return setter
        elif attr.startswith('_') and attr[1:] in self.__dict__.get('_props', {}):
            return self._host(attr[1:])       # reference code touches e.g. bundle._spectra
        raise AttributeError(attr)

    def has_property(self, propname):
        return '_' + propname in self._check_attr

    def is_set(self, propname):
        return propname in self._props

    def get_num_rays(self):
        v = self._props['vertices']
        return v.num() if isinstance(v, _DeviceColumns) else v.shape[1]

    def property_names(self):
        return [a[1:] for a in self._check_attr]

    # -- bundle algebra -----------------------------------------------------
    def inherit(self, selector=np.s_[:], vertices=None, direction=None, energy=None, parents=None,
                ref_index=None, **kwds):
        """New bundle: given properties as passed, the others copied from this
        bundle at ``selector`` (ray_bundle.py:117-143; note the ``direction`` spelling)."""
        vals = dict((name, None) for name in self.property_names())
        vals.update(vertices=vertices, directions=direction, energy=energy, parents=parents,
                    ref_index=ref_index)
        vals.update(kwds)
        for name in list(vals):
            if vals[name] is None and name in self._props:
                vals[name] = self._host(name)[..., selector]
        return RayBundle(**vals)

    def __add__(self, added):
        out = RayBundle()
        for attr in self._check_attr:
            name = attr[1:]
            if name in self._props and name in added._props:
                out._create_property(name, np.hstack((self._host(name), added._host(name))))
        return out

    @staticmethod
    def empty_bund():
        e = np.zeros((3, 0))
        return RayBundle(vertices=e, directions=e.copy(), energy=np.array([]),
                         parents=np.array([], dtype=int), ref_index=np.array([]))

    def delete_rays(self, selector):
        n = self.get_num_rays()
        if selector is None:
            selector = np.arange(n)
        return self.inherit(np.delete(np.arange(n), selector))


def concatenate_rays(bundles):
    """Merge bundles that share a property set (ray_bundle.py:197-223)."""
    if len(bundles) == 0:
        return RayBundle.empty_bund()
    out = RayBundle()
    for attr in bundles[0]._check_attr:
        name = attr[1:]
        if name in bundles[0]._props:
            out._create_property(name, np.hstack([b._host(name) for b in bundles]))
    return out
