"""Printing-medium containers ("vials").  Mirrors ``src/drtvam/geometry.py``:
same classes, parameters, defaults, errors and ``to_dict()`` keys
(``Container`` :4-72, ``IndexMatchedVial`` :75-96, ``CylindricalVial`` :142-183,
``SquareVial`` :186-219, ``DoubleCylindricalVial`` :222-308, registry :312-318).

The dicts describe the same shapes as the reference's Mitsuba dicts; here they
are lowered to closed-form interface kernels (``_scene.surface_from_shape_dict``)
instead of OptiX/Embree geometry.
"""
from __future__ import annotations

from .transform import ScalarTransform4f

MEDIUM_REF = {"type": "ref", "id": "printing_medium"}


def _tube(height, radius, bsdf, **extra):
    """An open cylinder about z, centred on the origin (no end caps)."""
    d = {'type': 'cylinder', 'p0': [0., 0., -0.5 * height], 'p1': [0., 0., 0.5 * height],
         'radius': radius, 'bsdf': bsdf}
    d.update(extra)
    return d


def _glass(int_ior, ext_ior=None):
    b = {'type': 'dielectric', 'int_ior': int_ior}
    if ext_ior is not None:
        b['ext_ior'] = ext_ior
    return b


class Container:
    """Base class: medium properties (IOR, extinction, albedo, phase) + occlusions."""

    def __init__(self, params):
        if 'medium' not in params.keys():
            raise ValueError(f"[{self.__class__.__name__}] Missing field 'medium'.")
        medium = params['medium']
        self.medium_ior = medium['ior']
        self.sigma_t = medium['extinction']
        self.albedo = medium['albedo']
        self.occlusions = params.get('occlusions', [])

        if 'phase' in medium.keys():
            self.medium_phase = medium['phase']
        elif self.albedo > 0.:
            raise ValueError(f"[{self.__class__.__name__}] Tried to load a scattering medium without specifying a phase function.")
        else:
            self.medium_phase = None

    def medium_dict(self):
        medium_dict = {
            'type': 'homogeneous',
            'sigma_t': self.sigma_t,
            'albedo': self.albedo,
        }
        if self.medium_phase is not None:
            medium_dict['phase'] = self.medium_phase
        return medium_dict

    def to_dict(self):
        raise NotImplementedError

    def add_occlusions(self, dd):
        default_bsf = {"type": "diffuse", "reflectance": {"type": "spectrum", "value": 0.0}}
        for occlusion in self.occlusions:
            dd["occlusion" + occlusion["filename"].replace("/", "_").replace(".", "_")] = {
                'type': 'ply',
                "face_normals": occlusion.get("face_normals", True),
                'filename': occlusion["filename"],
                'bsdf': occlusion.get("bsdf", default_bsf),
                'exterior': {"type": "ref", "id": "printing_medium"}
            }
        return dd


class IndexMatchedVial(Container):
    def __init__(self, params):
        super().__init__(params)
        self.r = params['r']
        self.height = params.get('height', 40.)

    def to_dict(self):
        return self.add_occlusions({
            'printing_medium': self.medium_dict(),
            'vial_exterior': _tube(self.height, self.r, {'type': 'null'}, interior=MEDIUM_REF),
        })


class CustomVial(Container):
    def __init__(self, params):
        super().__init__(params)
        if "filename_vial_outer" not in params.keys() or "filename_vial_inner" not in params.keys():
            raise ValueError(f"[{self.__class__.__name__}] Missing fields 'filename_vial_outer' or 'filename_vial_inner' for custom vial.")
        self.vial_ior = params['ior']
        self.filename_vial_outer = params["filename_vial_outer"]
        self.filename_vial_inner = params["filename_vial_inner"]

    def to_dict(self):
        d = {
            'printing_medium': self.medium_dict(),
            'vial_exterior': {
                'type': 'ply', "face_normals": True, 'filename': self.filename_vial_outer,
                'bsdf': {'type': 'dielectric', 'int_ior': self.vial_ior, 'ext_ior': "air"},
            },
            'vial_interior': {
                'type': 'ply', 'filename': self.filename_vial_inner, "face_normals": True,
                'bsdf': {'type': 'dielectric', 'ext_ior': self.vial_ior, 'int_ior': self.medium_ior},
                'interior': {"type": "ref", "id": "printing_medium"},
            }
        }
        return self.add_occlusions(d)


class CylindricalVial(Container):
    def __init__(self, params):
        super().__init__(params)
        self.r_int = params['r_int']
        self.r_ext = params['r_ext']
        self.height = params.get('height', 40.)
        self.vial_ior = params['ior']

    def to_dict(self):
        return self.add_occlusions({
            'printing_medium': self.medium_dict(),
            'vial_exterior': _tube(self.height, self.r_ext, _glass(self.vial_ior, "air")),
            'vial_interior': _tube(self.height, self.r_int, _glass(self.medium_ior, self.vial_ior),
                                   interior=MEDIUM_REF),
        })


class SquareVial(Container):
    def __init__(self, params):
        super().__init__(params)
        self.w_int = params['w_int']
        self.w_ext = params['w_ext']
        self.height = params.get('height', 100.)
        self.vial_ior = params['ior']

    def to_dict(self):
        d = {
            'printing_medium': self.medium_dict(),
            'vial_exterior': {
                'type': 'cube',
                'to_world': ScalarTransform4f().scale((0.5 * self.w_ext, 0.5 * self.w_ext, 0.5 * self.height)),
                'bsdf': {'type': 'dielectric', 'int_ior': self.vial_ior},
            },
            'vial_interior': {
                'type': 'cube',
                'to_world': ScalarTransform4f().scale((0.5 * self.w_int, 0.5 * self.w_int, 0.5 * 0.9 * self.height)),
                'bsdf': {'type': 'dielectric', 'ext_ior': self.vial_ior, 'int_ior': self.medium_ior},
                'interior': {"type": "ref", "id": "printing_medium"},
            }
        }
        return self.add_occlusions(d)


class DoubleCylindricalVial(Container):
    def __init__(self, params):
        super().__init__(params)
        self.r_ext_outer = params['r_ext_outer']
        self.r_int_outer = params['r_int_outer']
        self.r_ext_inner = params['r_ext_inner']
        self.r_int_inner = params['r_int_inner']
        self.height = params.get('height', 40.)
        self.vial_ior_inner = params['ior_inner']
        self.vial_ior_outer = params['ior_outer']
        self.inside_inner_ior = params['ior_inside_inner']

    def to_dict(self):
        h = self.height
        return self.add_occlusions({
            'printing_medium': self.medium_dict(),
            'outer_vial': _tube(h, self.r_ext_outer, _glass(self.vial_ior_outer)),
            'outer_vial_interior': _tube(h, self.r_int_outer, _glass(self.medium_ior, self.vial_ior_outer),
                                         interior=MEDIUM_REF),
            'inner_vial': _tube(h, self.r_ext_inner, _glass(self.vial_ior_inner, self.medium_ior),
                                exterior=MEDIUM_REF),
            'inner_vial_interior': _tube(h, self.r_int_inner, _glass(self.inside_inner_ior, self.vial_ior_inner)),
        })


# List of registered geometries
geometries = {
    'index_matched': IndexMatchedVial,
    'cylindrical': CylindricalVial,
    'square': SquareVial,
    'custom': CustomVial,
    'double_cylindrical': DoubleCylindricalVial
}
