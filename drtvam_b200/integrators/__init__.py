from .common import TVAMIntegrator
from .volume import VolumeIntegrator
from .radon import RadonIntegrator
from .filter_corner import CornerIntegrator

integrators = {
    'volume': VolumeIntegrator,
    'radon': RadonIntegrator,
    'corner': CornerIntegrator,
}
