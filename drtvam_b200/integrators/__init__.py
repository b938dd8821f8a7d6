from .common import TVAMIntegrator
from .volume import VolumeIntegrator

integrators = {
    'volume': VolumeIntegrator,
}
