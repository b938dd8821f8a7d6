"""Integrator base: option parsing and wavefront bookkeeping.  Mirrors
``src/drtvam/integrators/common.py`` (``__init__`` :6-22, ``prepare`` :41-68;
``sample_rays`` :70-116 runs on the GPU in ``csrc/raygen.cuh``).
"""
from __future__ import annotations


class TVAMIntegrator:
    def __init__(self, props):
        # Mitsuba's ADIntegrator reads these two (defaults 6 / 5; -1 = unbounded)
        max_depth = props.get('max_depth', 6)
        self.max_depth = max_depth if max_depth >= 0 else 2 ** 31 - 1
        self.rr_depth = props.get('rr_depth', 5)
        # Sample time stochastically ?
        self.sample_time = props.get('sample_time', False)
        self.print_time = props.get('print_time', 1.)
        # ID of the target shape, necessary when surface-aware discretization is enabled
        self.target_id = props.get('target_id', 'target')
        # Always sample only the transmission component at interfaces
        self.transmission_only = props.get('transmission_only', True)
        # Shoot from the center of the pixels only
        self.regular_sampling = props.get('regular_sampling', False)

    def prepare(self, emitter, seed: int = 0, spp: int = 0):
        """Returns the effective spp and wavefront size (common.py:41-68)."""
        if self.regular_sampling:
            spp = 1  # spp > 1 makes no sense when shooting from the pixel centre
        if spp == 0:
            spp = 4  # Mitsuba's independent sampler default sample_count
        wavefront_size = emitter.active_size() * spp
        if wavefront_size > 2 ** 32:
            raise Exception(
                "The total number of Monte Carlo samples required by this "
                "rendering task (%i) exceeds 2^32 = 4294967296. Please use "
                "fewer samples per pixel or render using multiple passes."
                % wavefront_size)
        return spp, wavefront_size
