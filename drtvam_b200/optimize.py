"""Pattern optimisation driver and ``drtvam`` CLI.  Mirrors
``src/drtvam/optimize.py``: ``load_scene`` (:15-79), ``optimize`` (:81-368),
``-D key=value`` overrides (:373-394) and ``main`` (:396-441).  The three
projections of every L-BFGS iteration (A p :294, A^T g :308, A d lbfgs.py:248)
run through the CUDA C-ABI.
"""
from __future__ import annotations

import argparse
import json
import os
import time as _time

import numpy as np
import torch

from . import parallel
from .geometry import geometries
from .integrators import CornerIntegrator, RadonIntegrator, VolumeIntegrator
from .lbfgs import LinearLBFGS
from .optimizers import SGD, Adam
from .loss import losses
from .scene import load_dict, render, traverse
from .transform import ScalarTransform4f
from .utils import discretize, load_mesh, save_img, save_vol


def load_scene(config):
    for key in ['target', 'vial', 'projector', 'sensor']:
        if key not in config:
            raise ValueError(f"Missing field '{key}' in the configuration file.")

    if 'type' not in config['vial']:
        raise ValueError("The vial geometry must have a 'type' field.")
    if config['vial']['type'] not in geometries.keys():
        raise ValueError(f"Unknown vial geometry: '{config['vial']['type']}'")
    # the reference resolves mesh paths against the working directory only (optimize.py:15-79); a config
    # loaded through the CLI also finds the vial and occluder meshes next to the config file
    def resolve(path):
        if not os.path.isabs(path) and not os.path.exists(path) and '_config_dir' in config:
            return os.path.join(config['_config_dir'], path)
        return path

    vial_cfg = dict(config['vial'])
    for key in ('filename_vial_outer', 'filename_vial_inner'):
        if key in vial_cfg:
            vial_cfg[key] = resolve(vial_cfg[key])
    if 'occlusions' in vial_cfg:
        vial_cfg['occlusions'] = [dict(o) | ({'filename': resolve(o['filename'])} if 'filename' in o else {})
                                  for o in vial_cfg['occlusions']]
    vial = geometries[config['vial']['type']](vial_cfg)

    if 'filename' not in config['target']:
        raise ValueError("Missing field 'filename' for the target shape.")

    # Scale and centre the target: T(center) S(size / max extent) T(-bbox centre)
    filename = resolve(config['target']['filename'])
    mesh_type = os.path.splitext(filename)[1][1:]
    verts, tris = load_mesh(filename)
    lo, hi = verts.min(0), verts.max(0)
    c = 0.5 * (lo + hi)
    size = config['target'].get('size', 1.)
    center_pos = [config['target'].get(f'box_center_{a}', 0.) for a in 'xyz']
    target_to_world = ScalarTransform4f().translate(center_pos) @ \
        ScalarTransform4f().scale(size / float(np.max(hi - lo))) @ ScalarTransform4f().translate(-c)

    def get_sensor_transform(sensor_dict):
        sx = sensor_dict.pop('scalex', 1.)
        sy = sensor_dict.pop('scaley', 1.)
        sz = sensor_dict.pop('scalez', 1.)
        return ScalarTransform4f().scale((sx, sy, sz))

    sensor_to_world = get_sensor_transform(config['sensor'])
    scene_dict = {
        'type': 'scene',
        'projector': config['projector'],
        'sensor': config['sensor'] | {'to_world': sensor_to_world},
        'target': {
            'type': mesh_type,
            'filename': filename,
            'to_world': target_to_world,
            'bsdf': {'type': 'null'},
        },
    } | vial.to_dict()

    if 'final_sensor' in config.keys():
        final_sensor_to_world = get_sensor_transform(config['final_sensor'])
        scene_dict['final_sensor'] = config['final_sensor'] | {'to_world': final_sensor_to_world}
    return scene_dict


def _dist_state():
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def optimize(config, patterns_fwd=None, progress=True):
    """Optimize the patterns for TVAM; returns the final rendered dose.
    If `patterns_fwd` is given the optimisation is skipped (forward mode)."""
    scene_dict = load_scene(config)
    scene = load_dict(scene_dict)
    params = traverse(scene)
    rank, world = _dist_state()

    output = config['output']
    spp = config.get('spp', 4)
    spp_ref = config.get('spp_ref', 16)
    spp_grad = config.get('spp_grad', spp)
    max_depth = config.get('max_depth', 6)
    rr_depth = config.get('rr_depth', 6)  # i.e. disabled by default
    time = config.get('time', 1.)  # print duration in seconds
    progressive = config.get('progressive', False)
    transmission_only = config.get('transmission_only', True)
    regular_sampling = config.get('regular_sampling', False)

    sensor = scene.sensor('sensor')
    final_sensor = scene._sensors.get('final_sensor', sensor)
    if final_sensor.film().surface_aware:
        raise ValueError("The final sensor is used to generate visualizations and metrics of the final simulated print. Therefore, it must not be surface-aware. If you are using the surface-aware discretization for optimization, please specify another sensor called 'final_sensor' in the configuration file.")

    surface_aware = sensor.film().surface_aware

    def keep_positive(image):
        # optimize.py:151-159 / 173-181: the pixels whose filter image is positive, patterns reset to zero
        active_pixels = torch.nonzero(image.reshape(-1) > 0.).reshape(-1).to(torch.int32)
        if active_pixels.numel() == 0:
            raise ValueError("No active pixels found in the Radon transform.")
        params['projector.active_pixels'] = active_pixels
        params['projector.active_data'] = torch.zeros(active_pixels.numel(), device=image.device,
                                                      dtype=torch.float32)
        params.update()

    if config.get('filter_radon', False) and patterns_fwd is None:
        # Deactivate pixels where the Radon transform is zero (optimize.py:143-163)
        radon_integrator = RadonIntegrator({'max_depth': 5})
        keep_positive(render(scene, integrator=radon_integrator, spp=config.get('spp_filter_radon', 4)))
    if 'filter_corner' in config and patterns_fwd is None:
        corner_integrator = CornerIntegrator({'regular_sampling': True} | config['filter_corner'])
        keep_positive(render(scene, integrator=corner_integrator, spp=1))

    integrator = VolumeIntegrator({
        'max_depth': 3 if progressive else max_depth,
        'rr_depth': rr_depth,
        'print_time': time,
        'transmission_only': transmission_only,
        'regular_sampling': regular_sampling,
    })

    if surface_aware:
        target = sensor.compute_volume(scene)
    else:
        target = discretize(scene, sensor=sensor)
    if rank == 0:
        os.makedirs(output, exist_ok=True)
        if surface_aware:   # optimize.py:131-134
            save_vol(target[..., 0, None], os.path.join(output, "target_in.exr"))
            save_vol(target[..., 1, None], os.path.join(output, "target_out.exr"))
        else:               # optimize.py:135-137
            save_vol(target, os.path.join(output, "target.exr"))
        np.save(os.path.join(output, "target.npy"), target.cpu().numpy())

    patterns_key = 'projector.active_data'

    if "loss" not in config.keys():
        print("No loss function specified. Using thresholded loss.")
        config['loss'] = {'type': 'threshold'}
    loss_cfg = dict(config['loss'])
    loss_type = loss_cfg.pop('type')
    if loss_type not in losses.keys():
        raise ValueError(f"Unknown loss type: '{loss_type}'. Available losses are: {list(losses.keys())}")
    loss_fn = losses[loss_type](loss_cfg)

    if 'optimizer' not in config.keys():
        print("No optimizer specified. Using linear L-BFGS.")
        config['optimizer'] = {'type': 'lbfgs'}
    opt_cfg = dict(config['optimizer'])
    optim_type = opt_cfg.pop('type')
    if optim_type not in ('lbfgs', 'adam', 'sgd'):
        raise ValueError(f"Unknown optimizer type: '{optim_type}'")

    if world > 1:
        scene.shard(rank, world)
        params = traverse(scene)
        loss_fn.pattern_allsum = parallel.make_sum(None, world)

    state = {'i': 0}

    def render_fn(vars):
        scene.projector.active_data = vars[patterns_key].detach()
        # same seed as the primal pass of this iteration: vol(p + a d) = vol + a dvol
        return render(scene, None, integrator=integrator, sensor=sensor, spp=spp, spp_grad=spp_grad, seed=state['i'])

    def loss_fn2(y, patterns):
        return loss_fn(y, target, patterns)
    loss_fn2.accepts_lazy = getattr(loss_fn, 'accepts_lazy', False)
    if getattr(loss_fn, 'armijo_sweep', None) is not None:
        loss_fn2.armijo_sweep = lambda x, dx, direction, *a: loss_fn.armijo_sweep(x, dx, target, direction, *a)

    consensus = parallel.make_consensus(None, world) if world > 1 else None
    if optim_type == 'adam':      # optimize.py:205-208
        if world > 1 and opt_cfg.get('uniform', False):
            raise NotImplementedError("Adam(uniform=True) on a sharded pattern vector")
        opt = Adam(**opt_cfg)
    elif optim_type == 'sgd':
        opt = SGD(**opt_cfg)
    else:
        opt = LinearLBFGS(loss_fn=loss_fn2, render_fn=render_fn, dot=parallel.make_dot(None, world) if world > 1 else None,
                          allsum=parallel.make_sum(None, world) if world > 1 else None, consensus=consensus, **opt_cfg)
    opt[patterns_key] = params[patterns_key]
    n_steps = config.get('n_steps', 40)
    loss_hist = np.zeros(n_steps)
    timing_hist = np.zeros((n_steps, 2))

    integrator_final = VolumeIntegrator({
        'max_depth': config.get('max_depth_ref', 16),
        'rr_depth': config.get('rr_depth_ref', 8),
        'transmission_only': transmission_only,
        'regular_sampling': regular_sampling,
        'print_time': time,
    })

    if patterns_fwd is not None:
        print("Using provided patterns for forward mode.")
        flat = torch.as_tensor(np.asarray(patterns_fwd, np.float32)).reshape(-1).to(scene.projector.active_data.device)
        if world > 1:
            raise NotImplementedError("forward mode runs on one GPU")
        params[patterns_key] = flat
        params.update()
    elif "psf_analysis" in config:
        # optimize.py:245-264: project a handful of single pixels (a sparse pixel list: the general tracer)
        if world > 1:
            raise NotImplementedError("psf_analysis runs on one GPU")
        print("\nPSF analysis enabled.")
        entries = config["psf_analysis"]
        print("Number of traced pixels:", len(entries))
        xres, yres = config["projector"]["resx"], config["projector"]["resy"]
        pixels, data = [], []
        for entry in entries:
            assert entry["x"] < xres, "Invalid entry in psf_analysis: x out of bounds. Please check the configuration file."
            assert entry["y"] < yres, "Invalid entry in psf_analysis: y out of bounds. Please check the configuration file."
            assert entry["index_pattern"] < config["projector"]["n_patterns"], \
                "Invalid entry in psf_analysis: index_pattern out of bounds. Please check the configuration file."
            pixels.append(xres * yres * entry["index_pattern"] + xres * entry["y"] + entry["x"])
            data.append(float(entry["intensity"]))
        dev = scene.projector.active_data.device
        params['projector.active_pixels'] = torch.tensor(pixels, dtype=torch.int64, device=dev).to(torch.int32)
        params['projector.active_data'] = torch.tensor(data, dtype=torch.float32, device=dev)
        params.update()
    else:
        if rank == 0:
            print("Optimizing patterns...")
        it = range(n_steps)
        if progress and rank == 0:
            try:
                from tqdm import trange
                it = trange(n_steps)
            except Exception:
                pass
        dev = scene.projector.active_data.device
        for i in it:
            state['i'] = i
            if progressive and i == 5:
                integrator.max_depth = max_depth
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
            ev[0].record()
            p = opt[patterns_key]
            scene.projector.active_data = p
            vol = _RenderWithGrad(scene, integrator, sensor, spp, spp_grad, i, p)
            loss = loss_fn(vol, target, p)
            # every rank holds the same all-reduced dose, but its loss sum uses atomics: the value that decides
            # to stop is the rank mean, identical everywhere
            loss_hist[i] = float(loss.detach() if consensus is None else consensus(loss.detach()))
            ev[1].record()
            loss.backward()
            if loss_hist[i] == 0.0:
                print("Converged")
                break
            if optim_type == 'lbfgs':
                opt.step(vol.detach(), loss.detach())
            else:
                opt.step()
            # clamp patterns (optimize.py:320)
            opt[patterns_key] = torch.clamp_min(opt[patterns_key].detach(), 0)
            ev[2].record()
            torch.cuda.synchronize(dev)
            timing_hist[i, 0] = ev[0].elapsed_time(ev[1]) * 1e-3
            timing_hist[i, 1] = ev[0].elapsed_time(ev[2]) * 1e-3
        scene.projector.active_data = opt[patterns_key].detach()

    if rank == 0:
        print("Rendering final state...")
    with torch.no_grad():
        vol_final = render(scene, None, spp=spp_ref, integrator=integrator_final, sensor=final_sensor)

    if rank == 0:
        np.save(os.path.join(output, "final.npy"), vol_final.cpu().numpy())
        save_vol(vol_final, os.path.join(output, "final.exr"))
        np.save(os.path.join(output, "loss.npy"), loss_hist)
        np.save(os.path.join(output, "timing.npy"), timing_hist)
    # the ranks' pattern shards -> the one pattern stack the reference saves (optimize.py:331-358)
    imgs_final = scene.full_patterns()
    if rank == 0:
        array = imgs_final.cpu().numpy()
        if config.get('save_pattern_images', True):
            print("Saving images...")
            for i in range(array.shape[0]):
                save_img(array[i], os.path.join(output, "patterns", f"{i:04d}.exr"))
        np.savez_compressed(os.path.join(output, "patterns.npz"), patterns=array)
        array_max = np.max(array)
        if array_max > 0:
            normalized_array = array / array_max
            np.savez_compressed(os.path.join(output, "patterns_normalized_uint8.npz"),
                                patterns=(normalized_array * 255).astype(np.uint8))
            efficiency = np.sum(normalized_array / normalized_array.size)
            print("Pattern efficiency {:.4f}".format(efficiency))
        if surface_aware:
            # a binary target on the final sensor's grid next to the fractional one (optimize.py:354-359)
            target_binary = discretize(scene, sensor=final_sensor)
            np.save(os.path.join(output, "target_binary.npy"), target_binary.cpu().numpy())
            save_vol(target_binary, os.path.join(output, "target_binary.exr"))
    return vol_final


def _RenderWithGrad(scene, integrator, sensor, spp, spp_grad, seed, p):
    from .scene import Params
    prm = Params(scene)
    dict.__setitem__(prm, 'projector.active_data', p)
    return render(scene, prm, integrator=integrator, sensor=sensor, spp=spp, spp_grad=spp_grad, seed=seed)


class OverrideAction(argparse.Action):
    def __init__(self, option_strings, dest, nargs=None, **kwargs):
        super().__init__(option_strings, dest, **kwargs)
        self.overrides = {}

    def __call__(self, parser, namespace, values, option_string=None):
        try:
            key, value = values.split('=')
        except ValueError:
            raise ValueError("Invalid parameter override. Use the format '-D key=value'")
        for cast in (int, float):
            try:
                value = cast(value)
                break
            except ValueError:
                continue
        self.overrides[key] = value
        setattr(namespace, self.dest, self.overrides)


def main(argv=None):
    parser = argparse.ArgumentParser("Optimize patterns for TVAM.")
    parser.add_argument("config", type=str, help="Path to the configuration file")
    parser.add_argument("-D", dest="overrides", metavar="key=value", action=OverrideAction,
                        help="Override/Add a parameter in the configuration dictionary. Nested keys are separated by dots.")
    parser.add_argument("--backend", type=str, default="cuda", choices=["cuda"],
                        help="Only the CUDA (sm_100a) backend exists in this build.")
    parser.add_argument("--forward_mode", action="store_true",
                        help="Just project the patterns without optimization. Needs --patterns (a .npz file).")
    parser.add_argument("--patterns", type=str, help="Path to the patterns file (.npz), forward mode only.")
    args = parser.parse_args(argv)

    with open(args.config, 'r') as f:
        config = json.load(f)
    if args.overrides is not None:
        for key, value in args.overrides.items():
            key = key.split('.')
            tmp = config
            for k in key[:-1]:
                tmp = tmp[k]
            tmp[key[-1]] = value

    config['_config_dir'] = os.path.dirname(os.path.abspath(args.config))
    if 'output' not in config:
        config['output'] = config['_config_dir']
    os.makedirs(os.path.join(config['output'], "patterns"), exist_ok=True)
    with open(os.path.join(config['output'], "opt_config.json"), 'w') as f:
        json.dump(config, f, indent=4)

    if args.forward_mode:
        if args.patterns is None:
            raise ValueError("In forward mode, you must specify the patterns file.")
        patterns = np.load(args.patterns)['patterns']
        optimize(config, patterns_fwd=patterns)
    else:
        optimize(config)


if __name__ == "__main__":
    main()
