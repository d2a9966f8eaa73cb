"""CPU suite: golden vectors produced by IMPORTING THE PYTHON REFERENCE (scripts/make_golden.py, run in the build
container where /root/reference exists) against (a) the host mirror's pure-Python logic and (b) the CPU oracle / port
that bench.py times as the reference's CPU path.  This is the pin of the oracle for every piece of the path the
reference can execute without its CUDA extensions."""
import os

import numpy as np
import pytest
import torch

from util import rel_err

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def gold(name):
    return np.load(os.path.join(GOLD, name), allow_pickle=False)


def test_level_table_golden(oracle):
    """GridEncoder.__init__ (grid.py:94-131): offsets, per_level_scale, parameter count — host mirror and oracle."""
    from nerf_navigation_b200.gridencoder import GridEncoder
    g = gold("level_table.npz")
    for bound in (1, 2):
        enc = GridEncoder(input_dim=3, num_levels=16, level_dim=2, base_resolution=16, log2_hashmap_size=19,
                          desired_resolution=2048 * bound)
        assert np.array_equal(enc.offsets.numpy(), g[f"offsets_b{bound}"])
        assert enc.offsets.dtype == torch.int32
        assert float(enc.per_level_scale) == float(g[f"per_level_scale_b{bound}"])
        assert enc.embeddings.numel() == int(g[f"n_params_b{bound}"]) and enc.output_dim == int(g[f"output_dim_b{bound}"])
        assert float(enc.embeddings.detach().abs().max()) <= 1e-4 and abs(float(g[f"emb_absmax_b{bound}"]) - 1e-4) < 1e-6
        off, pls = oracle.level_table(3, 16, 2, 2.0, 16, 19, 2048 * bound)
        assert np.array_equal(off, g[f"offsets_b{bound}"]) and pls == float(g[f"per_level_scale_b{bound}"])
    enc = GridEncoder(input_dim=2, num_levels=8, level_dim=4, base_resolution=8, log2_hashmap_size=12, desired_resolution=512)
    assert np.array_equal(enc.offsets.numpy(), g["offsets_2d"]) and float(enc.per_level_scale) == float(g["per_level_scale_2d"])


def test_ffmlp_init_golden():
    """FFMLP.reset_parameters (ffmlp.py:141-144): seed 42, U(+-sqrt(3/64)), flat fp16 weights of the reference sizes."""
    from nerf_navigation_b200.ffmlp import FFMLP
    g = gold("ffmlp_init.npz")
    for tag, (i, o, n) in {"sigma": (32, 16, 2), "color": (32, 3, 3)}.items():
        m = FFMLP(input_dim=i, output_dim=o, hidden_dim=64, num_layers=n)
        w = m.weights.detach().float().numpy()
        assert w.size == int(g[f"{tag}_numel"]) and str(m.weights.dtype) == str(g[f"{tag}_dtype"])
        assert np.array_equal(w[:256], g[f"{tag}_head"]) and np.array_equal(w[-256:], g[f"{tag}_tail"])
        assert w.astype(np.float64).sum() == float(g[f"{tag}_sum"])


def test_trunc_exp_golden():
    from nerf_navigation_b200.activation import trunc_exp
    g = gold("trunc_exp.npz")
    x = torch.from_numpy(g["x"]).requires_grad_(True)
    y = trunc_exp(x)
    y.backward(torch.from_numpy(g["g"]))
    assert np.array_equal(y.detach().numpy(), g["y"]) and np.array_equal(x.grad.numpy(), g["gx"])


def test_get_rays_golden():
    """closures.rays_from_poses (sampled pixels only) == nerf/utils.py:52-116 get_rays at the same indices."""
    from nerf_navigation_b200.closures import rays_from_poses
    g = gold("get_rays.npz")
    ro, rd = rays_from_poses(torch.from_numpy(g["poses"]), torch.from_numpy(g["inds"][0]), tuple(g["intrinsics"]), 800)
    assert rel_err(ro.numpy(), g["rays_o"]) == 0.0
    assert rel_err(rd.numpy(), g["rays_d"]) < 1e-6


def test_sample_pdf_golden():
    from nerf_navigation_b200.nerf.renderer import sample_pdf
    g = gold("sample_pdf.npz")
    s = sample_pdf(torch.from_numpy(g["bins"]), torch.from_numpy(g["weights"]), 48, det=True)
    assert rel_err(s.numpy(), g["samples"]) < 1e-6


def _cfg1_field():
    from oracle import cfg1_torch as C1
    g = gold("cfg1.npz")
    sd = {k[3:].replace("__", "."): g[k] for k in g.files if k.startswith("sd_")}
    return C1, g, C1.Cfg1Field(sd, bound=2.0, density_scale=1.0, min_near=0.2)


def test_cfg1_density_port_vs_reference(oracle):
    """The CPU port bench.py times (oracle/cfg1_torch.py) against the reference's own NeRFNetwork.density + backward
    (config 1a: 20 x 100 planner points).  Same torch ops on the same machine class: 1e-6."""
    C1, g, field = _cfg1_field()
    x = torch.from_numpy(g["x"]).requires_grad_(True)
    sigma, _ = field.density(x.reshape(-1, 3))
    (sigma ** 2).mean().backward()
    assert rel_err(sigma.detach().numpy(), g["sigma"]) < 1e-6
    assert rel_err(x.grad.numpy(), g["dx"]) < 1e-5


def test_cfg1_render_port_vs_reference(oracle):
    """Config 1b: 1024 rays x 512 samples through the uniform sampler (renderer.py:125-254), image/depth and the
    gradient of the photometric loss w.r.t. rays_o / rays_d (fp32 tolerance 1e-5; gradients 1e-4: cumprod chains)."""
    C1, g, field = _cfg1_field()
    ro = torch.from_numpy(g["rays_o"][0]).requires_grad_(True)
    rd = torch.from_numpy(g["rays_d"][0]).requires_grad_(True)
    image, depth = C1.render_run(field, ro, rd, num_steps=512, bg_color=1.0)
    loss = ((image - torch.from_numpy(g["target"][0])) ** 2).mean()
    loss.backward()
    assert rel_err(image.detach().numpy(), g["image"][0]) < 1e-5
    assert rel_err(depth.detach().numpy(), g["depth"][0]) < 1e-5
    assert abs(loss.item() - float(g["loss"])) < 1e-6 * float(g["loss"])
    assert rel_err(ro.grad.numpy(), g["d_rays_o"][0]) < 1e-4 and rel_err(rd.grad.numpy(), g["d_rays_d"][0]) < 1e-4


def test_hashgrid_torch_transcription_vs_oracle(oracle):
    """The torch transcription of the hash-grid lookup (north_star's second CPU baseline) == the C oracle."""
    from oracle import cfg1_torch as C1
    from util import hashgrid_config
    cfg = hashgrid_config(table_std=0.5)
    sc = oracle.level_scales(16, cfg["per_level_scale"], 16)
    rng = np.random.default_rng(3)
    x = rng.uniform(0, 1, (500, 3)).astype(np.float32)
    ref, _, _ = oracle.grid_encode_forward(x, cfg["table"], cfg["offsets"], sc)       # [L, B, C]
    out = C1.hashgrid_encode(torch.from_numpy(x), torch.from_numpy(cfg["table"]), cfg["offsets"], sc)
    assert rel_err(out.numpy().reshape(500, 16, 2).transpose(1, 0, 2), ref) < 1e-5


def test_estimator_pose_algebra_golden():
    """closures.state_to_pose (12-state -> NGP cam2world pose) and its gradient w.r.t. the state against the reference's own
    vec_to_rot_matrix / rot_x / nerf_matrix_to_ngp_torch chain (nav/estimator_helpers.py:293-327, nav/math_utils.py:26-38,
    159-174), incl. the zero rotation vector."""
    from nerf_navigation_b200 import closures
    g = np.load(os.path.join(GOLD, "pose_algebra.npz"))
    states = torch.from_numpy(g["states"]).requires_grad_(True)
    poses = closures.state_to_pose(states)
    assert poses.shape == (8, 4, 4)
    assert np.abs(poses.detach().numpy() - g["poses"]).max() < 1e-6
    (poses * torch.from_numpy(g["functional"])).sum().backward()
    assert np.abs(states.grad.numpy() - g["grad_states"]).max() < 1e-5 * max(1.0, np.abs(g["grad_states"]).max())
    assert (states.grad[:, 3:6] == 0).all() and (states.grad[:, 9:] == 0).all()     # velocities take no measurement gradient
