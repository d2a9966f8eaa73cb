"""Generate tests/golden/*.npz by IMPORTING THE PYTHON REFERENCE from /root/reference (runs only in the build container;
the reference cannot travel to the GPU box, the vectors can).  Nothing here copies reference source: the reference's
own modules are executed, with
  * MagicMock stand-ins for modules that are absent and irrelevant to the path (imageio, tensorboardX, matplotlib, cv2,
    trimesh, mcubes, torch_ema, lpips, dearpygui, tqdm, rich ...),
  * MagicMock `_backend`s for the CUDA extensions (gridencoder/ffmlp/... import them at module import time; none of the
    pure-Python logic captured below calls into them),
  * a stand-in `raymarching` module whose near_far_from_aabb is the CPU oracle (raymarching.cu:93-147) — the only
    raymarching op NeRFRenderer.run needs (renderer.py:141-142),
  * get_encoder('frequency') routed to the reference's own pure-torch encoding.FreqEncoder (encoding.py:5-43), which its
    factory leaves commented out in favour of the CUDA freqencoder (encoding.py:55).

Vectors written (what pins what):
  level_table.npz   GridEncoder.__init__ offsets / per_level_scale / param count (grid.py:94-131) for bound 1 and 2
  ffmlp_init.npz    FFMLP.reset_parameters (ffmlp.py:141-144), sigma and colour nets
  trunc_exp.npz     activation.py:5-17 forward / backward
  get_rays.npz      nerf/utils.py:52-116 for two poses, sampled pixel indices
  sample_pdf.npz    nerf/renderer.py:13-48 (det=True)
  cfg1.npz          BASELINE config 1: NeRFNetwork(frequency) state dict, density() on 2000 points with dL/dx, and
                    render(staged, 512 steps) of 1024 rays with the image, depth and dL/d(rays_o, rays_d)

  python scripts/make_golden.py
"""
import os
import sys
import types
from unittest.mock import MagicMock

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
OUT = os.path.join(ROOT, "tests", "golden")
sys.path.insert(0, ROOT)


def install_stubs():
    for name in ["imageio", "tensorboardX", "matplotlib", "matplotlib.pyplot", "cv2", "trimesh", "mcubes", "torch_ema",
                 "lpips", "dearpygui", "dearpygui.dearpygui", "tqdm", "rich", "rich.console", "turtle"]:
        sys.modules[name] = MagicMock()
    for ext in ["_gridencoder", "_ffmlp", "_shencoder", "_freqencoder", "_raymarching"]:
        sys.modules[ext] = MagicMock()
    from oracle import oracle as O
    rm = types.ModuleType("raymarching")

    def near_far_from_aabb(rays_o, rays_d, aabb, min_near=0.2):
        n, f = O.near_far_from_aabb(rays_o.detach().numpy().reshape(-1, 3), rays_d.detach().numpy().reshape(-1, 3),
                                    aabb.numpy(), min_near)
        return torch.from_numpy(n), torch.from_numpy(f)

    rm.near_far_from_aabb = near_far_from_aabb
    sys.modules["raymarching"] = rm
    sys.path.insert(0, REF)


def pose_algebra():
    """nav/estimator_helpers.py:293-327 measurement_fn, the pose part: state[6:9] -> vec_to_rot_matrix (nav/math_utils.py:159-174)
    -> rot_x(pi/2) @ R -> nerf_matrix_to_ngp_torch (math_utils.py:26-38) -> 4x4 cam2world handed to get_rays; plus the
    gradient of a fixed linear functional of the pose w.r.t. the 12-state (autograd through the reference's own code)."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("ref_math_utils", os.path.join(REF, "nav", "math_utils.py"))
    M = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(M)
    gen = torch.Generator().manual_seed(11)
    states = torch.randn(8, 12, generator=gen)
    states[:, 6:9] *= 0.7
    states[0, 6:9] = 0.0                        # zero rotation vector: the 1e-10 guard of vec_to_rot_matrix
    states[1, 6:9] = torch.tensor([1e-6, 0.0, 0.0])
    Wf = torch.randn(4, 4, generator=gen)
    poses, grads = [], []
    for st in states:
        st = st.clone().requires_grad_(True)
        R = M.vec_to_rot_matrix(st[6:9])
        rot = M.rot_x(torch.tensor(np.pi / 2)) @ R[:3, :3]
        pose, trans = M.nerf_matrix_to_ngp_torch(rot, st[:3])
        new_pose = torch.eye(4)
        new_pose[:3, :3] = pose
        new_pose[:3, 3] = trans
        (new_pose * Wf).sum().backward()
        poses.append(new_pose.detach().numpy())
        grads.append(st.grad.numpy())
    np.savez(os.path.join(OUT, "pose_algebra.npz"), states=states.numpy(), poses=np.stack(poses), functional=Wf.numpy(),
             grad_states=np.stack(grads))


def main():
    os.makedirs(OUT, exist_ok=True)
    if "--pose-only" in sys.argv:
        install_stubs()
        pose_algebra()
        return
    install_stubs()
    torch.set_num_threads(max(1, os.cpu_count() or 1))

    # ---------------------------------------------------------------- G1: level table
    from gridencoder.grid import GridEncoder
    rec = {}
    for bound in (1, 2):
        enc = GridEncoder(input_dim=3, num_levels=16, level_dim=2, base_resolution=16, log2_hashmap_size=19,
                          desired_resolution=2048 * bound)
        rec[f"offsets_b{bound}"] = enc.offsets.numpy().astype(np.int32)
        rec[f"per_level_scale_b{bound}"] = np.float64(enc.per_level_scale)
        rec[f"n_params_b{bound}"] = np.int64(enc.embeddings.numel())
        rec[f"emb_absmax_b{bound}"] = np.float32(enc.embeddings.abs().max().item())
        rec[f"output_dim_b{bound}"] = np.int64(enc.output_dim)
    enc = GridEncoder(input_dim=2, num_levels=8, level_dim=4, base_resolution=8, log2_hashmap_size=12, desired_resolution=512)
    rec["offsets_2d"] = enc.offsets.numpy().astype(np.int32)
    rec["per_level_scale_2d"] = np.float64(enc.per_level_scale)
    np.savez(os.path.join(OUT, "level_table.npz"), **rec)

    # ---------------------------------------------------------------- M3: FFMLP init
    from ffmlp.ffmlp import FFMLP
    rec = {}
    for tag, (i, o, n) in {"sigma": (32, 16, 2), "color": (32, 3, 3)}.items():
        m = FFMLP(input_dim=i, output_dim=o, hidden_dim=64, num_layers=n)
        w = m.weights.detach().float().numpy()
        rec[f"{tag}_numel"] = np.int64(w.size)
        rec[f"{tag}_head"] = w[:256].copy()
        rec[f"{tag}_tail"] = w[-256:].copy()
        rec[f"{tag}_sum"] = np.float64(w.astype(np.float64).sum())
        rec[f"{tag}_dtype"] = str(m.weights.dtype)
    np.savez(os.path.join(OUT, "ffmlp_init.npz"), **rec)

    # ---------------------------------------------------------------- M4: trunc_exp
    from activation import trunc_exp
    x = torch.linspace(-20, 20, 81, requires_grad=True)
    y = trunc_exp(x)
    g = torch.linspace(0.5, 1.5, 81)
    y.backward(g)
    np.savez(os.path.join(OUT, "trunc_exp.npz"), x=x.detach().numpy(), y=y.detach().numpy(), g=g.numpy(), gx=x.grad.numpy())

    # ---------------------------------------------------------------- get_rays / sample_pdf
    import nerf.utils as U
    import nerf.renderer as R
    gen = torch.Generator().manual_seed(3)
    poses = torch.eye(4).repeat(2, 1, 1)
    q, _ = torch.linalg.qr(torch.randn(2, 3, 3, generator=gen))
    poses[:, :3, :3] = q
    poses[:, :3, 3] = torch.randn(2, 3, generator=gen)
    intr = np.array([1111.1, 1111.1, 400.0, 400.0])
    torch.manual_seed(5)
    res = U.get_rays(poses, intr, 800, 800, N=257)
    np.savez(os.path.join(OUT, "get_rays.npz"), poses=poses.numpy(), intrinsics=intr, inds=res["inds"].numpy(),
             rays_o=res["rays_o"].numpy(), rays_d=res["rays_d"].numpy())
    bins = torch.sort(torch.rand(7, 33, generator=gen), dim=-1).values
    wts = torch.rand(7, 32, generator=gen)
    smp = R.sample_pdf(bins, wts, 48, det=True)
    np.savez(os.path.join(OUT, "sample_pdf.npz"), bins=bins.numpy(), weights=wts.numpy(), samples=smp.numpy())

    # ---------------------------------------------------------------- estimator pose algebra (12-state -> NGP pose)
    pose_algebra()

    # ---------------------------------------------------------------- config 1
    import encoding as E
    orig_get_encoder = E.get_encoder

    def get_encoder(enc_name, input_dim=3, multires=6, **kw):
        if enc_name == "frequency":
            e = E.FreqEncoder(input_dim=input_dim, max_freq_log2=multires - 1, N_freqs=multires, log_sampling=True)
            return e, e.output_dim
        return orig_get_encoder(enc_name, input_dim=input_dim, multires=multires, **kw)

    E.get_encoder = get_encoder
    import nerf.network as N
    N.get_encoder = get_encoder
    torch.manual_seed(0)
    model = N.NeRFNetwork(encoding="frequency", encoding_dir="frequency", bound=2, cuda_ray=False, density_scale=1,
                          min_near=0.2, density_thresh=10, bg_radius=-1)
    model.eval()
    rec = {"sd_" + k.replace(".", "__"): v.numpy() for k, v in model.state_dict().items()}
    gen = torch.Generator().manual_seed(1)
    x = (torch.rand(20, 100, 3, generator=gen) * 2 - 1).requires_grad_(True)
    sigma = model.density(x.reshape(-1, 3))["sigma"]
    (sigma ** 2).mean().backward()
    rec.update(x=x.detach().numpy(), sigma=sigma.detach().numpy(), dx=x.grad.numpy())
    # estimator: 1024 rays of an 800x800 pinhole at (0, 0, -1.5) looking +z (SURVEY §8d cfg 1)
    pose = torch.eye(4).unsqueeze(0)
    pose[0, 2, 3] = -1.5
    torch.manual_seed(7)
    rays = U.get_rays(pose, intr, 800, 800, N=1024)
    ro = rays["rays_o"].clone().requires_grad_(True)
    rd = rays["rays_d"].clone().requires_grad_(True)
    target = torch.rand(1, 1024, 3, generator=gen)
    out = model.render(ro, rd, staged=True, bg_color=1.0, perturb=False, num_steps=512, upsample_steps=0, max_ray_batch=4096)
    loss = ((out["image"] - target) ** 2).mean()
    loss.backward()
    rec.update(rays_o=ro.detach().numpy(), rays_d=rd.detach().numpy(), inds=rays["inds"].numpy(), target=target.numpy(),
               image=out["image"].detach().numpy(), depth=out["depth"].detach().numpy(), loss=np.float64(loss.item()),
               d_rays_o=ro.grad.numpy(), d_rays_d=rd.grad.numpy())
    # a second, short render with upsampling (sample_pdf, det=True in eval) for the host-logic parity of run()
    with torch.no_grad():
        out2 = model.render(ro[:, :64], rd[:, :64], staged=True, bg_color=1.0, perturb=False, num_steps=64,
                            upsample_steps=32, max_ray_batch=4096)
    rec.update(image_up=out2["image"].numpy(), depth_up=out2["depth"].numpy())
    np.savez_compressed(os.path.join(OUT, "cfg1.npz"), **rec)
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))


if __name__ == "__main__":
    main()
