"""Extra legs of bench.py (rank 0, N = 1): (a) the reference's OWN CUDA extensions (oracle/_ref, unmodified sources built for
sm_100a) timed with CUDA events on configs 2-4 beside ours, incl. the like-for-like pose-gradient comparator of BASELINE.md
§3; (b) the FFMLP tcgen05 kernels against the reference's wmma build and against the HBM roofline."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def _time(fn, reps, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def ref_cuda_build_leg(dev, points=1 << 22, train_rays=32768, run_rays=1024, frame_rays=640000):
    """Returns a dict of reference-CUDA-build timings, each with OUR number for the same work beside it."""
    from oracle import ref_pipeline as RP
    if not RP.available():
        return {"unavailable": "oracle/_ref extensions are not in the tree (oracle/build_ref.py builds them where /root/reference exists)"}
    from nerf_navigation_b200 import closures, synthetic
    from nerf_navigation_b200 import raymarching as rm
    from nerf_navigation_b200.field import render_rays_infer
    from nerf_navigation_b200.nerf.network import NeRFNetwork
    torch.manual_seed(0)
    ours = NeRFNetwork(encoding="hashgrid", bound=2, cuda_ray=True, density_scale=1, min_near=0.2, density_thresh=10).to(dev).eval()
    ours.encoder.embeddings.data.uniform_(-0.5, 0.5)
    synthetic.install_stonehenge(ours)
    emb = ours.encoder.embeddings.detach().clone().requires_grad_(True)
    sw = [l.weight.detach().clone().requires_grad_(True) for l in ours.sigma_net]
    cw = [l.weight.detach().clone().requires_grad_(True) for l in ours.color_net]
    field = RP.RefField(emb, ours.encoder.offsets, ours.encoder.per_level_scale, sw, cw, 2.0)
    out = {"what": "the reference's own gridencoder / raymarching / shencoder CUDA sources, unmodified, built for sm_100a "
                   "(oracle/build_ref.py) and driven in the call order of its Python wrappers (oracle/ref_pipeline.py); same box, "
                   "same run, CUDA events"}

    def zero():
        for p in field.params():
            p.grad = None

    # ---- config 2: density fwd + dL/dx bwd (the reference also scatters the table gradient: grid.py:74-81)
    x = torch.rand(points, 3, device=dev) * 2 - 1

    def ref2():
        xin = x.detach().requires_grad_(True)
        sigma, _ = field.density(xin)
        (sigma ** 2).mean().backward()
        zero()

    def our2():
        xin = x.detach().requires_grad_(True)
        (ours.density(xin)["sigma"] ** 2).mean().backward()

    r, o = _time(ref2, 3), _time(our2, 5)
    out["cfg2_field_fwd_bwd"] = {"points": points, "ref_ms": r, "ref_Msamples_per_s": points / r / 1e3, "ours_ms": o,
                                 "ours_Msamples_per_s": points / o / 1e3, "ours_vs_ref": r / o,
                                 "ours": "NeRFNetwork.density(x) + autograd (the drop-in call)"}
    del x

    # ---- config 3: photometric loss through the cuda_ray train branch
    gt = closures.look_at_pose((1.2, 0.0, 0.4))
    hyp = max(1, train_rays // 4096)
    poses = closures.perturb_poses(gt, hyp, generator=torch.Generator().manual_seed(0)).to(dev)
    pix = torch.randint(0, 640000, (4096,), generator=torch.Generator().manual_seed(0)).to(dev)
    ro, rd = closures.rays_from_poses(poses, pix, (1111.1, 1111.1, 400.0, 400.0), 800)
    ro, rd = ro.reshape(-1, 3).contiguous(), rd.reshape(-1, 3).contiguous()
    N = ro.shape[0]
    target = torch.rand(N, 3, device=dev)
    info = {}

    def ref3(backward=True):
        image, m = RP.ref_render_train(field, ro, rd, ours.density_bitfield, ours.cascade, ours.grid_size, ours.aabb_train, 0.2)
        info["m"] = m
        if backward:
            ((image - target) ** 2).mean().backward()
            zero()

    ours.train()   # run_cuda's train branch (march_rays_train / composite_rays_train); the field stays frozen
    for p in ours.parameters():
        p.requires_grad_(False)

    def our3(backward=True):
        a = ro.detach().requires_grad_(backward)
        b = rd.detach().requires_grad_(backward)
        image = ours.render(a[None], b[None], bg_color=1.0, perturb=False, force_all_rays=True)["image"][0]
        if backward:
            ((image - target) ** 2).mean().backward()

    with torch.no_grad():
        rf, of = _time(lambda: ref3(False), 3), _time(lambda: our3(False), 3)
    rfb, ofb = _time(ref3, 3), _time(our3, 3)
    out["cfg3_estimator_cuda_ray_train"] = {
        "rays": N, "samples_per_ray": info["m"] / N,
        "ref_fwd_rays_per_s": N / rf * 1e3, "ours_fwd_rays_per_s": N / of * 1e3,
        "ref_fwd_bwd_rays_per_s": N / rfb * 1e3, "ours_fwd_bwd_rays_per_s": N / ofb * 1e3,
        "ours_vs_ref_fwd": rf / of, "ours_vs_ref_fwd_bwd": rfb / ofb,
        "note": "different backward targets: the reference's cuda_ray path has NO pose gradient — its backward goes to sigma / rgb -> "
                "MLPs -> table atomics (nerf/utils.py:446-450); ours goes to rays_o / rays_d (the estimator's gradient)"}
    ours.eval()

    # ---- BASELINE.md §3 comparator: cuda_ray=False `run` (512 uniform samples per ray), fwd + bwd TO THE RAYS on both sides
    ro1, rd1 = ro[:run_rays].contiguous(), rd[:run_rays].contiguous()
    tgt1 = target[:run_rays]

    def ref_run():
        a, b = ro1.detach().requires_grad_(True), rd1.detach().requires_grad_(True)
        image, _ = RP.ref_run(field, a, b, ours.aabb_infer, 0.2, num_steps=512)
        ((image - tgt1) ** 2).mean().backward()
        zero()

    nr = NeRFNetwork(encoding="hashgrid", bound=2, cuda_ray=False, density_scale=1, min_near=0.2).to(dev).eval()
    nr.load_state_dict({k: v for k, v in ours.state_dict().items() if k in nr.state_dict()}, strict=False)

    def our_run():
        a, b = ro1.detach().requires_grad_(True), rd1.detach().requires_grad_(True)
        image = nr.render(a[None], b[None], staged=True, bg_color=1.0, perturb=False, num_steps=512, upsample_steps=0,
                          max_ray_batch=4096)["image"][0]
        ((image - tgt1) ** 2).mean().backward()

    rr, orr = _time(ref_run, 3), _time(our_run, 3)
    out["cfg1_shape_run_512_fwd_bwd_to_rays"] = {
        "rays": run_rays, "samples_per_ray": 512, "ref_ms": rr, "ours_ms": orr, "ref_rays_per_s": run_rays / rr * 1e3,
        "ours_rays_per_s": run_rays / orr * 1e3, "ours_vs_ref": rr / orr,
        "note": "the reference-shaped estimator loss (renderer.py:125-254 over its CUDA GridEncoder / SHEncoder), gradient to "
                "rays_o / rays_d on both sides: like for like"}

    # ---- config 4: one 800x800 frame through the reference's eval loop vs the fused kernel
    fo, fd = closures.rays_from_poses(gt[None].to(dev), torch.arange(frame_rays, device=dev), (1111.1, 1111.1, 400.0, 400.0), 800)
    fo, fd = fo.reshape(-1, 3).contiguous(), fd.reshape(-1, 3).contiguous()
    ev = {}

    def ref4():
        ev["n"] = RP.ref_render_infer(field, fo, fd, ours.density_bitfield, ours.cascade, ours.grid_size, ours.aabb_infer, 0.2)[2]

    def our4():
        ev["s"] = render_rays_infer(ours, fo, fd, ours.aabb_infer, 1.0, return_stats=True)[3]

    r4, o4 = _time(ref4, 2, 1), _time(our4, 3)
    out["cfg4_render_frame"] = {"rays": frame_rays, "ref_ms": r4, "ours_ms": o4, "ref_frames_per_s": 1e3 / r4,
                                "ours_frames_per_s": 1e3 / o4, "ours_vs_ref": r4 / o4,
                                "ours_samples": int(ev["s"][1].item()), "ref_samples_incl_padding": int(ev["n"])}
    return out


def ffmlp_leg(dev, hbm_gbs, B=1 << 22):
    """FFMLP (ffmlp.py:99-169): the sigma net (in 32, 2 hidden layers) and the colour net (in 32, 3 hidden layers) of
    network_ff.py at B rows, inference and training forward + backward, tcgen05 kernels vs the reference's wmma build."""
    from nerf_navigation_b200 import _lib as ngp
    from oracle import build_ref
    L = ngp.lib()
    ref = build_ref.load_ref("ffmlp")
    st = torch.cuda.current_stream().cuda_stream
    out = {"rows": B, "what": "ngp_ffmlp_forward / ngp_ffmlp_backward (tcgen05.mma kind::f16, TMEM accumulators) on fp16 rows resident "
                              "in HBM; algorithmic bytes = rows x (inputs + outputs + saved activations) x 2 B; reference = its own "
                              "ffmlp.cu (wmma + CUTLASS 2.8 split-K) built for sm_100a" if ref is not None else
                              "tcgen05 FFMLP kernels; the reference's ffmlp build is not in the tree"}
    if ref is not None:
        ref.allocate_splitk(8)
    for name, input_dim, n in (("sigma_net_32_64_64_16", 32, 2), ("color_net_32_64_64_64_16", 32, 3)):
        nw = 64 * (input_dim + 64 * (n - 1) + 16)
        g = torch.Generator(device=dev).manual_seed(3)
        x = (torch.rand(B, input_dim, device=dev, generator=g) * 2 - 1).half()
        w = ((torch.rand(nw, device=dev, generator=g) * 2 - 1) * float(np.sqrt(3 / 64))).half()
        grad = torch.zeros(B, 16, device=dev, dtype=torch.half)
        grad[:, :3] = 0.05
        outp = torch.empty(B, 16, device=dev, dtype=torch.half)
        fwd = torch.empty(n, B, 64, device=dev, dtype=torch.half)
        gx = torch.empty(B, input_dim, device=dev, dtype=torch.half)
        gw = torch.empty(nw, device=dev, dtype=torch.half)
        wsb = L.ngp_ffmlp_backward_workspace_bytes(B, input_dim, n)
        wsp = torch.empty(wsb, device=dev, dtype=torch.uint8)

        def inf():
            ngp.check(L.ngp_ffmlp_forward(x.data_ptr(), w.data_ptr(), B, input_dim, 16, 64, n, 0, 6, None, outp.data_ptr(), st))

        def tf():
            ngp.check(L.ngp_ffmlp_forward(x.data_ptr(), w.data_ptr(), B, input_dim, 16, 64, n, 0, 6, fwd.data_ptr(), outp.data_ptr(), st))

        def tb():
            ngp.check(L.ngp_ffmlp_backward(grad.data_ptr(), x.data_ptr(), w.data_ptr(), fwd.data_ptr(), B, input_dim, 16, 64, n, 0, 6,
                                           1, None, gx.data_ptr(), gw.data_ptr(), wsp.data_ptr(), wsb, st))

        t_inf, t_f, t_b = _time(inf, 10, 3), _time(tf, 10, 3), _time(tb, 10, 3)
        by_inf = B * (input_dim + 16) * 2
        by_f = B * (input_dim + 16 + 64 * n) * 2
        by_b = B * (16 + input_dim + 64 * n + input_dim) * 2     # grad, inputs, saved activations in; grad_inputs out
        rec = {"inference_ms": t_inf, "inference_Mrows_per_s": B / t_inf / 1e3, "inference_GBps": by_inf / t_inf / 1e6,
               "inference_hbm_frac": by_inf / t_inf / 1e6 / hbm_gbs,
               "train_fwd_ms": t_f, "train_fwd_GBps": by_f / t_f / 1e6, "train_fwd_hbm_frac": by_f / t_f / 1e6 / hbm_gbs,
               "train_bwd_ms": t_b, "train_bwd_GBps": by_b / t_b / 1e6, "train_bwd_hbm_frac": by_b / t_b / 1e6 / hbm_gbs,
               "fwd_bwd_Mrows_per_s": B / (t_f + t_b) / 1e3}
        if ref is not None:
            r_bb = torch.zeros(n, B, 64, device=dev, dtype=torch.half)
            r_inf = torch.empty(B, 64, device=dev, dtype=torch.half)

            def rinf():
                ref.ffmlp_inference(x, w, B, input_dim, 16, 64, n, 0, 6, r_inf, outp)

            def rtf():
                ref.ffmlp_forward(x, w, B, input_dim, 16, 64, n, 0, 6, fwd, outp)

            def rtb():
                ref.ffmlp_backward(grad, x, w, fwd, B, input_dim, 16, 64, n, 0, 6, True, r_bb, gx, gw)

            ri, rf_, rb = _time(rinf, 10, 3), _time(rtf, 10, 3), _time(rtb, 5, 2)
            rec.update({"ref_inference_ms": ri, "ref_train_fwd_ms": rf_, "ref_train_bwd_ms": rb,
                        "ours_vs_ref_inference": ri / t_inf, "ours_vs_ref_train_fwd": rf_ / t_f, "ours_vs_ref_train_bwd": rb / t_b})
            del r_bb, r_inf
        out[name] = rec
        del x, w, grad, outp, fwd, gx, gw, wsp
    # network_ff.density (nerf/network_ff.py:76-89): hash-grid encoder -> FFMLP sigma net, as one kernel and as the two ops
    from nerf_navigation_b200.nerf.network_ff import NeRFNetwork as NetFF
    mff = NetFF(encoding="hashgrid", bound=2, cuda_ray=False).to(dev).eval()
    pts = torch.rand(B, 3, device=dev) * 2 - 1
    with torch.no_grad():
        mff.fused_eval = True
        t_fused = _time(lambda: mff.density(pts), 10, 3)
        mff.fused_eval = False
        t_two = _time(lambda: mff.density(pts), 10, 3)
    out["network_ff_density"] = {"points": B, "fused_ms": t_fused, "fused_Msamples_per_s": B / t_fused / 1e3,
                                 "encoder_then_ffmlp_ms": t_two, "speedup": t_two / t_fused,
                                 "l2_gather_GBps": B * 1024 / t_fused / 1e6,
                                 "what": "ngp_field_query_ff_forward (k_ffield5_forward: gather warps write fp16 features into the "
                                         "FFMLP's shared-memory A tile) vs ngp_grid_encode_forward + .half() + ngp_ffmlp_forward"}
    del mff, pts
    return out
