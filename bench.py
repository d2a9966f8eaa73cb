#!/usr/bin/env python
"""bench.py — BASELINE.json configs[1]: hash-grid field query (16 levels x 2 feats, 2^19 table, 64-wide sigma MLP),
4,194,304 points per GPU, density forward + dL/dx backward (the planner's collision-cost step, SURVEY §8d cfg 2).

  python bench.py --gpus N --steps K --warmup W            # the B200 path (one rank per GPU under torchrun for N>1)
  python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU path (oracle port) on the host cores

Prints ONE JSON line (rank 0).  `value` = device-resident throughput (inputs in HBM), `e2e` = the same metric through
the public API (NeRFNetwork.density + autograd) with pinned HOST buffers and the H2D / D2H copies inside the timed
region.  Points are sharded by rank with replicated weights and no collective (weak scaling: 4M points per GPU).
"""
import argparse
import json
import os

os.environ.setdefault("OMP_WAIT_POLICY", "PASSIVE")  # two OpenMP runtimes share the host cores (torch, the C oracle)
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "field Msamples/s fwd+bwd"
UNIT = "Msamples/s"
WORKLOAD = "hashgrid_field_query_fwd_dLdx_bwd"
BOUND = 2.0
BYTES_PER_POINT_PASS = 16 * 8 * 2 * 4  # 1024 B of table gathers per point per pass


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--points", type=int, default=1 << 22, help="points per GPU per step")
    ap.add_argument("--cpu-points", type=int, default=None, help="points per CPU step (default: 2^18 for the cpu_baseline leg; "
                    "--impl reference uses the full --points when the run fits in ~4 min)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-estimator", action="store_true")
    ap.add_argument("--hypotheses", type=int, default=256)
    ap.add_argument("--estimator-steps", type=int, default=3)
    ap.add_argument("--no-render", action="store_true")
    ap.add_argument("--no-config1", action="store_true", help="skip the config-1 CPU legs (reference torch path)")
    ap.add_argument("--frames", type=int, default=64, help="config 4: 800x800 frames in the batch (whole job)")
    ap.add_argument("--frames-per-step", type=int, default=8, help="frames rendered per kernel launch on a rank")
    ap.add_argument("--render-steps", type=int, default=1)
    ap.add_argument("--e2e-chunk", type=int, default=1 << 20, help="points per pipelined chunk of the host-fed query")
    ap.add_argument("--no-train", action="store_true")
    ap.add_argument("--no-ref-cuda", action="store_true", help="skip the reference-CUDA-build leg (oracle/_ref)")
    ap.add_argument("--no-ffmlp", action="store_true")
    ap.add_argument("--train-steps", type=int, default=20)
    ap.add_argument("--train-rays", type=int, default=4096, help="config 5: rays per GPU per optimiser step")
    a = ap.parse_args()
    a.cpu_points_forced = a.cpu_points is not None
    if a.cpu_points is None:
        a.cpu_points = 1 << 18
    return a


def ncu_traffic(kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch of `kernel` from the committed `ncu --set full` capture
    of this same workload at the same 4,194,304 points (profiles/r2_dfield5_4m_ncu_keys.txt for the tcgen05 kernels,
    profiles/r1_field_4m_ncu_keys.txt for the mma.sync ones); None if the capture is not there."""
    cur, tot, launches = None, {}, {}
    for name in ("r2_dfield5_4m_ncu_keys.txt", "r1_field_4m_ncu_keys.txt"):
        p = os.path.join(ROOT, "profiles", name)
        if not os.path.exists(p):
            continue
        for line in open(p):
            if line.startswith("----"):
                cur = line
            elif cur and "dram__bytes_" in line and ".sum" in line:
                unit = line.split("[")[1].split("]")[0]
                val = float(line.split("=")[1].strip().replace(",", ""))
                mul = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)
                tot[cur] = tot.get(cur, 0.0) + val * mul
                if "dram__bytes_read" in line:                 # the capture may hold several launches of a kernel
                    launches[cur] = launches.get(cur, 0) + 1
    for k, v in tot.items():
        if kernel in k:
            return v / launches[k]                             # per launch, like `achieved`
    return None


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)), "measured"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}, "fallback"


# --------------------------------------------------------------------------------------------- CPU (oracle) leg
def cpu_field_setup(points, seed=0):
    from oracle import oracle as O
    O.set_num_threads(os.cpu_count() or 1)   # torchrun exports OMP_NUM_THREADS=1; the CPU arm uses every host core
    offsets, pls = O.level_table(3, 16, 2, 2.0, 16, 19, 2048 * BOUND)
    scales = O.level_scales(16, pls, 16)
    rng = np.random.default_rng(seed)
    table = rng.uniform(-1e-4, 1e-4, (int(offsets[-1]), 2)).astype(np.float32)
    W0 = (rng.uniform(-1, 1, (64, 32)) / np.sqrt(32)).astype(np.float32)
    W1 = (rng.uniform(-1, 1, (16, 64)) / np.sqrt(64)).astype(np.float32)
    x = rng.uniform(-1, 1, (points, 3)).astype(np.float32)
    gs = np.full(points, 2.0 / points, np.float32)

    def step():
        O.field_query(x, BOUND, table, offsets, scales, W0, W1, gs)

    return step, O.num_threads()


def cpu_config1(reps=5, warm=2, hash_points=1 << 18):
    """BASELINE.md §2 / north_star: the reference's CPU path (config 1) timed on this box's host cores —
    NeRFNetwork with frequency encoding + torch MLP + the non-cuda_ray renderer.run branch (oracle/cfg1_torch.py, a port
    pinned to the real reference by tests/golden/cfg1.npz), plus a torch transcription of the hash-grid lookup.
    2 warm-ups + 5 repetitions, median wall time, all host threads."""
    import torch
    from oracle import cfg1_torch as C1
    from oracle import oracle as O
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    field = C1.Cfg1Field.random(seed=0, bound=BOUND)
    g = torch.Generator().manual_seed(0)

    def med(fn, warm=warm, reps=reps):
        for _ in range(warm):
            fn()
        ts = []
        for _ in range(reps):
            t0 = time.perf_counter(); fn(); ts.append(time.perf_counter() - t0)
        return float(np.median(ts))

    x = torch.rand(2000, 3, generator=g) * 2 - 1

    def planner():
        xi = x.clone().requires_grad_(True)
        sigma, _ = field.density(xi)
        (sigma ** 2).mean().backward()

    # 1024 rays of an 800x800 pinhole at (0, 0, -1.5) looking +z
    pix = torch.randint(0, 640000, (1024,), generator=g)
    d = torch.stack((((pix % 800).float() + 0.5 - 400.0) / 1111.1, ((pix // 800).float() + 0.5 - 400.0) / 1111.1,
                     torch.ones(1024)), -1)
    d = d / d.norm(dim=-1, keepdim=True)
    o = torch.tensor([0.0, 0.0, -1.5]).expand(1024, 3).contiguous()
    target = torch.rand(1024, 3, generator=g)

    def estimator():
        oi, di = o.clone().requires_grad_(True), d.clone().requires_grad_(True)
        image, _ = C1.render_run(field, oi, di, num_steps=512, bg_color=1.0)
        ((image - target) ** 2).mean().backward()

    offsets, pls = O.level_table(3, 16, 2, 2.0, 16, 19, 2048 * BOUND)
    scales = O.level_scales(16, pls, 16)
    table = ((torch.rand(int(offsets[-1]), 2, generator=g) * 2 - 1) * 1e-4).requires_grad_(True)
    xh = torch.rand(hash_points, 3, generator=g)

    def hashgrid():
        xi = xh.clone().requires_grad_(True)
        C1.hashgrid_encode(xi, table, offsets, scales).sum().backward()
        table.grad = None

    tp, te, th = med(planner), med(estimator), med(hashgrid, 1, 3)
    return {"cores": cores, "kind": "port", "harness": f"{warm} warm-ups + {reps} reps, median wall, torch {torch.__version__} CPU",
            "planner_density": {"points": 2000, "ms": tp * 1e3, "Msamples_per_s": 2000 / tp / 1e6},
            "estimator_photometric": {"rays": 1024, "samples": 524288, "ms": te * 1e3, "rays_per_s": 1024 / te,
                                      "Msamples_per_s": 524288 / te / 1e6},
            "hashgrid_torch_transcription": {"points": hash_points, "ms": th * 1e3, "Msamples_per_s": hash_points / th / 1e6,
                                             "what": "encode fwd + autograd bwd to x and table; 1 warm-up + 3 reps"}}


def run_reference(args):
    """The reference's CPU implementation of the path = the oracle port (the reference's own CPU-capable code has no
    hash-grid path: its gridencoder is CUDA-only), all host threads, bounded sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cfg1 = None if args.no_config1 else cpu_config1()
    # points per step: the GPU arm's own 4,194,304 when the whole run fits in ~4 minutes on this host, else the largest
    # power-of-two fraction that does (the metric is per sample, so a bounded sample measures the same rate)
    probe_step, cores = cpu_field_setup(1 << 18)
    probe_step()
    t0 = time.perf_counter(); probe_step(); rate = (1 << 18) / (time.perf_counter() - t0)
    points = args.points
    while points > (1 << 18) and points / rate * (args.steps + max(args.warmup, 1)) > 240.0:
        points >>= 1
    if args.cpu_points_forced:
        points = args.cpu_points
    step, cores = cpu_field_setup(points)
    for _ in range(max(args.warmup, 1)):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = time.perf_counter() - t0
    val = points * args.steps / dt / 1e6
    sample = f"{points} of {args.points} uniform points per step, fwd + dL/dx bwd"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "points_per_gpu": args.points, "points_per_step": points, "levels": 16, "feats": 2,
                   "log2_table": 19, "mlp": "32-64-16", "bound": BOUND, "points": "uniform[-1,1]^3"},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "config1_cpu": cfg1,
    }))


# --------------------------------------------------------------------------------------------- clocks sampler
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap,power.draw")

    def __init__(self, index):
        self.samples = []
        self.proc = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "50", "-i", str(index)], stdout=subprocess.PIPE, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append((time.perf_counter(), line.strip()))

    def stop(self, t_begin, t_end):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.06)
        self.proc.terminate()
        mhz, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for t, line in self.samples:
            if not (t_begin <= t <= t_end + 0.05):
                continue
            f = [s.strip() for s in line.split(",")]
            try:
                mhz.append(float(f[0])); mx = float(f[1])
            except Exception:
                continue
            for nm, v in zip(names, f[2:6]):
                if v == "Active":
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(mhz)) if mhz else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(mhz)}


# --------------------------------------------------------------------------------------------- B200 leg
def run_b200(args):
    import torch
    import torch.distributed as dist
    from nerf_navigation_b200 import _lib as ngp
    from nerf_navigation_b200.nerf.network import NeRFNetwork

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py --impl b200 needs a CUDA device (there is no CPU fallback)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    torch.manual_seed(0)
    model = NeRFNetwork(encoding="hashgrid", bound=BOUND, cuda_ray=False).to(dev).eval()
    for p_ in model.parameters():          # the planner / estimator query a frozen field (gradients to the points only)
        p_.requires_grad_(False)
    enc = model.encoder
    B = args.points
    g = torch.Generator(device=dev).manual_seed(1234 + rank)  # each rank owns its own slice of the point set
    x = torch.rand(B, 3, device=dev, generator=g) * 2 - 1      # uniform in [-1,1]^3: worst-case locality
    sigma = torch.empty(B, device=dev); geo = torch.empty(B, 15, device=dev)
    mask = torch.empty(B, device=dev, dtype=torch.int64); gs = torch.empty(B, device=dev); gx = torch.empty(B, 3, device=dev)
    W0 = model.sigma_net[0].weight.detach().contiguous(); W1 = model.sigma_net[1].weight.detach().contiguous()
    tab = enc.embeddings.detach(); off = enc.offsets
    S = float(np.log2(enc.per_level_scale)); H = int(enc.base_resolution)
    L = ngp.lib()
    st = torch.cuda.current_stream().cuda_stream

    def fwd():
        ngp.check(L.ngp_field_query_forward(x.data_ptr(), BOUND, tab.data_ptr(), off.data_ptr(), S, H, 0, 0, W0.data_ptr(),
                                            W1.data_ptr(), sigma.data_ptr(), geo.data_ptr(), mask.data_ptr(), B, st))

    def bwd():
        ngp.check(L.ngp_field_query_backward(x.data_ptr(), BOUND, tab.data_ptr(), off.data_ptr(), S, H, 0, 0, W0.data_ptr(),
                                             W1.data_ptr(), sigma.data_ptr(), mask.data_ptr(), gs.data_ptr(), None,
                                             gx.data_ptr(), B, st))

    jac = torch.empty(B, 3, device=dev)

    def fwd_jac():   # what density(x) launches when x requires a gradient (field._field_query.forward)
        ngp.check(L.ngp_field_query_forward_jacobian(x.data_ptr(), BOUND, tab.data_ptr(), off.data_ptr(), S, H, 0, 0,
                                                     W0.data_ptr(), W1.data_ptr(), sigma.data_ptr(), geo.data_ptr(),
                                                     mask.data_ptr(), jac.data_ptr(), B, st))

    def bwd_jac():   # ... and its backward for a gradient arriving through sigma (field._field_query.backward)
        ngp.check(L.ngp_field_query_backward_jacobian(jac.data_ptr(), gs.data_ptr(), gx.data_ptr(), B, st))

    def step():      # one field query: sigma, geo_feat and dL/dx for loss = mean(sigma^2) (quad_plot.py:237 collision cost)
        fwd_jac()
        torch.mul(sigma, 2.0 / B, out=gs)
        bwd_jac()

    def two_pass_step():   # the same result with a second table pass (the reference's structure: forward op, backward op)
        fwd()
        torch.mul(sigma, 2.0 / B, out=gs)
        bwd()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    ev = lambda: torch.cuda.Event(enable_timing=True)
    for _ in range(args.warmup):
        step()
    barrier()
    sampler = ClockSampler(local) if rank == 0 else None
    t_begin = time.perf_counter()
    # ---- device-resident timed region: K steps, per-kernel events on the launching stream
    e_fj = [(ev(), ev()) for _ in range(args.steps)]
    e_bj = [(ev(), ev()) for _ in range(args.steps)]
    e0, e1 = ev(), ev()
    barrier()
    e0.record()
    for i in range(args.steps):
        e_fj[i][0].record(); fwd_jac(); e_fj[i][1].record()
        torch.mul(sigma, 2.0 / B, out=gs)
        e_bj[i][0].record(); bwd_jac(); e_bj[i][1].record()
    e1.record()
    barrier()
    t_dev = e0.elapsed_time(e1) / 1e3
    fj_ms = float(np.mean([a.elapsed_time(b) for a, b in e_fj]))
    bj_ms = float(np.mean([a.elapsed_time(b) for a, b in e_bj]))

    # ---- the two-table-pass form of the same step (forward kernel, backward kernel): round 1's headline
    for _ in range(3):
        two_pass_step()
    e_fwd = [(ev(), ev()) for _ in range(args.steps)]
    e_bwd = [(ev(), ev()) for _ in range(args.steps)]
    e2, e3 = ev(), ev()
    barrier()
    e2.record()
    for i in range(args.steps):
        e_fwd[i][0].record(); fwd(); e_fwd[i][1].record()
        torch.mul(sigma, 2.0 / B, out=gs)
        e_bwd[i][0].record(); bwd(); e_bwd[i][1].record()
    e3.record()
    barrier()
    t_two = e2.elapsed_time(e3) / 1e3
    fwd_ms = float(np.mean([a.elapsed_time(b) for a, b in e_fwd]))
    bwd_ms = float(np.mean([a.elapsed_time(b) for a, b in e_bwd]))

    # ---- the same step as ONE launch (ngp_field_query_value_grad: the upstream factor given in the affine form a + b sigma)
    def vg_step():
        ngp.check(L.ngp_field_query_value_grad(x.data_ptr(), BOUND, tab.data_ptr(), off.data_ptr(), S, H, 0, 0, W0.data_ptr(),
                                               W1.data_ptr(), None, None, 0.0, 2.0 / B, sigma.data_ptr(), geo.data_ptr(),
                                               gx.data_ptr(), B, st))
    for _ in range(3):
        vg_step()
    ev0, ev1 = ev(), ev()
    barrier()
    ev0.record()
    for _ in range(args.steps):
        vg_step()
    ev1.record()
    barrier()
    t_vg = ev0.elapsed_time(ev1) / 1e3

    # ---- locality legs (rank-0 figures, same kernels): SURVEY §8d config 2 (ii) trajectory-clustered points — 8-cubed body
    # clouds (quad_plot.py:224-250) around waypoints on a helix — and the uniform batch pre-sorted along a Morton curve
    # (the upper bound a caller gets by handing over spatially ordered queries; results are order-independent)
    x_uniform = x
    def timed_steps(pts):
        nonlocal x
        x = pts
        for _ in range(3):
            step()
        a, b = ev(), ev()
        barrier()
        a.record()
        for _ in range(args.steps):
            step()
        b.record()
        barrier()
        x = x_uniform
        return world * B * args.steps / (a.elapsed_time(b) / 1e3) / 1e6
    locality = None
    if B % 512 == 0:
        nw = B // 512
        k = torch.arange(nw, device=dev, dtype=torch.float32) / nw
        way = torch.stack([0.6 * torch.cos(6 * np.pi * k), 0.6 * torch.sin(6 * np.pi * k), 0.1 + 0.5 * k], -1)
        lx = torch.linspace(-0.05, 0.05, 8, device=dev); lz = torch.linspace(-0.02, 0.02, 8, device=dev)
        body = torch.stack(torch.meshgrid(lx, lx, lz, indexing="ij"), -1).reshape(-1, 3)
        xc = (way[:, None, :] + body[None, :, :]).reshape(-1, 3).contiguous()
        q = ((x_uniform + BOUND) / (2 * BOUND) * 256).long().clamp_(0, 255)
        key = torch.zeros(B, dtype=torch.long, device=dev)
        for i in range(8):
            for d in range(3):
                key |= ((q[:, d] >> i) & 1) << (3 * i + d)
        xs = x_uniform[torch.argsort(key)].contiguous()
        locality = {"unit": UNIT, "clustered": timed_steps(xc), "uniform_morton_sorted": timed_steps(xs),
                    "what": "the headline step on (a) trajectory-clustered body clouds (config 2 (ii)), (b) the uniform batch handed "
                            "over in Morton order (8 bits/axis; sorting not timed); `value` is the uniform batch in arrival order"}
        del xc, xs, q, key

    # ---- end-to-end: public API, pinned host input, H2D + D2H inside the timed region
    # The rank's thread is bound to the CPUs of its GPU's NUMA node for this leg (nerf_navigation_b200/hostmem.py): the
    # pinned buffers below and the ones the host-fed query allocates are first touched here, so they land on the
    # memory of the socket the GPU hangs off (eight ranks otherwise share one socket's controllers and the
    # inter-socket link).  Affinity is restored after the leg (the CPU legs use every core).
    from nerf_navigation_b200 import hostmem
    numa_info = hostmem.describe(dev)
    numa_ctx = hostmem.numa_local(dev)
    numa_ctx.__enter__()
    xh = x.cpu().pin_memory()
    gxh = torch.empty(B, 3).pin_memory()
    xd = torch.empty(B, 3, device=dev)

    def e2e_autograd_step():   # the reference-shaped call on device tensors, copies serialised around it
        xd.copy_(xh, non_blocking=True)
        xin = xd.detach().requires_grad_(True)
        s = model.density(xin)["sigma"]
        loss = (s ** 2).mean()
        loss.backward()
        gxh.copy_(xin.grad, non_blocking=True)
        return loss.item()

    def e2e_dropin_step():     # THE drop-in call: the reference-level API on points that live in pinned host memory
        xin = xh.detach().requires_grad_(True)
        s = model.density(xin)["sigma"]          # uploads pipelined with the forward kernels (field._field_query_host)
        loss = (s ** 2).mean()
        loss.backward()                          # backward kernels pipelined with the download of x.grad (pinned host)
        return loss.item(), xin.grad

    from nerf_navigation_b200.field import HostFieldQuery
    hq = HostFieldQuery(model, B, chunk=args.e2e_chunk)
    loss_acc = torch.zeros((), device=dev)

    def dloss(sig):            # loss = mean(sigma^2): accumulate the loss, return dL/dsigma for the chunk
        loss_acc.add_((sig * sig).sum())
        return sig * (2.0 / B)

    def sq_hook(sig):
        loss_acc.add_((sig * sig).sum())

    def e2e_step():            # host-resident points -> dL/dx on the host + the loss scalar, chunks pipelined over PCIe;
        loss_acc.zero_()       # loss = mean(sigma^2) given as dL/dsigma = 0 + (2/B) sigma -> one kernel per chunk
        hq(xh, (0.0, 2.0 / B), None, gxh, sigma_hook=sq_hook)
        return loss_acc.item() / B

    def e2e_callable_step():   # the same with an arbitrary elementwise upstream gradient (forward + backward kernels)
        loss_acc.zero_()
        hq(xh, dloss, None, gxh)
        return loss_acc.item() / B

    e2e_steps = max(3, min(args.steps, 30))
    for _ in range(3):
        e2e_step(); e2e_autograd_step(); e2e_callable_step(); e2e_dropin_step()
    barrier()
    ed0, ed1 = ev(), ev()
    ed0.record()
    for _ in range(e2e_steps):
        e2e_dropin_step()
    ed1.record()
    barrier()
    t_e2e_dropin = ed0.elapsed_time(ed1) / 1e3
    ec0, ec1 = ev(), ev()
    ec0.record()
    for _ in range(e2e_steps):
        e2e_callable_step()
    ec1.record()
    barrier()
    t_e2e_callable = ec0.elapsed_time(ec1) / 1e3
    ea0, ea1 = ev(), ev()
    ea0.record()
    for _ in range(e2e_steps):
        e2e_autograd_step()
    ea1.record()
    barrier()
    t_e2e_autograd = ea0.elapsed_time(ea1) / 1e3
    e2, e3 = ev(), ev()
    e2.record()
    for _ in range(e2e_steps):
        e2e_step()
    e3.record()
    barrier()
    t_e2e = e2.elapsed_time(e3) / 1e3
    numa_ctx.__exit__(None, None, None)
    t_end = time.perf_counter()
    clocks = sampler.stop(t_begin, t_end) if sampler else None

    # ---- BASELINE config 3 (second half of the metric): pose-estimator photometric loss, 256 hypotheses x 4096 rays
    # sharded by hypothesis over the ranks (strong scaling, no collective), forward + backward to the poses
    est = None
    if not args.no_estimator:
        sys.path.insert(0, os.path.join(ROOT, "scripts"))
        import bench_estimator as BE
        from nerf_navigation_b200.sharding import shard_range
        emodel, eposes, epix, etarget, ekw = BE.setup(dev, args.hypotheses, 4096)
        lo, hi = shard_range(args.hypotheses, rank, world)
        eposes = eposes[lo:hi]
        estates = emodel.estimator_states[lo:hi].contiguous()
        BE.step_states(emodel, estates, epix, etarget, ekw)
        barrier()
        e4, e5 = ev(), ev()
        e4.record()
        for _ in range(args.estimator_steps):
            elosses, egrads = BE.step_states(emodel, estates, epix, etarget, ekw)
        e5.record()
        barrier()
        t_est = e4.elapsed_time(e5) / 1e3 / args.estimator_steps
        with torch.no_grad():   # samples per ray of the scene (one chunk of hypotheses), for the roofline fraction
            from nerf_navigation_b200 import closures as _cl, raymarching as _rm
            _ro, _rd = _cl.rays_from_poses(eposes[:8], epix, BE.INTR, BE.W)
            _ro, _rd = _ro.reshape(-1, 3).contiguous(), _rd.reshape(-1, 3).contiguous()
            _n, _f = _rm.near_far_from_aabb(_ro, _rd, emodel.aabb_train, emodel.min_near)
            _x = _rm.march_rays_train(_ro, _rd, emodel.bound, emodel.density_bitfield, emodel.cascade, emodel.grid_size,
                                      _n, _f, None, -1, False, -1, True, 0, 1024)[0]
            est_spr = _x.shape[0] / _ro.shape[0]
            del _ro, _rd, _n, _f, _x
        del emodel

    # ---- BASELINE config 4: full-frame cuda_ray render, 800x800 x `frames` poses on an orbit, inference path through the
    # fused kernel (csrc/render.cu); frames sharded by rank (strong scaling, no collective)
    t_ren, ren_samples, ren_frames = 0.0, 0, 0
    if not args.no_render:
        sys.path.insert(0, os.path.join(ROOT, "scripts"))
        import bench_render as BR
        from nerf_navigation_b200.field import render_rays_infer
        from nerf_navigation_b200.sharding import shard_range
        flo, fhi = shard_range(args.frames, rank, world)
        ren_frames = fhi - flo
        chunks = [(a, min(a + args.frames_per_step, fhi)) for a in range(flo, fhi, args.frames_per_step)]
        rmodel, rays = None, []
        for a, b in chunks:   # rays of every chunk are resident before the timed region (12 + 12 B per ray)
            rmodel, ro, rd = BR.setup(dev, b - a, first=a, total=args.frames)
            rays.append((ro, rd))
        for ro, rd in rays[:1]:
            BR.step(rmodel, ro, rd)
        barrier()
        e6, e7 = ev(), ev()
        e6.record()
        for _ in range(args.render_steps):
            for ro, rd in rays:
                BR.step(rmodel, ro, rd)
        e7.record()
        barrier()
        t_ren = e6.elapsed_time(e7) / 1e3 / args.render_steps
        for ro, rd in rays:
            ren_samples += int(render_rays_infer(rmodel, ro, rd, rmodel.aabb_infer, 1.0, return_stats=True)[3][1].item())
        del rmodel, rays

    # ---- BASELINE config 5: training step, 4096 rays per GPU (weak scaling), table + MLP gradient all-reduce over NCCL
    t_trn, trn_loss, trn_parity = 0.0, (0.0, 0.0), None
    if not args.no_train:
        sys.path.insert(0, os.path.join(ROOT, "scripts"))
        import bench_train as BT
        tmodel, topt, tro, trd, ttarget = BT.setup(dev, args.train_rays, rank)
        first = [float(BT.step(tmodel, topt, tro, trd, ttarget, world)) for _ in range(5)][0]
        barrier()
        e8, e9 = ev(), ev()
        e8.record()
        for _ in range(args.train_steps):
            tl = BT.step(tmodel, topt, tro, trd, ttarget, world)
        e9.record()
        barrier()
        t_trn = e8.elapsed_time(e9) / 1e3 / args.train_steps
        trn_loss = (first, float(tl))
        # the same step with the reference's own table-gradient method (float atomics: not bit-reproducible), rank-0 figure
        from nerf_navigation_b200.gridencoder import table_grad
        table_grad.DETERMINISTIC = False
        try:
            for _ in range(3):
                BT.step(tmodel, topt, tro, trd, ttarget, world)
            barrier()
            e10, e11 = ev(), ev()
            e10.record()
            for _ in range(args.train_steps):
                BT.step(tmodel, topt, tro, trd, ttarget, world)
            e11.record()
            barrier()
            t_trn_atomic = e10.elapsed_time(e11) / 1e3 / args.train_steps
        finally:
            table_grad.DETERMINISTIC = True
        trn_bytes = sum(p.numel() for p in tmodel.parameters() if p.requires_grad) * 4
        del tmodel, topt
        trn_parity = BT.parity(dev, args.train_rays, rank, world) if world > 1 else None

    if world > 1:
        tt = torch.tensor([t_dev, t_e2e, 0.0 if args.no_estimator else t_est, t_ren, t_trn, t_e2e_dropin], device=dev,
                          dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        t_dev, t_e2e, t_est_max, t_ren, t_trn, t_e2e_dropin = tt.tolist()
        if not args.no_estimator:
            t_est = t_est_max
        cnt = torch.tensor([ren_samples], device=dev, dtype=torch.int64)
        dist.all_reduce(cnt)
        ren_samples = int(cnt.item())
    if not args.no_estimator:
        est = {"metric": "estimator rays/s (photometric loss fwd + bwd to the 12-state)", "value": args.hypotheses * 4096 / t_est,
               "unit": "rays/s", "ms_per_step": t_est * 1e3, "hypotheses": args.hypotheses, "rays_per_hypothesis": 4096,
               "scaling": "strong", "sharding": "pose hypotheses by rank, replicated field, no collective",
               "scene": "synthetic stonehenge occupancy (128^3 x 2 cascades), random-init field, ~75 samples/ray",
               "path": "12-states -> poses -> rays of the sampled pixels (one kernel) -> near_far -> march_rays_train -> fused tcgen05 field+colour "
                       "kernels (csrc/cfield5.cu) -> composite_rays_train -> per-hypothesis MSE, and backward to the 12-states; all "
                       "hypotheses of the rank in ONE launch chain (no Python loop over chunks)",
               "loss_gt_hypothesis": float(elosses[0]) if rank == 0 else None, "grad_state_abs_mean": float(egrads.abs().mean()),
               "samples_per_ray": est_spr, "reference_cuda_build_same_box": "ref_cuda_build / estimator_vs_ref_cuda in this line"}
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- L2-gather roofline measured live: independent 8-byte corner gathers from a table-sized buffer
    probe_buf = torch.randint(0, 2 ** 31 - 1, (tab.numel(),), device=dev, dtype=torch.int32)
    sink = torch.zeros(4, device=dev, dtype=torch.int32)
    pb_blocks, pb_iters = 148 * 16, 64
    probe = lambda: ngp.check(L.ngp_probe_gather(probe_buf.data_ptr(), probe_buf.numel() * 4, 8, pb_iters, 1, pb_blocks,
                                                 sink.data_ptr(), st))
    for _ in range(3):
        probe()
    pe = [(ev(), ev()) for _ in range(5)]
    for a, b in pe:
        a.record(); probe(); b.record()
    torch.cuda.synchronize()
    probe_ms = min(a.elapsed_time(b) for a, b in pe)
    gather_peak = pb_blocks * 256 * pb_iters * 8 * 8 / (probe_ms / 1e3) / 1e9
    probe0 = lambda: ngp.check(L.ngp_probe_gather(probe_buf.data_ptr(), probe_buf.numel() * 4, 8, pb_iters, 0, pb_blocks,
                                                  sink.data_ptr(), st))
    for _ in range(3):
        probe0()
    pe0 = [(ev(), ev()) for _ in range(5)]
    for a, b in pe0:
        a.record(); probe0(); b.record()
    torch.cuda.synchronize()
    random8_peak = pb_blocks * 256 * pb_iters * 8 * 8 / (min(a.elapsed_time(b) for a, b in pe0) / 1e3) / 1e9
    del probe_buf

    peaks, peak_kind = measured_peaks()
    value = world * B * args.steps / t_dev / 1e6
    e2e_val = world * B * e2e_steps / t_e2e_dropin / 1e6
    tc5 = bool(L.ngp_set_field_impl(-1))
    KF, KB = ("k_dfield5_forward", "k_dfield5_backward") if tc5 else ("k_field_forward", "k_field_backward")
    KJ = "k_dfield5_value_grad" if tc5 else "k_field_forward + k_field_value_grad"
    # dominant kernel of the headline step: the Jacobian-returning forward (one table pass per point)
    dom, dom_ms = KJ, fj_ms
    achieved = B * BYTES_PER_POINT_PASS * (1 if tc5 else 2.25) / (dom_ms / 1e3) / 1e9
    traffic = ncu_traffic(dom)
    # the kernel of this bench that sits lowest against the same roofline (not the longest one of the headline leg)
    fracs = {KJ: achieved / gather_peak,
             KF: B * BYTES_PER_POINT_PASS / (fwd_ms / 1e3) / 1e9 / gather_peak,
             KB: B * BYTES_PER_POINT_PASS / (bwd_ms / 1e3) / 1e9 / gather_peak}
    if not args.no_render and t_ren > 0:
        fracs["render step (k_r5_premarch + k_render5_rays_infer_lists)"] = ren_samples * BYTES_PER_POINT_PASS / t_ren / 1e9 / world / gather_peak
    if not args.no_estimator:
        fracs["estimator step (k_cfield5_forward + k_cfield5_backward + marcher / compositor)"] = \
            args.hypotheses * 4096 * est_spr * 2 * BYTES_PER_POINT_PASS / t_est / 1e9 / world / gather_peak
    low = min(fracs, key=fracs.get)
    lowest = {"kernel": low, "frac": fracs[low], "all": fracs}
    out = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": t_dev / args.steps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "points_per_gpu": B, "levels": 16, "feats": 2, "log2_table": 19,
                   "mlp": "32-64-16", "bound": BOUND, "points": "uniform[-1,1]^3", "sharding": "points by rank, replicated weights, no collective",
                   "l2": "streams ~400 MB/step (> 126 MB L2) of points/outputs; the 48 MiB table is L2-resident by design"},
        "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": B * 12, "d2h_bytes_per_step": B * 12 + 4,
                "steps": e2e_steps,
                "numa": dict(numa_info, what="rank 0's GPU -> NUMA node (sysfs); each rank binds its thread to its node's CPUs "
                                             "for this leg so the pinned buffers are first touched there (hostmem.numa_local)"),
                "api": "the reference-level call on host data: x = pinned host tensor; s = NeRFNetwork.density(x)['sigma']; "
                       "(s ** 2).mean().backward(); x.grad (pinned host) + loss.item() — uploads pipelined with the "
                       "(Jacobian-returning) forward kernels, downloads with the backward kernels, inside density() itself",
                "host_query_api": {"value": world * B * e2e_steps / t_e2e / 1e6,
                                   "api": f"field.HostFieldQuery(model)(x_pinned, (0, 2/B), None, grad_pinned, sigma_hook): chunks of "
                                          f"{args.e2e_chunk} points, H2D / single-pass value+grad kernel / D2H on 3 streams (the "
                                          "upstream gradient restricted to a0 + b0 sigma)"},
                "callable_api": {"value": world * B * e2e_steps / t_e2e_callable / 1e6,
                                 "api": "same call with dloss_dsigma as a Python callable: Jacobian-returning forward (one table pass) + dL/dsigma * jac per chunk"},
                "autograd_api": {"value": world * B * e2e_steps / t_e2e_autograd / 1e6,
                                 "api": "x.cuda() -> NeRFNetwork.density(x)['sigma'] -> mean(sigma^2).backward() -> x.grad.cpu()"}},
        "gpu_launches": 2 * args.steps,  # device-resident region: forward_jacobian + backward_jacobian per step
        "step": "ngp_field_query_forward_jacobian (sigma, geo_feat, ReLU bits, d sigma / d x: ONE table pass) -> dL/dsigma = "
                "2 sigma / B -> ngp_field_query_backward_jacobian (dL/dx = dL/dsigma * jac): the launches "
                "NeRFNetwork.density(x) + backward() make for a point tensor that requires a gradient",
        "kernels_ms": {KJ: fj_ms, "k_scale_rows3": bj_ms, KF: fwd_ms, KB: bwd_ms,
                       "ngp_field_query_value_grad": t_vg / args.steps * 1e3},
        "locality": locality,
        "two_pass": {"value": world * B * args.steps / t_two / 1e6, "unit": UNIT, "ms_per_step": t_two / args.steps * 1e3,
                     "what": "the same outputs from ngp_field_query_forward + ngp_field_query_backward (two table passes, "
                             "the reference's op structure and round 1's headline; rank-0 time); still what runs when a "
                             "gradient arrives through geo_feat"},
        "single_pass": {"value": world * B * args.steps / t_vg / 1e6, "unit": UNIT,
                        "what": "ngp_field_query_value_grad: sigma + dL/dx in ONE launch when the upstream gradient has the "
                                "form a_i + b_i sigma_i (rank-0 time)"},
        "roofline": {"bound": "l2_gather", "kernel": dom, "achieved": achieved, "peak": gather_peak, "unit": "GB/s",
                     "frac": achieved / gather_peak, "traffic": traffic,
                     "peak_source": "ngp_probe_gather live on this GPU: 8-byte hashed-cell corner gathers from a 48 MiB "
                                    "L2-resident buffer (the kernels' own access shape); MEASURED_PEAKS.json has no L2 figure",
                     "peak_random8B": random8_peak,
                     "hbm_peak": peaks["hbm_gbs"], "hbm_peak_source": peak_kind,
                     "hbm_frac": (traffic / (dom_ms / 1e3) / 1e9 / peaks["hbm_gbs"]) if traffic else None,
                     "note": "algorithmic bytes = 1024 B of table gathers per point per table pass (the headline step makes ONE "
                             "pass: value and Jacobian from the same corners), served by L2 (table resident): the "
                             "binding limit is the L1->L2 request rate, not HBM (DRAM traffic per launch = `traffic`, ~1.07x the "
                             "minimal point/output streaming, a few % of the HBM peak).  The peak is a PATTERN ceiling for "
                             "arrival-ordered uniform points: spatially clustered batches exceed it (locality.clustered)",
                     "lowest": lowest},
        "l2_gather_roofline": {
            "peak": gather_peak, "unit": "GB/s useful", "how": "ngp_probe_gather live: 8-byte hashed-cell corner gathers "
            "from a 48 MiB L2-resident buffer, all SMs (the binding unit is one L1-miss sector per clock per SM)",
            KJ: {"achieved": achieved, "frac": achieved / gather_peak},
            KF: {"achieved": B * BYTES_PER_POINT_PASS / (fwd_ms / 1e3) / 1e9,
                                "frac": B * BYTES_PER_POINT_PASS / (fwd_ms / 1e3) / 1e9 / gather_peak},
            KB: {"achieved": B * BYTES_PER_POINT_PASS / (bwd_ms / 1e3) / 1e9,
                                 "frac": B * BYTES_PER_POINT_PASS / (bwd_ms / 1e3) / 1e9 / gather_peak}},
        "clocks": clocks,
    }
    out["estimator"] = est
    if not args.no_train:
        out["train"] = {"metric": "training step (config 5)", "rays_per_gpu": args.train_rays, "ms_per_step": t_trn * 1e3,
                        "steps_per_s": 1.0 / t_trn, "rays_per_s": world * args.train_rays / t_trn, "scaling": "weak",
                        "loss_first": trn_loss[0], "loss_last": trn_loss[1], "parity": trn_parity,
                        "train_parity": None if trn_parity is None else trn_parity["train_parity"],
                        "collective": "none (1 GPU)" if world == 1 else f"NCCL all-reduce (sum) of {trn_bytes} gradient bytes per step, 1/world folded into the Adam kernel",
                        "path": "march_rays_train -> GridEncoder/SH kernels + torch Linear -> composite_rays_train; "
                                "backward incl. deterministic table gradient; fused Adam + EMA",
                        "atomic_table_grad": {"ms_per_step": t_trn_atomic * 1e3, "steps_per_s": 1.0 / t_trn_atomic,
                                              "what": "gridencoder.table_grad.DETERMINISTIC = False: float atomics like the "
                                                      "reference (not bit-reproducible), rank-0 time"}}
    if not args.no_render:
        nrays = args.frames * 800 * 800
        # per-sample bound of the fused kernel: 1024 B of L2 gathers (same unit as the field query)
        out["render"] = {
            "metric": "full-frame render (config 4)", "frames": args.frames, "rays": nrays, "ms_per_step": t_ren * 1e3,
            "ms_per_frame": t_ren * 1e3 / args.frames, "frames_per_s": args.frames / t_ren, "rays_per_s": nrays / t_ren,
            "samples": ren_samples, "samples_per_ray": ren_samples / nrays, "samples_per_s": ren_samples / t_ren,
            "scaling": "strong", "sharding": "frames by rank, replicated field + bitfield, no collective",
            "path": "field.render_rays_infer -> ngp_render_rays_infer_segments: thread-per-ray marching pre-pass at full occupancy "
                    "(records each ray's runs of samples), then near/far + field + colour + composite + post in one persistent "
                    "warp-specialised tcgen05 kernel whose feeder warps replay the runs (csrc/render5.cu), then the marcher-inside "
                    "kernel for rays with more than 32 runs (none here); no host read; the whole call is timed",
            "gpu_launches": 3 * args.render_steps * ((ren_frames + args.frames_per_step - 1) // args.frames_per_step),
            "l2_gather": {"achieved": ren_samples * BYTES_PER_POINT_PASS / t_ren / 1e9 / world, "peak": gather_peak,
                          "unit": "GB/s useful per GPU",
                          "frac": ren_samples * BYTES_PER_POINT_PASS / t_ren / 1e9 / world / gather_peak}}
    if world == 1 and not args.no_cpu_baseline:
        cfg1 = None if args.no_config1 else cpu_config1()
        cstep, cores = cpu_field_setup(args.cpu_points)
        cstep()
        reps = 8
        t0 = time.perf_counter()
        for _ in range(reps):
            cstep()
        dt = time.perf_counter() - t0
        out["cpu_baseline"] = {"value": args.cpu_points * reps / dt / 1e6, "unit": UNIT, "cores": cores, "kind": "port",
                               "sample": f"{reps} x {args.cpu_points} of the {B} points, fwd + dL/dx bwd, C/OpenMP oracle"}
        if cfg1 is not None:
            out["cpu_baseline"]["config1"] = cfg1
    # ---- strong-scaling legs as top-level keys (the headline `value` is weak scaling of independent points): whole-job
    # throughput at this N, so that value(N) / value(1) / N is the efficiency of each leg
    out["estimator_rays_per_s"] = None if est is None else est["value"]
    out["render_frames_per_s"] = None if args.no_render else out["render"]["frames_per_s"]
    out["train_rays_per_s"] = None if args.no_train else out["train"]["rays_per_s"]
    out["strong_scaling"] = {"estimator_rays_per_s": out["estimator_rays_per_s"], "render_frames_per_s": out["render_frames_per_s"],
                             "what": "fixed total work (256 hypotheses x 4096 rays; 64 frames) split over n_gpus ranks, no collective; "
                                     "efficiency = value(N) / (N x value(1))"}
    if world == 1:
        sys.path.insert(0, os.path.join(ROOT, "scripts"))
        import bench_legs as BL
        if not args.no_ref_cuda:
            try:
                out["ref_cuda_build"] = BL.ref_cuda_build_leg(dev)
                rc = out["ref_cuda_build"]
                if "cfg3_estimator_cuda_ray_train" in rc:
                    out["estimator_vs_ref_cuda"] = {
                        "cuda_ray_train_fwd_bwd": rc["cfg3_estimator_cuda_ray_train"]["ours_vs_ref_fwd_bwd"],
                        "cuda_ray_train_fwd": rc["cfg3_estimator_cuda_ray_train"]["ours_vs_ref_fwd"],
                        "run_512_fwd_bwd_to_rays": rc["cfg1_shape_run_512_fwd_bwd_to_rays"]["ours_vs_ref"],
                        "what": "ours / reference CUDA build, rays/s, same box and run; first = the cuda_ray train branch (ours "
                                "back-propagates to the rays, the reference to its parameters: it has no pose gradient there), last = "
                                "the reference-shaped run() with 512 samples per ray, gradient to the rays on both sides"}
            except Exception as e:   # a reported baseline must not take the bench line down
                out["ref_cuda_build"] = {"error": f"{type(e).__name__}: {e}"[:300]}
        if not args.no_ffmlp:
            try:
                out["ffmlp"] = BL.ffmlp_leg(dev, peaks["hbm_gbs"])
            except Exception as e:
                out["ffmlp"] = {"error": f"{type(e).__name__}: {e}"[:300]}
    print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
