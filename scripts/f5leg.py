import sys, json, torch; sys.path.insert(0, "scripts"); import bench_legs as BL
d = BL.ffmlp_leg(torch.device("cuda", 0), 6529.1)
for k, v in d.items():
    if isinstance(v, dict): print(k, {a: round(b, 4) for a, b in v.items() if a in ("inference_ms", "train_fwd_ms")})
