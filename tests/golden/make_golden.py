"""Generate tests/golden/*.npz by running the UNMODIFIED reference on seeded inputs.

Run in the build container only (``/root/reference`` must exist):
    python tests/golden/make_golden.py
The reference imports after stubbing two absent packages, ``nerfacc`` and
``viser``; tinycudann is absent so the reference picks
its own pure-PyTorch encoders.  Stratified jitter is injected by patching
``torch.rand_like`` for the duration of one call (ray_rendering.py:286).
"""
import os
import sys
import types
import warnings

import numpy as np
import torch

warnings.filterwarnings("ignore")
REF = "/root/reference/adaptive_nerf"
HERE = os.path.dirname(os.path.abspath(__file__))


def import_reference():
    na = types.ModuleType("nerfacc")
    na.OccGridEstimator = object
    v = types.ModuleType("viser")
    vt = types.ModuleType("viser.transforms")
    v.transforms = vt
    sys.modules.update({"nerfacc": na, "viser": v, "viser.transforms": vt})
    if REF not in sys.path:
        sys.path.insert(0, REF)


def npd(d):
    return {k: (v.detach().cpu().numpy() if torch.is_tensor(v) else np.asarray(v)) for k, v in d.items()}


class inject_u:
    def __init__(self, u):
        self.u = u

    def __enter__(self):
        self.orig = torch.rand_like
        torch.rand_like = lambda x, **kw: self.u
        return self

    def __exit__(self, *a):
        torch.rand_like = self.orig


def main():
    import_reference()
    sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
    from oracle import acnerf_oracle as O  # scene helpers only (inputs), never outputs
    from nerfs.scene_box import SceneBox
    from nerfs.ray_sampling import get_ray_directions, get_rays, clamp_rays_near_far
    from nerfs.ray_rendering import render_rays, volume_render, stratified_t_vals
    from models.encodings import HashGridEncoder, SHEncoder, FrequencyEncoder
    from models.inr.meta_container import MetaContainer
    from models.inr.meta_ngp import MetaNGP
    from models.inr.meta_vanilla import MetaNeRF
    from nerfs.losses import compute_mse_loss

    g = torch.Generator().manual_seed(1234)
    aabb = torch.tensor(O.GLOBAL_AABB)

    # ---------------- rays (a1-a3) ----------------
    H, W, fx, fy, cx, cy = 12, 16, 20.0, 21.0, 7.7, 6.1
    c2w = O.ring_pose(3, 40)
    out = {"intr": np.array([H, W, fx, fy, cx, cy]), "c2w": c2w.numpy(), "aabb": aabb.numpy()}
    for cp in (True, False):
        dirs = get_ray_directions(H, W, fx, fy, cx, cy, center_pixels=cp, device="cpu")
        rays = get_rays(dirs, c2w, scene_box=SceneBox(aabb))
        out[f"dirs_cp{int(cp)}"] = dirs.numpy()
        out[f"rays_cp{int(cp)}"] = rays.numpy()
    flat = torch.from_numpy(out["rays_cp1"]).view(-1, 8)
    # add axis-parallel, zero-component and missing rays
    extra_o = torch.tensor([[0.2, 0.0, 0.0], [0.2, 0.0, 0.0], [5.0, 5.0, 5.0], [0.2, 0.3, -3.0], [0.1, 0.1, 0.1]])
    extra_d = torch.tensor([[1.0, 0.0, 0.0], [0.0, -0.0, 1.0], [1.0, 0.0, 0.0], [0.0, 0.0, 1.0], [1e-9, -1e-9, 1.0]])
    out["extra_o"], out["extra_d"] = extra_o.numpy(), extra_d.numpy()
    tn, tf = SceneBox(aabb).ray_aabb_intersect(extra_o, extra_d)
    out["extra_near"], out["extra_far"] = tn.numpy(), tf.numpy()
    r2, valid2 = clamp_rays_near_far(flat, near_far_override=(0.05, 1.2))
    r3, valid3 = clamp_rays_near_far(flat, near_far_override=None)
    r4, valid4 = clamp_rays_near_far(flat, near_far_override=(None, 0.4))
    out.update(clamp_a=r2.numpy(), valid_a=valid2.numpy(), valid_none=valid3.numpy(), clamp_b=r4.numpy(), valid_b=valid4.numpy())
    np.savez_compressed(os.path.join(HERE, "rays.npz"), **out)

    # ---------------- stratified t (a4) ----------------
    near = torch.rand(33, generator=g) * 0.5
    far = near + torch.rand(33, generator=g) * 2 + 0.01
    out = {"near": near.numpy(), "far": far.numpy()}
    for S in (2, 7, 64, 96):
        u = torch.rand(33, S, generator=g)
        with inject_u(u):
            tt = stratified_t_vals(near, far, S, randomized=True)
        te = stratified_t_vals(near, far, S, randomized=False)
        out[f"u{S}"], out[f"t_train{S}"], out[f"t_eval{S}"] = u.numpy(), tt.numpy(), te.numpy()
    np.savez_compressed(os.path.join(HERE, "t_vals.npz"), **out)

    # ---------------- routing (a6) on the example-like 4-expert layout ----------------
    _, cent4, boxes4 = O.grid_scene(2, 2)
    pts = torch.rand(4096, 3, generator=g) * (aabb[1] - aabb[0]) + aabb[0]
    # include points exactly on / next to the bisectors
    pts[:64, 1] = 0.0
    pts[64:128, 2] = 0.0
    pts[128:160, 1:] = 0.0
    out = {"pts": pts.numpy(), "centroids": cent4.numpy()}
    for tag, margin, c2d in (("soft2d", 1.05, True), ("hard2d", 1.0, True), ("soft3d", 1.2, False), ("hard3d", 1.0, False)):
        m = MetaContainer.__new__(MetaContainer)
        torch.nn.Module.__init__(m)
        m.register_buffer("centroids", cent4)
        m.boundary_margin, m.cluster_2d = margin, c2d
        m._coord_idx = (1, 2) if c2d else (0, 1, 2)
        wts, hard = m._routing(pts)
        out[tag] = (wts if wts is not None else hard).numpy()
    np.savez_compressed(os.path.join(HERE, "routing.npz"), **out)

    # ---------------- encodings (a9-a11) ----------------
    x01 = torch.rand(300, 3, generator=g)
    x01[:4] = torch.tensor([[1e-6, 1e-6, 1e-6], [1 - 1e-6] * 3, [0.5, 0.5, 0.5], [0.25, 1 - 1e-6, 1e-6]])
    out = {"x01": x01.numpy()}
    for tag, log2T, interp in (("lin8", 8, "Linear"), ("smooth8", 8, "Smoothstep"), ("near8", 8, "Nearest")):
        torch.manual_seed(7)
        enc = HashGridEncoder(levels=16, min_res=16, max_res=4096, log2_hashmap_size=log2T, features_per_level=2,
                              implementation="torch", interpolation=interp)
        with torch.no_grad():
            enc.hash_table.normal_(0, 0.5)
        xin = x01.clone()
        y = enc(xin)
        gy = torch.randn(y.shape, generator=g)
        (y * gy).sum().backward()
        out[f"{tag}_table"], out[f"{tag}_y"] = enc.hash_table.detach().numpy(), y.detach().numpy()
        out[f"{tag}_gy"], out[f"{tag}_gtable"] = gy.numpy(), enc.hash_table.grad.numpy()
    # full-size index check (T=2^20): floor/ceil corner rows incl. level offset
    enc = HashGridEncoder(levels=16, min_res=16, max_res=4096, log2_hashmap_size=20, features_per_level=2, implementation="torch")
    out["res"] = enc.level_resolutions.numpy()
    sc = x01[:, None, :] * enc.level_resolutions.float().view(1, 16, 1)
    fl = torch.floor(sc).to(torch.int64)
    ce = fl + 1
    idx = []
    for bx in (0, 1):
        for by in (0, 1):
            for bz in (0, 1):
                a = (ce if bx else fl)[..., 0]
                b = (ce if by else fl)[..., 1]
                c = (ce if bz else fl)[..., 2]
                idx.append(enc._hash(a, b, c) + enc.level_offsets)
    out["idx20"] = torch.stack(idx, -1).to(torch.int32).numpy()
    d = torch.randn(200, 3, generator=g)
    d[0] = torch.tensor([0.0, 0.0, 0.0])
    out["dirs"] = d.numpy()
    out["sh16"] = SHEncoder(levels=4, implementation="torch")(d).numpy()
    out["freq10"] = FrequencyEncoder(3, 10)(d).numpy()
    out["freq4"] = FrequencyEncoder(3, 4)(d).numpy()
    np.savez_compressed(os.path.join(HERE, "encodings.npz"), **out)

    # ---------------- compositing (a17) ----------------
    N, S = 37, 24
    rs = torch.rand(N, S, 4, generator=g) * 1.4 - 0.2
    rs[..., 3] = torch.randn(N, S, generator=g).exp() * 3 - 0.5
    rs[0, :, 3] = 0.0
    rs[1, :, 3] = 1e4
    t = torch.sort(torch.rand(N, S, generator=g) * 2, dim=1).values
    t[2, 5] = t[2, 4]  # zero-length interval -> 1e-4 floor
    bg = torch.rand(N, 3, generator=g)
    rs.requires_grad_(True)
    bg.requires_grad_(True)
    rgb, depth, w, acc = volume_render(rs, t, bg_rgb=bg, sigma_scale=1.0)
    gr, gd, ga = torch.randn(N, 3, generator=g), torch.randn(N, generator=g), torch.randn(N, generator=g)
    ((rgb * gr).sum() + (depth * gd).sum() + (acc * ga).sum()).backward()
    out = npd(dict(rgb_sigma=rs, t=t, bg=bg, rgb=rgb, depth=depth, weights=w, acc=acc, g_rgb=gr, g_depth=gd, g_acc=ga,
                   d_rgb_sigma=rs.grad, d_bg=bg.grad))
    rgb2, depth2, w2, acc2 = volume_render(rs.detach(), t, bg_rgb=None, sigma_scale=2.5)
    out.update(npd(dict(rgb_s25=rgb2, depth_s25=depth2, acc_s25=acc2)))
    np.savez_compressed(os.path.join(HERE, "composite.npz"), **out)

    # ---------------- expert + container + render_rays (a7-a8, a12-a14, a16, whole path) ----------------
    cfg = dict(O.HASH_CFG)
    cfg["log2_hashmap_size"] = 9
    _, cent2, boxes2 = O.grid_scene(1, 2)
    torch.manual_seed(0)
    model = MetaContainer(2, cent2, aabb, "instant", boundary_margin=1.05, cluster_2d=True, use_bg_nerf=True,
                          occ_conf={"use_occ": False}, expert_box_list=[SceneBox(b) for b in boxes2], hidden=64,
                          sigma_depth=2, color_depth=2, color_hidden=64, dir_encoding="spherical",
                          use_sigmoid_rgb=True, hash_enc_conf=cfg)
    with torch.no_grad():
        for sub in model.submodules:
            sub.xyz_encoder.hash_table.normal_(0, 0.3)
            sub.sigma_head.bias.fill_(0.5)
    out = {"state/" + k: v.detach().numpy() for k, v in model.state_dict().items()}
    out["box_min"] = torch.stack([b[0] for b in boxes2]).numpy()
    out["box_max"] = torch.stack([b[1] for b in boxes2]).numpy()
    out["log2_T"] = np.array(9)
    Nr, S = 96, 12
    rays = O.aerial_rays(Nr, aabb, seed=5)
    u = torch.rand(Nr, S, generator=g)
    gt = torch.rand(Nr, 3, generator=g)
    out.update(rays=rays.numpy(), u=u.numpy(), gt=gt.numpy())
    # expert 0 alone on arbitrary points
    x6 = torch.cat([torch.rand(200, 3, generator=g) * (aabb[1] - aabb[0]) + aabb[0], torch.randn(200, 3, generator=g)], -1)
    out["x6"] = x6.numpy()
    out["expert0_y"] = model.submodules[0](x6).detach().numpy()
    out["container_y"] = model(x6).detach().numpy()
    out["bg_rgb"] = model.background_color(rays[:, 3:6]).detach().numpy()

    class P:
        ray_samples, chunk_points, color_space, fim = S, 1 << 20, "linear", False

    for tag, active in (("full", None), ("act1", 1)):
        model.train()
        model.zero_grad()
        with inject_u(u):
            rgb, depth, w, acc = render_rays(model, rays, ray_samples=S, active_module=active)
        with inject_u(u):
            loss = compute_mse_loss(P, model, {"rays": rays, "rgbs": gt}, active_module=active)
        loss.backward()
        out.update(npd({f"{tag}_rgb": rgb, f"{tag}_depth": depth, f"{tag}_w": w, f"{tag}_acc": acc, f"{tag}_loss": loss}))
        for n, p in model.named_parameters():
            if p.grad is not None:
                out[f"{tag}_grad/{n}"] = p.grad.numpy().copy()
        model.eval()
        with torch.no_grad():
            rgb, depth, w, acc = render_rays(model, rays, ray_samples=S, active_module=active, bg_color_default="black")
        out.update(npd({f"{tag}_eval_rgb": rgb, f"{tag}_eval_depth": depth, f"{tag}_eval_acc": acc}))
    model.use_bg_nerf = False
    with torch.no_grad():
        rgb, *_ = render_rays(model, rays, ray_samples=S, bg_color_default="black")
    out["full_eval_rgb_blackbg"] = rgb.numpy()
    np.savez_compressed(os.path.join(HERE, "ngp_path.npz"), **out)

    # ---------------- PE-MLP expert (a15, config 1) ----------------
    torch.manual_seed(3)
    pe = MetaNeRF(hidden=64, sigma_depth=8, color_hidden=32, encoding_dir="frequency")
    out = {"state/" + k: v.detach().numpy() for k, v in pe.state_dict().items()}
    x = torch.rand(150, 3, generator=g) * 2 - 1
    dd = torch.nn.functional.normalize(torch.randn(150, 3, generator=g), dim=-1)
    r = pe(x, dd)
    gs, gc = torch.randn(150, 1, generator=g), torch.randn(150, 3, generator=g)
    ((r["sigma"] * gs).sum() + (r["rgb"] * gc).sum()).backward()
    out.update(npd(dict(x=x, d=dd, sigma=r["sigma"], rgb=r["rgb"], g_sigma=gs, g_rgb=gc)))
    for n, p in pe.named_parameters():
        out[f"grad/{n}"] = p.grad.numpy().copy()
    np.savez_compressed(os.path.join(HERE, "pe_mlp.npz"), **out)
    for f in sorted(os.listdir(HERE)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(HERE, f)))


if __name__ == "__main__":
    main()
