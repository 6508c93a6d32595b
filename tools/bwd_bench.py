"""mlp_bwd device time on the bench's 8-expert step (8192 rays x 96 samples), one vs two tiles in flight per CTA
(ACN_BWD_TILES=1|2 in the environment: each setting runs in its own process), plus the parameter / gradient agreement of
one step between the two settings.   python tools/bwd_bench.py            (GPU)

ACN_BWD_TILES / ACN_BWD2_SKIP are read by the EXPERIMENTAL kernel of profiles/r2_bwd_two_tiles.patch only (measured, not kept:
profiles/r2_bwd_two_tiles.txt); on the product library every setting runs the one-tile kernel."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def child():
    import torch
    sys.path.insert(0, ROOT)
    import bench
    from nerf_sys_b200 import ops
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    model, aabb = bench.build_container(dev)
    model.train()
    model.set_deferred_routing(True)
    from nerf_sys_b200.nerfs.losses import compute_mse_loss
    from nerf_sys_b200 import synthetic as SY
    rays = SY.aerial_rays(bench.N_RAYS, aabb, seed=5).to(dev)
    rgbs = torch.rand(bench.N_RAYS, 3, generator=torch.Generator().manual_seed(1)).to(dev)
    params = [p for p in model.parameters() if p.requires_grad]

    def step():
        for p in params:
            p.grad = None
        loss = compute_mse_loss(bench.P, model, {"rays": rays, "rgbs": rgbs})
        loss.backward()
        return loss
    for _ in range(3):
        step()
    timing = ops.kernel_timing(True)
    for _ in range(10):
        loss = step()
    torch.cuda.synchronize()
    ops.kernel_timing(False)
    for k in ("mlp_bwd", "mlp_fwd", "hash_bwd"):
        ts = [a.elapsed_time(b) for a, b in timing.get(k, [])]
        if ts:
            print(f"{k:9s} {1000 * sum(ts) / len(ts):8.1f} us per call ({len(ts)} calls)")
    # gradients of the MLP weights and a table, for the cross-setting comparison
    out = {n: p.grad.detach().float().cpu() for n, p in model.named_parameters() if p.grad is not None and ("submodules.3." in n or "submodules.0.sigma" in n)}
    out["loss"] = loss.detach().cpu()
    torch.save(out, sys.argv[2])


def main():
    outs = {}
    # (tiles in flight, timing-only probe: 1 = no bias MMAs, 2 = no weight-gradient MMAs at all -> wrong gradients, issue-rate probe)
    for tiles, skip in ((1, 0), (2, 0)):
        f = f"/tmp/bwd_bench_{tiles}_{skip}.pt"
        env = dict(os.environ, ACN_BWD_TILES=str(tiles), ACN_BWD2_SKIP=str(skip))
        print(f"--- ACN_BWD_TILES={tiles} ACN_BWD2_SKIP={skip}", flush=True)
        r = subprocess.run([sys.executable, __file__, "--child", f], env=env, timeout=300)
        if r.returncode != 0:
            print("child failed", r.returncode)
            return 1
        outs[(tiles, skip)] = f
    import torch
    a, b = torch.load(outs[(1, 0)]), torch.load(outs[(2, 0)])
    for k in a:
        d = (a[k].double() - b[k].double()).norm() / max(a[k].double().norm().item(), 1e-30)
        print(f"  {k:60s} |one| {a[k].double().norm():.3e}  rel-L2(one, two) {float(d):.2e}")
    return 0


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "--child":
        child()
    else:
        sys.exit(main())
