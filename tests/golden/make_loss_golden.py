"""Golden vectors for the fused loss from the reference's OWN `get_loss`.

Run in the authoring container (needs /root/reference; CPU only):  python tests/golden/make_loss_golden.py

`slam/Loss.py` is imported UNMODIFIED from the reference checkout.  Its package imports drag in dependencies that are
not installed here and that get_loss never touches (pytorch3d, plyfile, simple_knn, the NetVLAD descriptor): they are
replaced by empty stub modules before the import.  get_loss itself allocates its accumulator with
`torch.tensor(0., device='cuda')` (slam/Loss.py:60); the module's view of `torch` is wrapped so that this one call
lands on the CPU.  Nothing else is altered: masks, means / sums and weights are the reference's code.

What the reference does BEFORE get_loss -- the weight normalisation of the rendered depth in Renderer_mapping /
Renderer_tracking (render/__init__.py:46-49) -- cannot be imported (those functions call the CUDA rasterizer), so these
three lines are restated here, verbatim, on the synthetic rasterizer outputs.

Output: tests/golden/loss_<case>.npz with in_* (color [3,H,W], allmap [7,H,W], gt_color [H,W,3], gt_depth [H,W,1]) and,
for mode in (mapping, tracking): ref_loss_<mode>, ref_dcolor_<mode>, ref_dallmap_<mode> (autograd through the restated
weight normalisation and the reference's get_loss).
"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REFERENCE = "/root/reference"

CONFIG = dict(
    loss=dict(use_normal_loss=False, ignore_outliners=False, silmask_th=0.90,          # configs/replica/config.py:92-106
              tracking=dict(color=0.5, depth=1.0), mapping=dict(color=0.5, depth=1.0, dist=0.1)),
    densify=dict(use_edge_growth=False, edge_thres=0.5),
    render=dict(use_weight_norm=True, eps=1e-6, depth_near=1e-2, depth_far=1e2))


def import_reference_loss():
    def stub(name, **attrs):
        m = types.ModuleType(name)
        sys.modules[name] = m
        for k, v in attrs.items():
            setattr(m, k, v)
        return m
    stub("pytorch3d")
    sys.modules["pytorch3d"].transforms = stub("pytorch3d.transforms")
    stub("plyfile", PlyData=object, PlyElement=object)
    stub("simple_knn")
    stub("simple_knn._C", distCUDA2=None)
    stub("utils.descriptor", MyDesc=object)
    sys.path.insert(0, REFERENCE)
    from slam import Loss

    class TorchOnCpu:
        """The module's `torch`, with tensor(..., device='cuda') redirected to the CPU (slam/Loss.py:60)."""
        def __getattr__(self, name):
            return getattr(torch, name)

        @staticmethod
        def tensor(*a, **kw):
            kw.pop("device", None)
            return torch.tensor(*a, **kw)
    Loss.torch = TorchOnCpu()
    return Loss.get_loss


def inputs(H, W, seed, nasty):
    rng = np.random.default_rng(seed)
    color = rng.uniform(0, 1, (3, H, W)).astype(np.float32)
    allmap = rng.normal(0, 1, (7, H, W)).astype(np.float32)
    alpha = rng.uniform(0.0, 1.0, (H, W)).astype(np.float32)
    alpha[rng.uniform(size=(H, W)) < 0.5] = 0.95                       # half the pixels pass the silhouette mask
    depth = rng.uniform(0.5, 5.0, (H, W)).astype(np.float32)
    allmap[1] = alpha
    allmap[0] = depth * alpha
    allmap[6] = rng.uniform(0, 0.05, (H, W)).astype(np.float32)
    gt_color = rng.uniform(0, 1, (H, W, 3)).astype(np.float32)
    gt_depth = rng.uniform(0.5, 5.0, (H, W, 1)).astype(np.float32)
    if nasty:
        gt_depth[rng.uniform(size=(H, W)) < 0.1] = 0.0                 # holes in the sensor depth
        allmap[1][rng.uniform(size=(H, W)) < 0.05] = 0.0               # nothing rendered: depth = D / eps -> outlier
        color[:, rng.uniform(size=(H, W)) < 0.02] = np.nan
        color[0][rng.uniform(size=(H, W)) < 0.01] = np.inf
        allmap[6][rng.uniform(size=(H, W)) < 0.02] = np.inf
        allmap[0][0, :4] = 1e6                                         # beyond depth_far
        color[1][1, :4] = gt_color[1, :4, 1]                           # exact zeros of |c - gt|
    return color, allmap, gt_color, gt_depth


def outlier_inputs(H, W, seed):
    """Rendered depth close to the sensor depth except for 6 % gross outliers, 8 % sensor holes: the median of the
    masked depth error is small, so `depth_error < 10 * median` (slam/Loss.py:37-40) removes the outliers."""
    rng = np.random.default_rng(seed)
    color, allmap, gt_color, gt_depth = inputs(H, W, seed, False)
    alpha = allmap[1]
    depth = gt_depth[..., 0] + rng.normal(0, 0.02, (H, W)).astype(np.float32)
    gross = rng.uniform(size=(H, W)) < 0.06
    depth[gross] += rng.uniform(0.5, 2.0, int(gross.sum())).astype(np.float32)
    allmap[0] = depth * alpha
    gt_depth[rng.uniform(size=(H, W)) < 0.08] = 0.0
    return color, allmap, gt_color, gt_depth


def reference_loss(get_loss, color, allmap, gt_color, gt_depth, tracking, config=None):
    global CONFIG
    saved = CONFIG
    if config is not None:
        CONFIG = config
    try:
        return _reference_loss(get_loss, color, allmap, gt_color, gt_depth, tracking)
    finally:
        CONFIG = saved


def _reference_loss(get_loss, color, allmap, gt_color, gt_depth, tracking):
    c = torch.tensor(color, requires_grad=True)
    a = torch.tensor(allmap, requires_grad=True)
    r = CONFIG["render"]
    render_pkg = dict(render_color=c, render_alpha=a[1:2], render_depth=a[0:1], render_normal=a[2:5], render_dist=a[6:7])
    # render/__init__.py:46-49, verbatim
    render_pkg['render_depth'] = render_pkg['render_depth'] / (render_pkg['render_alpha'] + r['eps'])
    outliner_mask = torch.logical_or(render_pkg['render_depth'] > r['depth_far'], render_pkg['render_depth'] < r['depth_near'])
    render_pkg['render_depth'][outliner_mask] = 0
    camera = types.SimpleNamespace(gt_color=torch.tensor(gt_color), gt_depth=torch.tensor(gt_depth))
    loss = get_loss(CONFIG, render_pkg, camera, tracking=tracking)
    loss.backward()
    return np.float32(loss.detach().numpy()), c.grad.numpy(), a.grad.numpy()


CASES = dict(small=(24, 40, 0, True), ragged=(17, 33, 1, True), clean=(32, 48, 2, False))
# densify.use_edge_growth = True, edge_thres = 0.4 (the values of configs/*/config.py when edge growth is switched on):
# colour and distortion are averaged over alpha > 0.4, depth over the depth mask (slam/Loss.py:53)
EDGE_CASES = dict(edge=(24, 40, 5, True), edge_ragged=(19, 35, 6, True))


def main():
    get_loss = import_reference_loss()
    for name, (H, W, seed, nasty) in CASES.items():
        x = inputs(H, W, seed, nasty)
        out = dict(in_color=x[0], in_allmap=x[1], in_gt_color=x[2], in_gt_depth=x[3])
        for mode, tracking in (("mapping", False), ("tracking", True)):
            loss, dc, da = reference_loss(get_loss, *x, tracking=tracking)
            out[f"ref_loss_{mode}"], out[f"ref_dcolor_{mode}"], out[f"ref_dallmap_{mode}"] = loss, dc, da
            print(name, mode, float(loss), int(np.isnan(da).sum()), "NaN in dL/dallmap")
        np.savez_compressed(os.path.join(HERE, f"loss_{name}.npz"), **out)
    import copy
    cfg = copy.deepcopy(CONFIG)
    cfg["densify"] = dict(use_edge_growth=True, edge_thres=0.4)
    for name, (H, W, seed, nasty) in EDGE_CASES.items():
        x = inputs(H, W, seed, nasty)
        out = dict(in_color=x[0], in_allmap=x[1], in_gt_color=x[2], in_gt_depth=x[3], edge_thres=np.float32(0.4))
        loss, dc, da = reference_loss(get_loss, *x, tracking=False, config=cfg)
        out["ref_loss_mapping"], out["ref_dcolor_mapping"], out["ref_dallmap_mapping"] = loss, dc, da
        print(name, "mapping + edge growth", float(loss), int(np.isnan(da).sum()), "NaN in dL/dallmap")
        np.savez_compressed(os.path.join(HERE, f"loss_{name}.npz"), **out)
    # the median-based outlier mask of the tracking loss, as configs/replica/splatam.py switches it on
    # (ignore_outliners = True, silmask_th = 0.99)
    import copy
    cfg = copy.deepcopy(CONFIG)
    cfg["loss"]["ignore_outliners"] = True
    cfg["loss"]["silmask_th"] = 0.99
    x = outlier_inputs(28, 44, 3)
    x[1][1][x[1][1] == np.float32(0.95)] = 0.995          # the pixels meant to pass the silhouette mask pass 0.99 too
    x[1][0][...] = (x[1][0] / 0.95 * 0.995).astype(np.float32) * (x[1][1] == np.float32(0.995)) + x[1][0] * (x[1][1] != np.float32(0.995))
    out = dict(in_color=x[0], in_allmap=x[1], in_gt_color=x[2], in_gt_depth=x[3], silmask_th=np.float32(0.99))
    loss, dc, da = reference_loss(get_loss, *x, tracking=True, config=cfg)
    out["ref_loss_tracking_inliers"], out["ref_dcolor_tracking_inliers"], out["ref_dallmap_tracking_inliers"] = loss, dc, da
    plain, _, _ = reference_loss(get_loss, *x, tracking=True, config={**cfg, "loss": {**cfg["loss"], "ignore_outliners": False}})
    out["ref_loss_tracking_all"] = plain
    print("outliers tracking", float(loss), "without the outlier mask", float(plain))
    np.savez_compressed(os.path.join(HERE, "loss_outliers.npz"), **out)


if __name__ == "__main__":
    main()
