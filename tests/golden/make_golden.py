"""Generate the golden fixtures from the UNMODIFIED reference.

Run in the build container only (needs /root/reference):

    python tests/golden/make_golden.py

It imports the reference's own ``get_points``, ``backproject``,
``get_extrinsics`` and ``ImVoxelNet._compute_projection`` through
``oracle/ref_loader.py`` (nothing of the reference is copied), feeds them
seeded synthetic inputs and stores inputs + outputs as small ``.npz`` files
beside this script.  ``tests/test_oracle_golden.py`` pins the C oracle to
these files; ``tests/test_backproject_gpu.py`` pins the CUDA kernels to them.

The scene reduction of ``extract_feat`` (detectors/imvoxelnet.py:70-74) is a
method body, not a function, so it is driven here with the very same five
torch statements applied to the reference ``backproject`` output.
"""
import hashlib
import os
import sys
import warnings

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.abspath(os.path.join(HERE, "..", "..")))

from imvoxelnet_b200 import scenes  # noqa: E402
from oracle import ref_loader  # noqa: E402

warnings.filterwarnings("ignore")
torch.set_num_threads(8)
det, _ = ref_loader.load()


def reference_scene_mean(features, points, projection):
    volume, valid = det.backproject(features, points, projection)
    per_view_volume, per_view_valid = volume, valid
    volume = volume.sum(dim=0)
    valid = valid.sum(dim=0)
    volume = volume / valid
    valid = valid > 0
    volume[:, ~valid[0]] = .0
    return per_view_volume, per_view_valid, volume, valid


def reference_indices(features, points, projection):
    """x, y, valid exactly as backproject computes them (:148-154); they are
    locals there, so the same statements are evaluated on the same inputs."""
    n_images = features.shape[0]
    height, width = features.shape[-2:]
    p = points.view(1, 3, -1).expand(n_images, 3, -1)
    p = torch.cat((p, torch.ones_like(p[:, :1])), dim=1)
    p3 = torch.bmm(projection, p)
    x = (p3[:, 0] / p3[:, 2]).round().long()
    y = (p3[:, 1] / p3[:, 2]).round().long()
    valid = (x >= 0) & (y >= 0) & (x < width) & (y < height) & (p3[:, 2] > 0)
    return x, y, valid


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def scene_inputs(scene, seed, channels=None, channels_last=False):
    feats = scenes.random_features(scene, seed=seed, channels_last=channels_last, channels=channels)
    h, w = scene.feat_hw
    origin = torch.tensor(scene.img_meta['lidar2img']['origin'])
    points = det.get_points(torch.tensor(scene.n_voxels), torch.tensor(scene.voxel_size), origin)
    projection = det.ImVoxelNet._compute_projection(scene.img_meta, scenes.STRIDE, None)
    return feats, feats[:, :, :h, :w], points, projection


def save_small(name, features_full, hw, points, projection, note):
    h, w = hw
    feats = features_full[:, :, :h, :w]
    pv, pval, vol, val = reference_scene_mean(feats, points, projection)
    x, y, v2 = reference_indices(feats, points, projection)
    assert torch.equal(v2.view_as(pval), pval)
    np.savez(os.path.join(HERE, name + ".npz"),
             features=features_full.numpy(), h=h, w=w, points=points.numpy(),
             projection=projection.numpy(), x=x.numpy(), y=y.numpy(),
             valid_views=pval.numpy(), volume_views=pv.numpy(),
             volume=vol.numpy(), valid=val.numpy(), note=note)
    print(f"{name}: V={feats.shape[0]} C={feats.shape[1]} map={tuple(features_full.shape[-2:])} "
          f"hw=({h},{w}) N={points[0].numel()} valid={pval.float().mean():.3f} "
          f"covered={val.float().mean():.3f}")


def small_scene(builder, img_shape, ori_shape, n_voxels, voxel_size, channels, map_hw, seed, **kw):
    s = builder(**kw)
    s.img_meta['img_shape'] = img_shape
    s.img_meta['ori_shape'] = ori_shape
    s.n_voxels, s.voxel_size, s.channels, s.map_hw = n_voxels, voxel_size, channels, map_hw
    full, _, points, projection = scene_inputs(s, seed)
    return s, full, points, projection


def main():
    # 1. multi-view ring, image smaller than the padded map (strided slice, :69)
    s, full, pts, P = small_scene(scenes.scannet, (120, 156, 3), (480, 624, 3), (12, 10, 6),
                                  (.5, .55, .45), 8, (32, 40), seed=11, n_views=5)
    save_small("bp_ring5", full, s.feat_hw, pts, P, "5-view ring, non-cubic voxels, cropped map")

    # 2. KITTI-like forward camera, V=1
    s, full, pts, P = small_scene(scenes.kitti, (96, 320, 3), (375, 1242, 3), (27, 31, 3),
                                  (2.56, 2.56, 1.28), 4, (24, 80), seed=12)
    save_small("bp_kitti_small", full, s.feat_hw, pts, P, "forward camera, V=1")

    # 3. six-camera rig, intrinsic = identity, lidar2img in the extrinsic
    s, full, pts, P = small_scene(scenes.nuscenes, (90, 160, 3), (900, 1600, 3), (20, 20, 4),
                                  (5., 5., 2.), 4, (24, 40), seed=13)
    save_small("bp_nus6", full, s.feat_hw, pts, P, "6 cameras, K=I")

    # 4. hand-made ties / borders / z<=0.  P0 maps (x,y,z)->(x, y, 1): pixel = point
    #    coordinate, so .5 coordinates are exact ties; P1 uses z as depth with
    #    zeros and negatives; P2 is a scaled copy giving inf/NaN lanes at z == 0.
    w, h = 12, 9
    xs = torch.tensor([-1.5, -0.5, 0.5, 1.5, 2.5, 3.0, 10.5, 11.0, 11.5, 12.5, -0.49999997, 11.499999])
    ys = torch.tensor([-0.5, 0.5, 7.5, 8.0, 8.5, 8.499999, 3.5])
    zs = torch.tensor([-1.0, 0.0, 1.0, 2.0, -0.0])
    gx, gy, gz = torch.meshgrid(xs, ys, zs, indexing='ij')
    pts = torch.stack([gx, gy, gz]).contiguous()
    P = torch.tensor([
        [[1., 0, 0, 0], [0, 1., 0, 0], [0, 0, 0, 1.]],
        [[1., 0, 0, 0], [0, 1., 0, 0], [0, 0, 1., 0]],
        [[2., 0, 0, 0], [0, 2., 0, 0], [0, 0, 2., 0]],
        [[0., 0, 0, 3.], [0, 0, 0, 4.], [0, 0, 1., 0]],
    ])
    g = torch.Generator().manual_seed(14)
    full = torch.randn(4, 6, h + 3, w + 4, generator=g)
    save_small("bp_ties", full, (h, w), pts, P, "round-half-even ties, borders, z<=0, inf/NaN lanes")

    # 5. host-side set-up: projections, extrinsics-from-angles, grid corners
    out = {}
    for name in scenes.BUILDERS:
        sc = scenes.build(name)
        out[f"P_{name}"] = det.ImVoxelNet._compute_projection(sc.img_meta, 4, None).numpy()
        p = det.get_points(torch.tensor(sc.n_voxels), torch.tensor(sc.voxel_size),
                           torch.tensor(sc.img_meta['lidar2img']['origin'])).numpy()
        out[f"points_sha_{name}"] = sha(p)
        out[f"points_first_{name}"] = p[:, 0, 0, 0]
        out[f"points_last_{name}"] = p[:, -1, -1, -1]
    angles = torch.tensor([[0.13, -0.07], [0., 0.], [-0.4, 0.25]])
    out["angles"] = angles.numpy()
    out["extrinsics"] = torch.stack([det.get_extrinsics(a) for a in angles]).numpy()
    sc = scenes.build("total_sunrgbd")
    out["P_total_angles"] = det.ImVoxelNet._compute_projection(sc.img_meta, 4, angles[:1]).numpy()
    np.savez(os.path.join(HERE, "host_setup.npz"), **out)

    make_digests(FULLSIZE)


FULLSIZE = ("sunrgbd_fast", "kitti", "scannet_fast", "nuscenes", "total_sunrgbd", "nuscenes_refgrid")


def make_digests(names):
    """6. full-size configurations: digests of the reference outputs.  Entries of configurations not
    in `names` are kept as they are in the file (`--digests NAME ...` adds one without redoing all)."""
    path = os.path.join(HERE, "fullsize_digests.npz")
    dig = {}
    if os.path.isfile(path):
        with np.load(path) as old:
            dig = {k: old[k] for k in old.files}
    for name in names:
        sc = scenes.build(name)
        full, feats, pts, P = scene_inputs(sc, seed=100)
        x, y, valid = reference_indices(feats, pts, P)
        _, _, vol, val = reference_scene_mean(feats, pts, P)
        x, y, valid = x.numpy(), y.numpy(), valid.numpy()
        dig[f"{name}_valid_sha"] = sha(valid)
        dig[f"{name}_x_sha"] = sha(np.where(valid, x, 0))
        dig[f"{name}_y_sha"] = sha(np.where(valid, y, 0))
        dig[f"{name}_n_valid"] = int(valid.sum())
        dig[f"{name}_covered"] = int(val.sum())
        v = vol.numpy().reshape(vol.shape[0], -1)
        dig[f"{name}_chan_sum"] = v.astype(np.float64).sum(1)
        dig[f"{name}_chan_abs_sum"] = np.abs(v.astype(np.float64)).sum(1)
        rs = np.random.default_rng(5).integers(0, v.size, 4096)
        dig[f"{name}_sample_idx"] = rs
        dig[f"{name}_sample_val"] = v.reshape(-1)[rs]
        dig[f"{name}_features_sha"] = sha(full.numpy())
        print(name, "n_valid", dig[f"{name}_n_valid"], "covered", dig[f"{name}_covered"])
    np.savez(path, **dig)


if __name__ == "__main__" and "--digests" in sys.argv:
    make_digests(sys.argv[sys.argv.index("--digests") + 1:])
elif __name__ == "__main__" and not {"--necks", "--heads"} & set(sys.argv):
    main()


# --------------------------------------------------------------------------
# 3D necks: state dict + input + outputs of the reference modules (small sizes)
# --------------------------------------------------------------------------
from tests.golden.cases import NECK_CASES  # noqa: E402


def make_necks():
    from oracle import torch_necks
    _, necks = ref_loader.load()
    for i, (name, (kw, shape)) in enumerate(NECK_CASES.items()):
        torch.manual_seed(100 + i)
        module = getattr(necks, name)(**kw).eval()
        torch_necks.randomize_norms(module, seed=200 + i)      # non-trivial BN, non-zero bn2.weight
        x = torch.randn(shape)
        with torch.no_grad():
            outs = module(x)
        blob = {f"sd/{k}": v.numpy() for k, v in module.state_dict().items()}
        blob["x"] = x.numpy()
        for j, o in enumerate(outs):
            blob[f"out{j}"] = o.contiguous().numpy()
        np.savez(os.path.join(HERE, f"neck_{name}.npz"), **blob)
        print(name, [tuple(o.shape) for o in outs], sum(v.size for k, v in blob.items() if k.startswith("sd/")), "params")


if __name__ == "__main__" and "--necks" in sys.argv:
    make_necks()


# --------------------------------------------------------------------------
# dense heads: the reference classes' forward(), the valid-mask resampling and
# the pre-NMS scores (imvoxel_head_v2.py:91-93, 262-270), small sizes
# --------------------------------------------------------------------------
from tests.golden.cases import HEAD_CASES, MASK_CASES  # noqa: E402


def make_heads():
    from oracle import torch_heads
    v1, v2 = ref_loader.load_heads()
    for i, (name, (kw, _, shapes)) in enumerate(HEAD_CASES.items()):
        torch.manual_seed(300 + i)
        head = getattr(v2 if name.endswith("V2") else v1, name)(**kw).eval()
        torch_heads.randomise(head, seed=400 + i)
        xs = [torch.randn(s) for s in shapes]
        with torch.no_grad():
            centernesses, bbox_preds, cls_scores = head(xs)
        blob = {f"sd/{k}": v.numpy() for k, v in head.state_dict().items()}
        for j, x in enumerate(xs):
            blob[f"x{j}"] = x.numpy()
            blob[f"centerness{j}"] = centernesses[j].numpy()
            blob[f"bbox{j}"] = bbox_preds[j].numpy()
            blob[f"cls{j}"] = cls_scores[j].numpy()
        # scores of scene 0, level 0, exactly the statements of _get_bboxes_single (:262-270)
        valid = torch.rand(shapes[0][2:], generator=torch.Generator().manual_seed(i))[None] > 0.3
        n_classes = kw["n_classes"]
        centerness = centernesses[0][0].permute(1, 2, 3, 0).reshape(-1).sigmoid()
        scores = cls_scores[0][0].permute(1, 2, 3, 0).reshape(-1, n_classes).sigmoid()
        scores = scores * centerness[:, None] * valid.permute(1, 2, 3, 0).reshape(-1)[:, None]
        max_scores, _ = scores.max(dim=1)
        blob["score_valid"], blob["scores"], blob["max_scores"] = valid.numpy(), scores.numpy(), max_scores.numpy()
        np.savez_compressed(os.path.join(HERE, f"head_{name}.npz"), **blob)
        print(name, [tuple(o.shape) for o in bbox_preds])
    for i, (name, (shape, sizes)) in enumerate(MASK_CASES.items()):
        valid = (torch.rand(shape, generator=torch.Generator().manual_seed(500 + i)) > 0.45).float()
        blob = {"valid": valid.numpy().astype(np.bool_)}
        for j, size in enumerate(sizes):   # the reference's statement, imvoxel_head_v2.py:93
            blob[f"out{j}"] = torch.nn.Upsample(size=size, mode='trilinear')(valid).round().bool().numpy()
        np.savez_compressed(os.path.join(HERE, f"mask_{name}.npz"), **blob)
        print(name, [int(blob[f"out{j}"].sum()) for j in range(len(sizes))])


if __name__ == "__main__" and "--heads" in sys.argv:
    make_heads()
