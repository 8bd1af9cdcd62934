"""The drop-in detector: builds from the reference's ``model`` dicts (config
contract) and its ``extract_feat`` equals the reference's loop
(torch port of backproject + view mean, oracle neck) on the same 2D features."""
import glob
import os
import warnings

import numpy as np
import pytest
import torch

from imvoxelnet_b200 import detector, necks, scenes
from imvoxelnet_b200.registry import DETECTORS, NECKS, build_from
from oracle import ref_loader
from tests.util import assert_volume_close

KITTI_MODEL = dict(
    type='ImVoxelNet', pretrained='torchvision://resnet50',
    backbone=dict(type='ResNet', depth=50, num_stages=4, out_indices=(0, 1, 2, 3), frozen_stages=1,
                  norm_cfg=dict(type='BN', requires_grad=False), norm_eval=True, style='pytorch'),
    neck=dict(type='FPN', in_channels=[256, 512, 1024, 2048], out_channels=64, num_outs=4),
    neck_3d=dict(type='KittiImVoxelNeck', in_channels=64, out_channels=256),
    bbox_head=dict(type='Anchor3DHead', num_classes=1, in_channels=256, feat_channels=256),
    n_voxels=(216, 248, 12), voxel_size=(.32, .32, .32))


def _build(model_cfg, **kw):
    cfg = dict(model_cfg)
    cfg.update(kw)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        return build_from(DETECTORS, cfg)


def test_registry_names_match_the_reference():
    for name in ("FastIndoorImVoxelNeck", "ImVoxelNeck", "KittiImVoxelNeck", "NuScenesImVoxelNeck"):
        assert NECKS.get(name) is getattr(necks, name)
    assert DETECTORS.get("ImVoxelNet") is detector.ImVoxelNet
    assert callable(detector.get_points)      # heads import it from the detector module


def test_detector_builds_from_a_reference_model_dict():
    m = _build(KITTI_MODEL, train_cfg=None, test_cfg=dict(nms_pre=100))
    assert isinstance(m.neck_3d, necks.KittiImVoxelNeck)
    assert m.n_voxels == (216, 248, 12) and m.voxel_size == (.32, .32, .32)
    assert len(m.neck_3d.state_dict()) == 57


@pytest.mark.skipif(not ref_loader.available(), reason="reference tree not present on this machine")
def test_every_reference_config_builds():
    """tests/test_config.py:19 of the reference builds every config's detector;
    here: every imvoxelnet config's ``model`` dict builds the drop-in detector."""
    paths = sorted(glob.glob(os.path.join(ref_loader.REFERENCE_ROOT, "configs/imvoxelnet/*.py")))
    assert len(paths) == 14
    for path in paths:
        scope = {}
        exec(compile(open(path).read(), path, "exec"), scope)          # plain python-dict configs
        with torch.device("meta"):
            m = _build(scope["model"], train_cfg=scope.get("train_cfg"), test_cfg=scope.get("test_cfg"))
        assert type(m.neck_3d).__name__ == scope["model"]["neck_3d"]["type"], path


@pytest.mark.gpu
@pytest.mark.parametrize("precision", ["fp32", "tf32", "bf16"])
def test_extract_feat_matches_the_reference_loop(precision):
    from oracle import torch_necks, torch_port
    from imvoxelnet_b200 import projection as pj
    dev = "cuda:0"
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.manual_seed(0)
    sc = scenes.scannet(n_views=3, seed=2)
    meta2 = scenes.scannet(n_views=3, seed=5).img_meta
    for meta in (sc.img_meta, meta2):
        meta['img_shape'] = (128, 160, 3)
        meta['ori_shape'] = (480, 600, 3)
    meta2['lidar2img']['origin'] = np.array([.1, -.2, .6], np.float32)
    metas = [sc.img_meta, meta2]
    n_voxels, voxel_size = (16, 16, 8), (.4, .4, .4)
    m = _build(dict(KITTI_MODEL, neck=dict(type='FPN', in_channels=[256, 512, 1024, 2048], out_channels=64,
                                           num_outs=4),
                    neck_3d=dict(type='FastIndoorImVoxelNeck', in_channels=64, n_blocks=[1, 1, 1],
                                 out_channels=64),
                    n_voxels=n_voxels, voxel_size=voxel_size))
    torch_necks.randomize_norms(m.neck_3d, 3)
    m = m.eval().to(dev)
    m.backbone.to(memory_format=torch.channels_last), m.neck.to(memory_format=torch.channels_last)
    m.set_precision(precision)
    img = torch.randn(2, 3, 3, 128, 160, device=dev)
    with torch.no_grad():
        outs, valids, feats2d = m.extract_feat(img, metas, 'test')
        # the reference's loop on the same FPN features
        x = m.neck(m.backbone(img.reshape(-1, 3, 128, 160)))[0].float().unflatten(0, (2, 3))
        ref_neck = torch_necks.FastIndoorImVoxelNeck(64, [1, 1, 1], 64).eval().to(dev)
        ref_neck.load_state_dict(m.neck_3d.state_dict())
        vols, vals = [], []
        for feature, meta in zip(x, metas):
            P = pj.compute_projection(meta, 4).to(dev)
            pts = pj.get_points(n_voxels, voxel_size, meta['lidar2img']['origin']).to(dev)
            hh, ww = meta['img_shape'][0] // 4, meta['img_shape'][1] // 4
            v, s = torch_port.scene_volume(feature[:, :, :hh, :ww], pts, P)
            vols.append(v), vals.append(s)
        want = ref_neck(torch.stack(vols))
        want_valid = torch.stack(vals)
    assert feats2d is None and torch.equal(valids, want_valid)
    rel = {"fp32": 1e-5, "tf32": 5e-3, "bf16": 1e-2}[precision]
    for a, b in zip(outs, want):
        assert a.shape == b.shape
        assert_volume_close(a.float().cpu().numpy(), b.cpu().numpy(), rel=rel, atol=1e-6 if precision == "fp32" else 1e-3)


@pytest.mark.gpu
@pytest.mark.parametrize("precision,channels", [("fp32", 16), ("tf32", 64)])
def test_extract_feat_is_differentiable_in_training_mode(precision, channels):
    """images -> 2D net -> fused back-projection -> neck, all under autograd: the loss
    gradient reaches the 2D backbone and the neck parameters (forward_train, :82-87); in tf32 the
    neck's convolutions and their gradients run on the tensor cores."""
    dev = "cuda:0"
    torch.manual_seed(0)
    sc = scenes.scannet(n_views=2, seed=2)
    sc.img_meta['img_shape'] = (96, 128, 3)
    sc.img_meta['ori_shape'] = (480, 640, 3)
    m = _build(dict(KITTI_MODEL, neck=dict(type='FPN', in_channels=[256, 512, 1024, 2048], out_channels=channels,
                                           num_outs=4),
                    neck_3d=dict(type='FastIndoorImVoxelNeck', in_channels=channels, n_blocks=[1, 1, 1],
                                 out_channels=channels // 2 if channels < 64 else 64),
                    n_voxels=(8, 8, 8), voxel_size=(.6, .6, .4)))
    m = m.train().to(dev)
    m.set_precision(precision)
    img = torch.randn(2, 2, 3, 96, 128, device=dev)
    outs, valids, _ = m.extract_feat(img, [sc.img_meta, sc.img_meta], 'train')
    assert valids.any()
    sum(o.square().mean() for o in outs).backward()
    g_neck = m.neck_3d.down_layer_0[0].conv1.weight.grad
    g_2d = m.backbone.conv1.weight.grad
    assert g_neck is not None and torch.isfinite(g_neck).all() and float(g_neck.abs().max()) > 0
    assert g_2d is not None and torch.isfinite(g_2d).all() and float(g_2d.abs().max()) > 0
