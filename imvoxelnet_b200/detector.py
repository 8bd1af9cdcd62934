"""Drop-in ``ImVoxelNet`` detector: same registry name, constructor, and
``extract_feat`` / ``forward_train`` / ``forward_test`` / ``simple_test``
signatures as mmdet3d/models/detectors/imvoxelnet.py:8-112, with the hot path
(per-scene loop of :57-78 and the 3D neck of :79) replaced by one fused
back-projection launch per batch and the CUDA neck kernels.

2D backbone, FPN and heads are not part of this library: they are built from
the config through mmdet when it is installed; otherwise the backbone/FPN
stand-ins of ``backbone2d`` are used and heads that cannot be built are left
out (``extract_feat`` still works; ``simple_test`` then raises).
"""
import warnings

import torch
from torch import nn

from . import backbone2d, ops
from . import necks as _necks  # noqa: F401  (importing it registers the four 3D necks)
from .ops import backproject  # noqa: F401  (module-level function of detectors/imvoxelnet.py:145)
from .projection import compute_projection, compute_projection_batch, get_extrinsics, get_points  # noqa: F401  (re-exported as in detectors/imvoxelnet.py:133, :164; heads import get_points from here, imvoxel_head.py:7)
from .registry import DETECTORS, HAVE_MMDET, NECKS, build_from, register

if HAVE_MMDET:  # pragma: no cover
    from mmdet.models import build_backbone, build_head, build_neck
    from mmdet.models.detectors import BaseDetector as _Base
else:
    _Base = nn.Module

    def build_backbone(cfg):
        cfg = dict(cfg)
        typ = cfg.pop('type')
        if typ != 'ResNet':
            raise KeyError(f"backbone {typ}: only ResNet has a stand-in without mmdet")
        return backbone2d.ResNet(**cfg)

    def build_neck(cfg):
        if cfg.get('type') == 'FPN':
            cfg = dict(cfg)
            cfg.pop('type')
            return backbone2d.FPN(**cfg)
        return build_from(NECKS, cfg)

    def build_head(cfg):
        warnings.warn(f"head {cfg.get('type')} is the reference's PyTorch (mmdet3d) and cannot be built "
                      "without mmdet/mmdet3d; the detector is built without it")
        return None


def _bbox3d2result(bboxes, scores, labels):
    try:  # pragma: no cover
        from mmdet3d.core import bbox3d2result
        return bbox3d2result(bboxes, scores, labels)
    except ImportError:
        return dict(boxes_3d=bboxes.to('cpu'), scores_3d=scores.cpu(), labels_3d=labels.cpu())


@register(DETECTORS)
class ImVoxelNet(_Base):
    #: dtype of the voxel volume handed to the 3D neck ("fp32" parity mode, or
    #: "bf16" together with ``neck_3d.precision = "bf16"``)
    precision = "fp32"

    def __init__(self, backbone, neck, neck_3d, bbox_head, n_voxels, voxel_size, head_2d=None,
                 train_cfg=None, test_cfg=None, pretrained=None):
        super().__init__()
        self.backbone = backbone if isinstance(backbone, nn.Module) else build_backbone(backbone)
        self.neck = neck if isinstance(neck, nn.Module) else build_neck(neck)
        self.neck_3d = neck_3d if isinstance(neck_3d, nn.Module) else build_from(NECKS, neck_3d) \
            if not HAVE_MMDET else build_neck(neck_3d)
        if isinstance(bbox_head, nn.Module) or bbox_head is None:
            self.bbox_head = bbox_head
        else:
            bbox_head = dict(bbox_head)
            bbox_head.update(train_cfg=train_cfg, test_cfg=test_cfg)
            self.bbox_head = build_head(bbox_head)
        if self.bbox_head is not None:
            self.bbox_head.voxel_size = voxel_size
            from . import heads
            if heads.supports(self.bbox_head):      # ImVoxelHead / ImVoxelHeadV2: fused convolution tail
                heads.accelerate(self.bbox_head)
        self.head_2d = (head_2d if isinstance(head_2d, nn.Module) else build_head(head_2d)) \
            if head_2d is not None else None
        self.n_voxels = n_voxels
        self.voxel_size = voxel_size
        self.train_cfg = train_cfg
        self.test_cfg = test_cfg
        self.init_weights(pretrained=pretrained)

    def init_weights(self, pretrained=None):
        if HAVE_MMDET:  # pragma: no cover
            super().init_weights(pretrained)
        for m in (self.backbone, self.neck, self.neck_3d, self.bbox_head, self.head_2d):
            if m is not None and hasattr(m, 'init_weights'):
                try:
                    m.init_weights(pretrained=pretrained) if m is self.backbone else m.init_weights()
                except TypeError:
                    m.init_weights()

    @staticmethod
    def _compute_projection(img_meta, stride, angles):
        """Same static method as detectors/imvoxelnet.py:114-129: [V,3,4] fp32 matrices on the CPU."""
        return compute_projection(img_meta, stride, angles)

    def set_precision(self, precision):
        """"fp32" (1e-5 of the reference) or "bf16" (tensor-core neck, 1e-2)."""
        self.precision = precision
        self.neck_3d.precision = precision
        return self

    def extract_feat(self, img, img_metas, mode):
        """img [B,V,3,H,W] -> (list of neck outputs, valids [B,1,X,Y,Z] bool, features_2d)."""
        batch_size = img.shape[0]
        img = img.reshape([-1] + list(img.shape)[2:])
        x = self.backbone(img)
        features_2d = self.head_2d.forward(x[-1], img_metas) if self.head_2d is not None else None
        x = self.neck(x)[0]
        stride = img.shape[-1] / x.shape[-1]
        assert stride == 4  # as upstream (:54)
        stride = int(stride)
        x = x.unflatten(0, (batch_size, -1))                     # [B,V,C,Hf,Wf], strides kept

        # predicted pitch and roll for SUN RGB-D Total at test time (:59-60)
        angles = features_2d[0] if features_2d is not None and mode == 'test' else None
        hws = [(m['img_shape'][0] // stride, m['img_shape'][1] // stride) for m in img_metas]
        origins = [m['lidar2img']['origin'] for m in img_metas]
        n_views = {len(m['lidar2img']['extrinsic']) for m in img_metas}
        if angles is None and x.is_cuda and len(n_views) == 1:
            projection = compute_projection_batch(img_metas, stride, x.device)   # one copy + one launch
        else:
            projection = torch.stack([compute_projection(m, stride, angles) for m in img_metas])
        out_dtype = torch.bfloat16 if self.precision == "bf16" else torch.float32
        # one launch for the whole batch: get_points, backproject, sum / count / mean / mask, stack
        if not (x.dtype == torch.bfloat16 and self.precision == "bf16" and x.stride(2) == 1
                and not torch.is_grad_enabled()):
            x = x.float()       # bf16 maps are consumed as they are in bf16 inference
        volume, valids, _ = ops.backproject_scenes(x, hws, projection, self.n_voxels,
                                                   self.voxel_size, origins, out_dtype=out_dtype)
        x = self.neck_3d(volume)
        return x, valids, features_2d

    def forward(self, img, img_metas, return_loss=True, **kwargs):
        if HAVE_MMDET:  # pragma: no cover
            return super().forward(img, img_metas, return_loss=return_loss, **kwargs)
        if return_loss:
            return self.forward_train(img, img_metas, **kwargs)
        return self.forward_test(img, img_metas, **kwargs)

    def forward_train(self, img, img_metas, gt_bboxes_3d, gt_labels_3d, **kwargs):
        """As upstream (:82-87).  The hot path is differentiable: the fused back-projection has
        a backward kernel and the necks run batch-statistics BatchNorm + autograd through this
        library's fp32 kernels (``precision`` must be "fp32"); the heads are the reference's."""
        if self.bbox_head is None:
            raise RuntimeError("no bbox_head was built (mmdet3d heads are not available here)")
        x, valids, features_2d = self.extract_feat(img, img_metas, 'train')
        losses = self.bbox_head.forward_train(x, valids.float(), img_metas, gt_bboxes_3d, gt_labels_3d)
        if self.head_2d is not None:
            losses.update(self.head_2d.loss(*features_2d, img_metas))
        return losses

    def forward_test(self, img, img_metas, **kwargs):
        return self.simple_test(img, img_metas)

    def simple_test(self, img, img_metas):
        if self.bbox_head is None:
            raise RuntimeError("no bbox_head was built (mmdet3d heads are not available here)")
        x, valids, features_2d = self.extract_feat(img, img_metas, 'test')
        x = self.bbox_head(x)
        bbox_list = self.bbox_head.get_bboxes(*x, valids.float(), img_metas)
        bbox_results = [_bbox3d2result(b, s, l) for b, s, l in bbox_list]
        if self.head_2d is not None:
            angles, layouts = self.head_2d.get_bboxes(*features_2d, img_metas)
            for i in range(len(img)):
                bbox_results[i]['angles'] = angles[i]
                bbox_results[i]['layout'] = layouts[i]
        return bbox_results

    def aug_test(self, imgs, img_metas):
        pass

    def show_results(self, *args, **kwargs):
        pass
