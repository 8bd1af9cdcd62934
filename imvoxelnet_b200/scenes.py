"""Seeded synthetic scenes for the five BASELINE.json configurations.

There are no datasets on the build or GPU machines, so tests and ``bench.py``
use camera rigs that look like the real ones.  Every
builder returns a ``Scene`` whose ``img_meta`` has exactly the keys the
reference reads (``img_shape``, ``ori_shape``, ``lidar2img.{intrinsic,
extrinsic, origin}``; detectors/imvoxelnet.py:65-68, 117-126).
"""
from dataclasses import dataclass, field

import numpy as np

STRIDE = 4


@dataclass
class Scene:
    name: str
    n_views: int
    channels: int
    n_voxels: tuple
    voxel_size: tuple
    img_meta: dict
    map_hw: tuple            # padded stride-4 feature map (Hf, Wf)
    neck: str = ""
    extra: dict = field(default_factory=dict)

    @property
    def feat_hw(self):       # un-padded extent the reference slices to (:67-69)
        h, w = self.img_meta['img_shape'][:2]
        return h // STRIDE, w // STRIDE

    @property
    def n_total_voxels(self):
        return int(np.prod(self.n_voxels))

    @property
    def voxel_views(self):
        return self.n_views * self.n_total_voxels


def _look_at(eye, target, up=(0., 0., 1.)):
    """World->camera [4,4]: camera z forward, x right, y down (OpenCV)."""
    eye, target, up = (np.asarray(a, np.float64) for a in (eye, target, up))
    fwd = target - eye
    fwd /= np.linalg.norm(fwd)
    right = np.cross(fwd, up)
    right /= np.linalg.norm(right)
    down = np.cross(fwd, right)
    rot = np.stack([right, down, fwd])
    ext = np.eye(4)
    ext[:3, :3] = rot
    ext[:3, 3] = -rot @ eye
    return ext.astype(np.float32)


def _pinhole(f, cx, cy):
    k = np.eye(4, dtype=np.float32)
    k[0, 0] = k[1, 1] = f
    k[0, 2], k[1, 2] = cx, cy
    return k


def _pad(x, m):
    return (x + m - 1) // m * m


def sunrgbd(n_voxels=(40, 40, 16), voxel_size=(.16, .16, .16), channels=256,
            neck="FastIndoorImVoxelNeck", name="imvoxelnet_sunrgbd_fast"):
    """cfg1 / cfg5: one 640x480 pinhole looking along +y (depth frame)."""
    meta = dict(img_shape=(480, 640, 3), ori_shape=(480, 640, 3),
                lidar2img=dict(intrinsic=_pinhole(529.5, 320., 240.),
                               extrinsic=[_look_at((0, 0, 0), (0, 1, 0))],
                               origin=np.array([0., 3., -1.], np.float32)))
    return Scene(name, 1, channels, tuple(n_voxels), tuple(voxel_size), meta, (120, 160), neck)


def total_sunrgbd():
    """cfg5: 80x80x32 grid at 0.08 m, 64 channels, ImVoxelNeck."""
    return sunrgbd((80, 80, 32), (.08, .08, .08), 64, "ImVoxelNeck", "imvoxelnet_total_sunrgbd")


def kitti():
    """cfg2: 1280x384 forward camera, 216x248x12 grid at 0.32 m, 64 channels."""
    pcr = np.array([0, -39.68, -3, 69.12, 39.68, 1], np.float32)
    ext = np.array([[0, -1, 0, 0.], [0, 0, -1, -0.08], [1, 0, 0, -0.27], [0, 0, 0, 1]], np.float32)
    k = _pinhole(721.5377, 609.5593, 172.854)
    meta = dict(img_shape=(384, 1280, 3), ori_shape=(375, 1242, 3),
                lidar2img=dict(intrinsic=k, extrinsic=[ext],
                               origin=((pcr[:3] + pcr[3:]) / 2).astype(np.float32)))
    return Scene("imvoxelnet_kitti", 1, 64, (216, 248, 12), (.32, .32, .32), meta,
                 (96, 320), "KittiImVoxelNeck")


def scannet(n_views=50, seed=0):
    """cfg3: ``n_views`` cameras on a 2.5 m ring at 1.5 m, looking at the room."""
    rng = np.random.default_rng(seed)
    exts = []
    for i in range(n_views):
        th = 2 * np.pi * (i + 0.25 * rng.random()) / n_views
        eye = (2.5 * np.cos(th), 2.5 * np.sin(th), 1.5 + 0.1 * rng.standard_normal())
        exts.append(_look_at(eye, (0.3 * rng.standard_normal(), 0.3 * rng.standard_normal(), 0.5)))
    meta = dict(img_shape=(480, 640, 3), ori_shape=(480, 640, 3),
                lidar2img=dict(intrinsic=_pinhole(580., 320., 240.), extrinsic=exts,
                               origin=np.array([0., 0., .5], np.float32)))
    return Scene("imvoxelnet_scannet_fast", n_views, 256, (40, 40, 16), (.16, .16, .16),
                 meta, (120, 160), "FastIndoorImVoxelNeck")


def nuscenes(reference_grid=False):
    """cfg4: six 1600x900 cameras at 60 degree yaw spacing; intrinsic = I and the
    full lidar->image matrix rides in ``extrinsic`` (as the reference's
    nuScenes dataset does, nuscenes_monocular_dataset.py:19-22).  BASELINE's
    grid is 200x200x4 (no reference neck accepts Z=4); ``reference_grid``
    selects the reference's own 312x312x12 at 0.32 m."""
    k = _pinhole(1266.4, 816.3, 491.5)
    exts = []
    for i in range(6):
        th = np.deg2rad(60. * i)
        eye = (0.9 * np.cos(th), 0.9 * np.sin(th), 1.6)
        tgt = (eye[0] + np.cos(th), eye[1] + np.sin(th), 1.55)
        exts.append((k @ _look_at(eye, tgt)).astype(np.float32))
    if reference_grid:
        nv, vs, neck = (312, 312, 12), (.32, .32, .32), "NuScenesImVoxelNeck"
        origin = np.array([0., 0., -1.08], np.float32)
    else:
        nv, vs, neck = (200, 200, 4), (.5, .5, 2.), ""
        origin = np.array([0., 0., -1.], np.float32)
    meta = dict(img_shape=(900, 1600, 3), ori_shape=(900, 1600, 3),
                lidar2img=dict(intrinsic=np.eye(4, dtype=np.float32), extrinsic=exts, origin=origin))
    return Scene("imvoxelnet_nuscenes" + ("_refgrid" if reference_grid else ""), 6, 64, nv, vs,
                 meta, (_pad(900, 32) // STRIDE, 400), neck)


BUILDERS = {
    "sunrgbd_fast": sunrgbd,
    "kitti": kitti,
    "scannet_fast": scannet,
    "nuscenes": nuscenes,
    "nuscenes_refgrid": lambda: nuscenes(True),
    "total_sunrgbd": total_sunrgbd,
}


def build(name, **kw):
    return BUILDERS[name](**kw)


def random_features(scene, seed=0, batch=None, channels_last=True, map_hw=None, channels=None):
    """Seeded N(0,1) fp32 feature maps ``[V, C, Hf, Wf]`` (or ``[B, V, ...]``).

    Always created in NCHW order from the same generator stream so that the
    values do not depend on the physical layout; ``channels_last`` only changes
    strides (the layout the product hands to the kernel)."""
    import torch
    g = torch.Generator().manual_seed(seed)
    hf, wf = map_hw or scene.map_hw
    c = channels or scene.channels
    lead = (scene.n_views,) if batch is None else (batch * scene.n_views,)
    x = torch.randn(lead + (c, hf, wf), generator=g, dtype=torch.float32)
    if channels_last:
        x = x.contiguous(memory_format=torch.channels_last)
    if batch is not None:
        x = x.view(batch, scene.n_views, c, hf, wf) if not channels_last else \
            x.as_strided((batch, scene.n_views, c, hf, wf),
                         (scene.n_views * x.stride(0),) + tuple(x.stride()))
    return x
