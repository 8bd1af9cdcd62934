"""Construction arguments and input shapes of the neck fixtures (shared by the
generator, which needs the reference tree, and the tests, which must not)."""

NECK_CASES = {
    "FastIndoorImVoxelNeck": (dict(in_channels=8, n_blocks=[1, 1, 1], out_channels=8), (2, 8, 8, 12, 4)),
    "ImVoxelNeck": (dict(channels=[4, 8, 16, 32], out_channels=4, down_layers=[1, 2, 3, 4],
                         up_layers=[3, 2, 1], conditional=False), (1, 4, 16, 8, 8)),
    "KittiImVoxelNeck": (dict(in_channels=4, out_channels=8), (2, 4, 9, 10, 12)),
    "NuScenesImVoxelNeck": (dict(in_channels=4, out_channels=8), (1, 4, 12, 10, 12)),
}

# dense heads: reference class -> (constructor kwargs, oracle class + flags, level shapes [B,C,X,Y,Z])
HEAD_CASES = {
    "ScanNetImVoxelHeadV2": (dict(n_classes=5, n_channels=8, n_reg_outs=6, n_scales=3, limit=27),
                             ("HeadV2", dict(n_classes=5, n_channels=8, n_reg_outs=6, n_scales=3, box_angle=False)),
                             [(2, 8, 8, 12, 4), (2, 8, 4, 6, 2), (2, 8, 2, 3, 1)]),
    "SunRgbdImVoxelHeadV2": (dict(n_classes=3, n_channels=8, n_reg_outs=7, n_scales=3, limit=27),
                             ("HeadV2", dict(n_classes=3, n_channels=8, n_reg_outs=7, n_scales=3, box_angle=True)),
                             [(1, 8, 8, 8, 4), (1, 8, 4, 4, 2), (1, 8, 2, 2, 1)]),
    "ScanNetImVoxelHead": (dict(n_classes=4, n_channels=8, n_convs=2, n_reg_outs=6),
                           ("HeadV1", dict(n_classes=4, n_channels=8, n_convs=2, n_reg_outs=6, box_angle=False)),
                           [(1, 8, 8, 8, 8), (1, 8, 4, 4, 4), (1, 8, 2, 2, 2)]),
    "SunRgbdImVoxelHead": (dict(n_classes=6, n_channels=4, n_convs=1, n_reg_outs=7),
                           ("HeadV1", dict(n_classes=6, n_channels=4, n_convs=1, n_reg_outs=7, box_angle=True)),
                           [(2, 4, 6, 8, 4), (2, 4, 3, 4, 2), (2, 4, 2, 2, 1)]),
}
# valid-mask resampling: (input [B,1,X,Y,Z], target sizes); integer ratios as in the configs, and ragged ones
MASK_CASES = {
    "indoor_fast": ((2, 1, 40, 40, 16), [(40, 40, 16), (20, 20, 8), (10, 10, 4)]),
    "ragged": ((1, 1, 13, 9, 7), [(13, 9, 7), (6, 5, 3), (3, 2, 2), (20, 11, 9)]),
}
