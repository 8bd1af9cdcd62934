"""Construction arguments and input shapes of the neck fixtures (shared by the
generator, which needs the reference tree, and the tests, which must not)."""

NECK_CASES = {
    "FastIndoorImVoxelNeck": (dict(in_channels=8, n_blocks=[1, 1, 1], out_channels=8), (2, 8, 8, 12, 4)),
    "ImVoxelNeck": (dict(channels=[4, 8, 16, 32], out_channels=4, down_layers=[1, 2, 3, 4],
                         up_layers=[3, 2, 1], conditional=False), (1, 4, 16, 8, 8)),
    "KittiImVoxelNeck": (dict(in_channels=4, out_channels=8), (2, 4, 9, 10, 12)),
    "NuScenesImVoxelNeck": (dict(in_channels=4, out_channels=8), (1, 4, 12, 10, 12)),
}
