"""B200-native image->voxel path of ImVoxelNet (backproject + 3D neck)."""
__version__ = "0.1.0"
