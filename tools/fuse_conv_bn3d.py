#!/usr/bin/env python
"""Fold BatchNorm3d into the 3D convolutions of a checkpoint's ``neck_3d`` (the 3D
counterpart of the reference's tools/fuse_conv_bn.py, which only handles 2D pairs).

    python tools/fuse_conv_bn3d.py in.pth out.pth --neck KittiImVoxelNeck \
        --neck-args '{"in_channels": 64, "out_channels": 256}' [--prefix neck_3d.]

Reads ``in.pth`` (a state dict, or an mmcv-style {'state_dict': ...} checkpoint), writes the
same file with the ``prefix`` keys replaced by their folded counterparts.  Load the result into
a neck on which ``imvoxelnet_b200.checkpoint.fold_conv_bn3d`` has been called.
"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), "..")))
from imvoxelnet_b200 import checkpoint, necks  # noqa: E402


def main():
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("checkpoint")
    ap.add_argument("out")
    ap.add_argument("--neck", required=True, choices=["FastIndoorImVoxelNeck", "ImVoxelNeck", "KittiImVoxelNeck",
                                                      "NuScenesImVoxelNeck"])
    ap.add_argument("--neck-args", required=True, help="JSON object of the neck's constructor arguments")
    ap.add_argument("--prefix", default="neck_3d.")
    args = ap.parse_args()
    blob = torch.load(args.checkpoint, map_location="cpu")
    wrapped = isinstance(blob, dict) and "state_dict" in blob
    sd = blob["state_dict"] if wrapped else blob
    kw = json.loads(args.neck_args)
    folded = checkpoint.fold_state_dict(lambda: getattr(necks, args.neck)(**kw), sd, prefix=args.prefix)
    if wrapped:
        blob = dict(blob, state_dict=folded)
    else:
        blob = folded
    torch.save(blob, args.out)
    n_before = sum(1 for k in sd if k.startswith(args.prefix))
    n_after = sum(1 for k in folded if k.startswith(args.prefix))
    print(f"{args.neck}: {n_before} -> {n_after} tensors under '{args.prefix}', written to {args.out}")


if __name__ == "__main__":
    main()
