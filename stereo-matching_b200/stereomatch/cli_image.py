#!/usr/bin/env python
"""Disparity map of one stereo pair from image files -- headless.

Arguments and options of reference stereomatch/cli_image.py:34-62 (left-image right-image
max-disparity output-depthmap, -cm/-am/-dm, -c).  What differs: the pipeline always runs on
CUDA (-c is accepted and implied: there is no CPU path), nothing is shown (no matplotlib; the
-sd / -fig options of the reference are accepted and ignored), and the output file is the
disparity map itself as 16-bit grayscale (PNG through PIL, or binary PGM "P5" with maxval
65535 when the name ends in .pgm) instead of a colour-mapped figure.

    python -m stereomatch.cli_image left.png right.png 128 disparity.png -cm ssd -am sgm -dm wta
"""
import argparse
import sys

import numpy as np
import torch

from .cli_common import AGGREGATION_METHODS, COST_METHODS, DISPARITY_METHODS, create_pipeline


def read_gray_f32(path: str) -> torch.Tensor:
    """Image file -> [H,W] float32 in 0..255, the reference's conversion
    (cli_image.py:71-75: PIL `convert('L')`, ITU-R 601-2 luma, then `.float()`)."""
    from PIL import Image
    with Image.open(path) as im:
        return torch.from_numpy(np.array(im.convert("L"))).float()


def write_u16(path: str, disparity: np.ndarray) -> None:
    """[H,W] integer disparity map -> 16-bit grayscale file (reference casts to uint16 too,
    cli_image.py:81-82)."""
    d16 = np.ascontiguousarray(disparity.astype(np.uint16))
    if path.lower().endswith(".pgm"):
        h, w = d16.shape
        with open(path, "wb") as f:
            f.write(f"P5\n{w} {h}\n65535\n".encode("ascii"))
            f.write(d16.astype(">u2").tobytes())  # PGM is big-endian
        return
    from PIL import Image
    Image.fromarray(d16).save(path)  # mode I;16


def read_u16(path: str) -> np.ndarray:
    """Inverse of write_u16 (used by the tests and by callers that post-process)."""
    if path.lower().endswith(".pgm"):
        with open(path, "rb") as f:
            assert f.readline().strip() == b"P5"
            w, h = (int(v) for v in f.readline().split())
            maxval = int(f.readline())
            dt = ">u2" if maxval > 255 else "u1"
            return np.frombuffer(f.read(), dtype=dt, count=w * h).reshape(h, w).astype(np.uint16)
    from PIL import Image
    with Image.open(path) as im:
        return np.array(im).astype(np.uint16)


def build_parser() -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    p.add_argument("left_image", metavar="left-image", help="Left image")
    p.add_argument("right_image", metavar="right-image", help="Right image")
    p.add_argument("max_disparity", metavar="max-disparity", type=int,
                   help="Maximum disparity for stereo matching.")
    p.add_argument("output_depthmap", metavar="output-depthmap", help="Output file (.png or .pgm, 16 bit).")
    p.add_argument("-cm", "--cost-method", choices=list(COST_METHODS), default="ssd")
    p.add_argument("-am", "--aggregation-method", choices=list(AGGREGATION_METHODS), default=None)
    p.add_argument("-dm", "--disparity-method", choices=list(DISPARITY_METHODS), default="wta")
    p.add_argument("-c", "--cuda-on", action="store_true", help="Accepted for compatibility; always on.")
    p.add_argument("-sd", "--show-depthmap", action="store_true", help="Ignored (headless).")
    p.add_argument("-fig", "--figure", action="store_true", help="Ignored (headless).")
    p.add_argument("--device", default="cuda:0")
    return p


def main(argv=None) -> int:
    args = build_parser().parse_args(argv)
    pipeline = create_pipeline(args.cost_method, args.disparity_method, args.aggregation_method,
                               args.max_disparity)
    left = read_gray_f32(args.left_image)
    right = read_gray_f32(args.right_image)
    if left.shape != right.shape:
        print(f"error: image sizes differ: {tuple(left.shape)} vs {tuple(right.shape)}", file=sys.stderr)
        return 2
    disparity = pipeline.estimate(left.to(args.device), right.to(args.device))
    write_u16(args.output_depthmap, disparity.cpu().numpy())
    return 0


if __name__ == "__main__":
    sys.exit(main())
