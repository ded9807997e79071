"""Middlebury-2021 style datasets on disk (API of reference stereomatch/data.py:14-93).

    <dataset_dir>/<scene>/im0.png  im1.png  disp0.pfm  disp1.pfm  calib.txt

Same class and method names and the same dictionaries as the reference.  What differs: scene
directories are ordered by a natural sort implemented here (the reference uses the `natsort`
package) and the PFM ground truth is read by `read_pfm` below (the reference calls
`cv2.imread(..., IMREAD_UNCHANGED)`, data.py:86; both return rows top to bottom with +inf where
the disparity is unknown).  `write_pfm` is new: it lets tests and tools write scenes in this layout.
"""
from __future__ import annotations

import re
from pathlib import Path
from typing import Optional

import numpy as np
import torch


def _parse_middlebury_calib(filepath) -> dict:
    """calib.txt: `name=value` lines; width, height and ndisp are kept (reference data.py:14-24)."""
    props = {}
    with open(filepath, "r", encoding="ascii") as file:
        for line in file:
            if "=" not in line:
                continue
            prop_name, prop_value = line.split("=", 1)
            props[prop_name.strip()] = prop_value.strip()
    return dict(width=int(props["width"]), height=int(props["height"]), ndisp=int(props["ndisp"]))


def read_pfm(path) -> np.ndarray:
    """Portable float map -> float32 [H, W] (`Pf`) or [H, W, 3] (`PF`), rows TOP to bottom.

    Header: type line, `width height`, scale (negative = little-endian samples); the samples
    follow bottom row first.  Unknown disparities are +inf in Middlebury ground truth."""
    with open(path, "rb") as f:
        kind = f.readline().strip()
        if kind not in (b"Pf", b"PF"):
            raise ValueError(f"{path}: not a PFM file")
        dims = f.readline()
        while dims.startswith(b"#"):
            dims = f.readline()
        m = re.match(rb"^\s*(\d+)\s+(\d+)\s*$", dims)
        if not m:
            raise ValueError(f"{path}: malformed PFM header")
        width, height = int(m.group(1)), int(m.group(2))
        scale = float(f.readline().strip())
        channels = 3 if kind == b"PF" else 1
        data = np.frombuffer(f.read(width * height * channels * 4), dtype="<f4" if scale < 0 else ">f4")
    if data.size != width * height * channels:
        raise ValueError(f"{path}: truncated PFM data")
    img = data.reshape(height, width, channels) if channels == 3 else data.reshape(height, width)
    return np.ascontiguousarray(img[::-1]).astype(np.float32)


def write_pfm(path, image: np.ndarray, scale: float = 1.0) -> None:
    """float32 [H, W] or [H, W, 3] -> little-endian PFM (bottom row first, negative scale)."""
    image = np.asarray(image, dtype=np.float32)
    if image.ndim not in (2, 3) or (image.ndim == 3 and image.shape[2] != 3):
        raise ValueError("write_pfm: expected [H, W] or [H, W, 3]")
    with open(path, "wb") as f:
        f.write(b"PF\n" if image.ndim == 3 else b"Pf\n")
        f.write(f"{image.shape[1]} {image.shape[0]}\n".encode("ascii"))
        f.write(f"{-abs(float(scale))}\n".encode("ascii"))
        f.write(np.ascontiguousarray(image[::-1]).astype("<f4").tobytes())


def _natural_key(path: Path):
    return [int(t) if t.isdigit() else t.lower() for t in re.split(r"(\d+)", path.name)]


class MiddleburyDataset:
    """Pytorch-like dataset over a directory of Middlebury scenes (reference data.py:27-93)."""

    def __init__(self, dataset_dir: str, max_size: Optional[int] = None):
        dataset_dir = Path(dataset_dir)
        if not dataset_dir.is_dir():
            raise RuntimeError(f"MiddleburyDataset: {dataset_dir} must be a directory")
        self.images = []
        self.disps = []
        self.calibs = []
        sample_dirs = sorted(dataset_dir.iterdir(), key=_natural_key)
        if max_size is not None:
            sample_dirs = sample_dirs[:max_size]
        for sample_dir in sample_dirs:
            if not sample_dir.is_dir():
                continue
            self.images.append((sample_dir / "im0.png", sample_dir / "im1.png"))
            self.disps.append((sample_dir / "disp0.pfm", sample_dir / "disp1.pfm"))
            self.calibs.append(_parse_middlebury_calib(sample_dir / "calib.txt"))

    def __getitem__(self, idx: int):
        item = self.get_stereo_pair(idx)
        item.update(self.get_ground_truth(idx))
        return item

    def get_stereo_pair(self, idx: int):
        """dict(stereo_name, left, right, max_disparity); images as stored (uint8 [H,W,3] or [H,W])."""
        from PIL import Image
        left_img_path, right_img_path = self.images[idx]
        with Image.open(left_img_path) as im:
            left_img = np.array(im)
        with Image.open(right_img_path) as im:
            right_img = np.array(im)
        return dict(stereo_name=left_img_path.parent.name, left=torch.from_numpy(left_img),
                    right=torch.from_numpy(right_img), max_disparity=self.calibs[idx]["ndisp"])

    def get_ground_truth(self, idx: int):
        """dict(stereo_name, gt_disparity float32 [H,W] with +inf where unknown, max_disparity)."""
        disp_path = self.disps[idx][0]
        return dict(stereo_name=disp_path.parent.name, gt_disparity=torch.from_numpy(read_pfm(disp_path)),
                    max_disparity=self.calibs[idx]["ndisp"])

    def __len__(self):
        return len(self.images)
