"""Matching-cost volumes (API of reference stereomatch/cost.py)."""
from typing import Optional

import torch

from ._common import empty_tensor
from ._cstereomatch import CostOps as _CostOps
from .cuda_texture import CUDATexture


class SSD:
    """Windowed sum of squared differences (reference stereomatch/cost.py:13-48).

    Attributes (public and mutable, as in the reference): max_disparity, kernel_size
    (half-width of the half-open 2k x 2k window, reference src/ssd.cu:36-42),
    cost_volume_dtype (float32 is the only dtype the reference's dispatch accepts).
    """

    def __init__(self, max_disparity: int, kernel_size: int = 7,
                 cost_volume_dtype: torch.dtype = torch.float):
        self.max_disparity = max_disparity
        self.kernel_size = kernel_size
        self.cost_volume_dtype = cost_volume_dtype

    def __call__(self, left_image: torch.Tensor, right_image: torch.Tensor,
                 cost_volume: Optional[torch.Tensor] = None) -> torch.Tensor:
        """[H,W] x [H,W] float32 CUDA images -> [H,W,max_disparity] cost volume
        (+inf where d > x)."""
        cost_volume = empty_tensor(left_image.size(0), left_image.size(1), self.max_disparity,
                                   dtype=self.cost_volume_dtype, device=left_image.device,
                                   reuse_tensor=cost_volume)
        _CostOps.compute_ssd(left_image, right_image, cost_volume, self.kernel_size)
        return cost_volume

    def cost_and_reduce(self, left_image: torch.Tensor, right_image: torch.Tensor,
                        disparity_img: Optional[torch.Tensor] = None) -> Optional[torch.Tensor]:
        """New: SSD cost followed by winner-takes-all in one pass; the cost volume is never written
        (28 bytes of HBM traffic per pixel instead of 8 per pixel-disparity).  Returns the int32
        disparity map -- exactly WinnerTakesAll()(self(left, right)) -- or None when the shape is
        not supported (kernel_size not in 1..10, max_disparity % 4 != 0), in which case the caller
        runs the two stages."""
        if self.cost_volume_dtype not in (torch.float, torch.float32):
            return None
        disparity_img = empty_tensor(left_image.size(0), left_image.size(1), dtype=torch.int32,
                                     device=left_image.device, reuse_tensor=disparity_img)
        ok = _CostOps.compute_ssd_wta(left_image, right_image, disparity_img, self.max_disparity,
                                      self.kernel_size)
        return disparity_img if ok else None


class SSDTexture:
    """SSD over CUDA texture inputs (reference stereomatch/cost.py:51-77): same volume as SSD on
    the images the textures hold.  Float32, unnormalised textures of equal size."""

    def __init__(self, max_disparity: int, kernel_size: int = 7):
        self.max_disparity = max_disparity
        self.kernel_size = kernel_size

    def __call__(self, left_image: CUDATexture, right_image: CUDATexture,
                 cost_volume: Optional[torch.Tensor] = None) -> torch.Tensor:
        if left_image.use_normalized_coords or right_image.use_normalized_coords:
            raise RuntimeError("Texture coordinates can't be normalized for this implementation")
        # the reference hard-codes cuda:0 (cost.py:72); here the volume goes where the texture is
        cost_volume = empty_tensor(left_image.height, left_image.width, self.max_disparity,
                                   dtype=torch.float32, device=left_image._device,
                                   reuse_tensor=cost_volume)
        _CostOps.compute_ssd(left_image, right_image, cost_volume, self.kernel_size)
        return cost_volume


class Birchfield:
    """Birchfield-Tomasi sampling-insensitive cost, 1-D window of 8 taps
    (reference stereomatch/cost.py:80-101; the reference ignores kernel_size and always
    uses 4, reference src/birchfield_cost.cu:159).  Entries d > x, which the reference
    leaves unwritten, are +inf."""

    def __init__(self, max_disparity: int, kernel_size: int = 7):
        self.max_disparity = max_disparity
        self.kernel_size = kernel_size
        self.kerne_size = kernel_size  # the reference's (misspelt) attribute, cost.py:90

    def __call__(self, left_image: torch.Tensor, right_image: torch.Tensor,
                 cost_volume: Optional[torch.Tensor] = None) -> torch.Tensor:
        cost_volume = empty_tensor(left_image.size(0), left_image.size(1), self.max_disparity,
                                   dtype=torch.float32, device=left_image.device,
                                   reuse_tensor=cost_volume)
        _CostOps.compute_birchfield(left_image.float(), right_image.float(), cost_volume, 4)
        return cost_volume

    def cost_and_reduce(self, left_image: torch.Tensor, right_image: torch.Tensor,
                        disparity_img: Optional[torch.Tensor] = None) -> Optional[torch.Tensor]:
        """New: Birchfield cost followed by winner-takes-all in one kernel; the cost volume is never
        written (4 bytes of HBM traffic per pixel instead of 8 per pixel-disparity).  Returns the
        int32 disparity map -- exactly WinnerTakesAll()(self(left, right)) -- or None when the shape
        is not supported (max_disparity % 32 != 0), in which case the caller runs the two stages."""
        disparity_img = empty_tensor(left_image.size(0), left_image.size(1), dtype=torch.int32,
                                     device=left_image.device, reuse_tensor=disparity_img)
        ok = _CostOps.compute_birchfield_wta(left_image.float(), right_image.float(), disparity_img,
                                             self.max_disparity)
        return disparity_img if ok else None
