"""CUDA textures as image inputs (API of reference stereomatch/cuda_texture.py:8-40).

CUDA texture objects are compute API: the native class wraps a cudaArray + cudaTextureObject_t
(csrc/texture.cu behind stm_texture_* of include/stm_abi.h)."""
import torch

from ._cstereomatch import CUDATexture as _CCUDATexture


class CUDATexture(_CCUDATexture):
    """A CUDA texture holding an [H, W] or [H, W, C] image (C in 1, 2, 4; float32 or uint8)."""

    def __del__(self):
        try:
            self.release()
        except Exception:   # interpreter shutdown / context already gone
            pass

    @classmethod
    def from_tensor(cls, tensor: torch.Tensor, normalized_coords: bool = False):
        """tensor: [H, W] or [H, W, C] image on the host or on a CUDA device;
        normalized_coords: texture reads use coordinates in [0, 1) (power-of-two sizes only)."""
        sizes = tensor.size()
        texture = cls()
        texture.copy_from_tensor(tensor.reshape(sizes[0], sizes[1], -1), normalized_coords)
        return texture

    def __str__(self):
        return (f"CUDATexture(texture_object={self.cuda_texture_object}, "
                f"width={self.width}, height={self.height}, channels={self.channels})")

    def __repr__(self):
        return str(self)
