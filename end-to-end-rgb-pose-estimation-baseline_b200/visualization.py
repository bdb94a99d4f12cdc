"""Drop-in for the per-pixel half of the reference's visualization.py: create_silhouette (ref:
visualization.py:19-31) and blend_imgs (:34-37), the only per-pixel host loops left in validation /
visualisation once rendering is on the GPU.  Same signatures, numpy in / numpy out;
the work runs as CUDA kernels (csrc/aux_kernels.cu), there is no CPU path.  Figure assembly and file
output (matplotlib / cv2, :43-201) stay the reference's."""
import numpy as np

import rac


def create_silhouette(img, silhouette_only=False):
    """ring = dilate5x5(mean(img,2) > 0) - mask.  silhouette_only: -> (H,W) mask of 0/1; otherwise the ring is painted
    red INTO `img` (the reference assigns in place, visualization.py:30) and `img` is returned."""
    if silhouette_only:
        return rac.silhouette_device(img, True).cpu().numpy().astype(np.float64)
    painted = rac.silhouette_device(img, False).cpu().numpy()
    if isinstance(img, np.ndarray):
        img[...] = painted.astype(img.dtype, copy=False)
        return img
    return painted      # LazyArray / tensor input: nothing to assign into


def blend_imgs(im1, im2, alpha):
    """np.asarray(Image.blend(uint8(im1*255), uint8(im2*255), alpha)) / 255.0 -> f64 ndarray."""
    return rac.blend_device(im1, im2, alpha).cpu().numpy().astype(np.float64)
