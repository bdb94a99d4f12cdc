"""numpy restatement of the reference's input assembly.  TEST INFRASTRUCTURE ONLY.

Follows data_loaders.py:69-77 (normalize_depth), :79-82 (normalize_img) and :158-177
(prepare_model_input) statement by statement; `ToTensorV2` (albumentations, not installed) is
an HWC -> CHW transpose.  Parity: these three functions cannot be imported from the reference
here (data_loaders.py imports trimesh / albumentations / pyrender), so they are restated; they
are ~30 lines of numpy with no third-party arithmetic inside.
"""
import numpy as np
import torch


def normalize_depth(depth_img):
    # data_loaders.py:70-71
    if np.max(depth_img) <= 0.1:
        return depth_img
    sel = depth_img > 0.01
    # data_loaders.py:72-73
    mu = np.mean(depth_img[sel])
    sd = np.std(depth_img[sel])
    # data_loaders.py:74-75
    if np.isclose(sd, 0.0):
        return depth_img
    # data_loaders.py:76-77
    return np.where(sel, (depth_img - mu) / sd, 0.0).astype(np.float32)


def normalize_img(img):
    # data_loaders.py:80-82
    sd = np.std(img)
    mu = np.mean(img)
    return (img - mu) / sd


def _to_chw_tensor(img):
    return torch.from_numpy(np.ascontiguousarray(np.transpose(img, (2, 0, 1))))


def prepare_model_input(init_imgs, gt_imgs, depths, use_norm_depth=False):
    batch = []
    for i in range(len(init_imgs)):
        gt = normalize_img(gt_imgs[i])                                   # :163 (render is NOT normalised, :165)
        gt_t = _to_chw_tensor(gt.astype(np.float32))                     # :166
        init_t = _to_chw_tensor(np.asarray(init_imgs[i]).astype(np.float32))   # :167
        if not use_norm_depth:
            batch.append(torch.cat([init_t, gt_t]))                      # :169
        else:
            nd = torch.tensor(np.expand_dims(normalize_depth(depths[i]), 0))   # :171-173
            batch.append(torch.cat([nd, init_t, gt_t]))                  # :174
    return torch.stack(batch)                                            # :176
