"""Golden vectors for the data path: the reference's own operations (cv2.resize, torchvision normalize) run in the
build container on seeded inputs.  Restates model/CrowdDataset.py:38-66 line by line with the real cv2 / torchvision
calls (the Dataset class itself needs matplotlib and files on disk).  Output: tests/golden/crowd_data_golden.npz"""
import os

import cv2
import numpy as np
import torch
from torchvision import transforms

HERE = os.path.dirname(os.path.abspath(__file__))


def reference_sample(img_u8, gt_dmap, flip, ds=8):
    img = img_u8 / 255                                              # CrowdDataset.py:38-39
    if len(img.shape) == 2:                                         # :41-43
        img = img[:, :, np.newaxis]
        img = np.concatenate((img, img, img), 2)
    if flip:                                                        # :48-50
        img = img[:, ::-1]
        gt_dmap = gt_dmap[:, ::-1]
    ds_rows = int(img.shape[0] // ds)                               # :53-62
    ds_cols = int(img.shape[1] // ds)
    img = cv2.resize(img, (ds_cols * ds, ds_rows * ds))
    img = img.transpose((2, 0, 1))
    gt_dmap = cv2.resize(gt_dmap, (ds_cols, ds_rows))
    gt_dmap = gt_dmap[np.newaxis, :, :] * ds * ds
    img_tensor = torch.tensor(img, dtype=torch.float)               # :64-67
    img_tensor = transforms.functional.normalize(img_tensor, mean=[0.485, 0.456, 0.406], std=[0.229, 0.224, 0.225])
    gt_dmap_tensor = torch.tensor(gt_dmap, dtype=torch.float)
    return img_tensor.numpy(), gt_dmap_tensor.numpy()


def main():
    rng = np.random.default_rng(2024)
    out = {}
    cases = [("rgb_odd", 61, 83, 3, 0), ("rgb_flip", 70, 45, 3, 1), ("gray", 40, 57, 1, 0), ("mult8", 64, 96, 3, 1),
             ("rgb_big", 97, 131, 3, 0)]
    for name, h, w, c, flip in cases:
        img = rng.integers(0, 256, (h, w) if c == 1 else (h, w, 3), dtype=np.uint8)
        gt = (rng.random((h, w)) ** 8).astype(np.float32)
        ri, rg = reference_sample(img, np.ascontiguousarray(gt), flip)
        out[f"{name}.img"], out[f"{name}.gt"], out[f"{name}.flip"] = img, gt, np.int32(flip)
        out[f"{name}.out_img"], out[f"{name}.out_gt"] = ri, rg
    np.savez_compressed(os.path.join(HERE, "crowd_data_golden.npz"), **out)
    print("wrote", len(cases), "cases; cv2", cv2.__version__)


if __name__ == "__main__":
    main()
