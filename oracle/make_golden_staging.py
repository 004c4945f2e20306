"""ORACLE (test infrastructure).  Generates tests/golden/staging_small.pt with torchvision's OWN transforms composed
exactly as the reference composes them (dataset/chalearn_dataset.py:41-46, :82-87), on seeded uint8 frames.  Run in
the authoring container:   python -m oracle.make_golden_staging
Contents: the uint8 clips, the RandomCrop windows torchvision drew, the fp32 batch the reference loader would yield
(with and without augmentation) and the normalised value of each of the 256 byte values."""
import os
import sys

import numpy as np
import torch
from torchvision import transforms

_HERE = os.path.dirname(os.path.abspath(__file__))
_REPO = os.path.dirname(_HERE)

N, T, S, C = 3, 2, 20, 21


def main():
    rng = np.random.default_rng(5)
    frames = rng.integers(0, 256, size=(N, T, S, S, C), dtype=np.uint8)
    frames[0, 0, 0, :16, 0] = np.arange(16) * 17  # a few exact corner values incl. 0 and 255
    preprocess = transforms.Compose([
        transforms.ToTensor(),
        transforms.Normalize(mean=[0.45] * C, std=[0.225] * C),
    ])
    plain = torch.stack([torch.stack([preprocess(frames[n, t]) for t in range(T)]) for n in range(N)])  # N,T,C,H,W
    padding = S // 10
    torch.manual_seed(11)
    aug, windows = [], []
    for n in range(N):
        # the draw RandomCrop.forward makes (pad, then get_params on the padded clip), recorded
        padded = transforms.functional.pad(plain[n], padding, 0, "constant")
        i, j, h, w = transforms.RandomCrop.get_params(padded, (S, S))
        windows.append((i, j))
        aug.append(transforms.functional.crop(padded, i, j, h, w))
    # and once through RandomCrop.forward itself with a replayed generator state, to make sure the recorded windows are
    # what the transform applies
    torch.manual_seed(11)
    aug2 = [transforms.RandomCrop(S, padding)(plain[n]) for n in range(N)]
    assert all(torch.equal(a, b) for a, b in zip(aug, aug2)), "recorded windows differ from RandomCrop.forward"
    ramp = np.arange(256, dtype=np.uint8).reshape(16, 16, 1)
    lut = preprocess(np.repeat(ramp, C, axis=2))[0].reshape(256)
    out = dict(frames=torch.from_numpy(frames), plain=plain, aug=torch.stack(aug), windows=windows, padding=padding,
               lut=lut, torchvision=__import__("torchvision").__version__)
    path = os.path.join(_REPO, "tests", "golden", "staging_small.pt")
    torch.save(out, path)
    print("wrote", path, os.path.getsize(path), "bytes; windows", windows)


if __name__ == "__main__":
    sys.exit(main())
