"""CPU: the oracle's restatement of the DataLoader transforms (reference src/mngrdata.py:139-165)
against hand-computed values, and the host-side draw order of PrefetchLoader's augmentation."""
import numpy as np
import torch

from oracle import ops as O

MEAN = (0.47889522, 0.47227842, 0.43047404)
STD = (0.24205776, 0.23828046, 0.25874835)


def test_to_tensor_normalize_known_values():
    img = np.zeros((1, 2, 2, 3), dtype=np.uint8)
    img[0, 0, 0] = (0, 128, 255)
    img[0, 1, 1] = (255, 1, 17)
    out = O.to_tensor_normalize(img, MEAN, STD)
    assert out.shape == (1, 3, 2, 2) and out.dtype == torch.float32
    f = np.float32
    for (h, w, c) in [(0, 0, 0), (0, 0, 1), (0, 0, 2), (1, 1, 0), (1, 1, 1), (1, 1, 2), (0, 1, 0)]:
        want = (f(img[0, h, w, c]) / f(255) - f(MEAN[c])) / f(STD[c])
        assert out[0, c, h, w].item() == float(want)


def test_random_crop_flip_indexing():
    img = np.arange(2 * 4 * 4 * 1, dtype=np.uint8).reshape(2, 4, 4, 1)
    aug = np.array([[0, 0, 0], [2, 1, 1]], dtype=np.int32)
    out = O.random_crop_flip(img, 4, 1, aug)
    # image 0: top-left corner of the padded image -> first row / column are the zero border
    assert (out[0, 0] == 0).all() and (out[0, :, 0] == 0).all()
    assert out[0, 1, 1, 0] == img[0, 0, 0, 0] and out[0, 3, 3, 0] == img[0, 2, 2, 0]
    # image 1: rows 1..4 of the image (last one is border), columns 0..3, then mirrored
    assert (out[1, 3] == 0).all()
    assert out[1, 0, 3, 0] == img[1, 1, 0, 0] and out[1, 0, 0, 0] == img[1, 1, 3, 0]


def test_augmentation_draw_order_matches_torchvision_get_params():
    """RandomCrop.get_params draws top then left with torch.randint, RandomHorizontalFlip one
    torch.rand(1): per image, in that order (the transforms run per sample)."""
    from convnets_b200.loader import PrefetchLoader
    ld = PrefetchLoader.__new__(PrefetchLoader)
    ld.crop, ld.padding, ld.flip_p = 32, 4, 0.5
    ld.generator = torch.Generator().manual_seed(11)
    aug = ld.draw_augmentation(6, 32, 32)
    g = torch.Generator().manual_seed(11)
    for i in range(6):
        top = int(torch.randint(0, 9, (1,), generator=g))
        left = int(torch.randint(0, 9, (1,), generator=g))
        flip = int(float(torch.rand(1, generator=g)) < 0.5)
        assert aug[i].tolist() == [top, left, flip]
