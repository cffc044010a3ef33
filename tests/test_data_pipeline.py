"""Input pipeline: idx-ubyte parsing on the CPU; the device-resident dataset with the
gather + ToTensor/Normalize kernel on the GPU, bit-identical to torchvision's CPU transforms."""
import gzip
import os
import struct

import numpy as np
import pytest
import torch

from pytorch_operator_b200 import data as D


def _write_idx(path, arr):
    arr = np.ascontiguousarray(arr, dtype=np.uint8)
    hdr = struct.pack(">HBB", 0, 0x08, arr.ndim) + struct.pack(">" + "I" * arr.ndim, *arr.shape)
    opener = gzip.open if path.endswith(".gz") else open
    with opener(path, "wb") as f:
        f.write(hdr + arr.tobytes())


def _fake_mnist(root, n_train=300, n_test=100, seed=0):
    rng = np.random.default_rng(seed)
    raw = os.path.join(root, "MNIST", "raw")
    os.makedirs(raw)
    sets = {}
    for stem, n, ext in (("train", n_train, ""), ("t10k", n_test, ".gz")):
        img, lab = rng.integers(0, 256, (n, 28, 28), dtype=np.uint8), rng.integers(0, 10, (n,), dtype=np.uint8)
        _write_idx(os.path.join(raw, f"{stem}-images-idx3-ubyte{ext}"), img)
        _write_idx(os.path.join(raw, f"{stem}-labels-idx1-ubyte{ext}"), lab)
        sets[stem] = (img, lab)
    return sets


def test_idx_reader_roundtrip_and_layout_discovery(tmp_path):
    sets = _fake_mnist(str(tmp_path))
    tr, te = D.find_mnist(str(tmp_path), True), D.find_mnist(str(tmp_path), False)
    assert tr and te and te[0].endswith(".gz")
    assert np.array_equal(D.read_idx(tr[0]), sets["train"][0]) and np.array_equal(D.read_idx(tr[1]), sets["train"][1])
    assert np.array_equal(D.read_idx(te[0]), sets["t10k"][0])
    assert D.find_mnist(str(tmp_path / "nowhere"), True) is None
    bad = tmp_path / "bad-idx"
    bad.write_bytes(b"\x00\x00\x0b\x01" + b"\x00" * 8)
    with pytest.raises(ValueError):
        D.read_idx(str(bad))


@pytest.mark.gpu
def test_device_dataset_matches_torchvision_transform_bit_exact(tmp_path):
    sets = _fake_mnist(str(tmp_path))
    img, lab = sets["train"]
    dl = D.DeviceMNIST.from_files(*D.find_mnist(str(tmp_path), True), batch_size=64, device="cuda:0", shuffle=True, seed=3)
    assert len(dl) == 5 and len(dl.dataset) == 300
    # the reference's own transform objects (mnist.py:121-124), applied the way datasets.MNIST does: PIL image in
    from PIL import Image
    from torchvision import transforms
    tf = transforms.Compose([transforms.ToTensor(), transforms.Normalize((0.1307,), (0.3081,))])
    want_all = torch.stack([tf(Image.fromarray(im, mode="L")) for im in img])
    # and its arithmetic spelled out: uint8 -> float32 / 255, then (x - mean) / std
    assert torch.equal(want_all, (torch.from_numpy(img).float().div(255).sub_(D.MNIST_MEAN).div_(D.MNIST_STD)).unsqueeze(1))
    seen = []
    for epoch in range(2):
        order = []
        for data, target in dl:
            assert data.is_cuda and data.shape[1:] == (1, 28, 28) and target.dtype == torch.int64
            d, t = data.cpu(), target.cpu()
            # identify each sample by exact match of its label + pixels
            for i in range(d.shape[0]):
                cand = np.nonzero(lab == int(t[i]))[0]
                hit = [c for c in cand if torch.equal(d[i], want_all[c])]
                assert len(hit) >= 1, "pixels are not bit-identical to ToTensor+Normalize"
                order.append(hit[0])
        assert sorted(order) == list(range(300)), "an epoch must visit every sample exactly once"
        seen.append(order)
    assert seen[0] != seen[1] and seen[0] != list(range(300))  # reshuffled every epoch
    fixed = D.DeviceMNIST(img, lab, batch_size=100, device="cuda:0", shuffle=False)
    data, target = next(iter(fixed))
    assert torch.equal(data.cpu(), want_all[:100]) and torch.equal(target.cpu(), torch.from_numpy(lab[:100].astype(np.int64)))


@pytest.mark.gpu
def test_payload_trains_from_idx_files_on_disk(tmp_path):
    import subprocess
    import sys
    _fake_mnist(str(tmp_path / "data"), n_train=640, n_test=200)
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-m", "pytorch_operator_b200.mnist", "--data-root", str(tmp_path / "data"),
                        "--test-batch-size", "100"], env=dict(os.environ, PYTHONPATH=root, WORLD_SIZE="1"),
                       cwd=str(tmp_path), capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    assert "Train Epoch: 1 [0/640 (0%)]\tloss=" in r.stdout and "\naccuracy=0." in r.stdout
