"""CPU: the SRN data path (SRNdataset.py:12-95, sampling.py:28-48,164-182) on a small synthetic SRN-shaped tree: same items,
same draws and bit-identical tensors as the reference's own loader (when /root/reference is present), the pre-tensorised
shard format serves the same items, PNG dump round-trips."""
import importlib.util
import os
import pickle
import random

import numpy as np
import pytest
import torch

from distributed_3d_diffusion_pytorch_b200 import srn_data

REF = "/root/reference/SRNdataset.py"


@pytest.fixture(scope="module")
def srn_tree(tmp_path_factory):
    from PIL import Image
    root = tmp_path_factory.mktemp("srn")
    rng = np.random.default_rng(0)
    table = {}
    for o in range(12):
        oid = f"obj{o:02d}"
        frames = [f"{v:06d}.png" for v in range(5)]
        table[oid] = frames
        for sub in ("rgb", "pose", "intrinsics"):
            os.makedirs(root / "cars" / oid / sub)
        for f in frames:
            Image.fromarray(rng.integers(0, 256, (128, 128, 4), dtype=np.uint8), "RGBA").save(root / "cars" / oid / "rgb" / f)
            pose = np.eye(4)
            pose[:3, :4] = rng.normal(size=(3, 4))
            (root / "cars" / oid / "pose" / (f[:-4] + ".txt")).write_text(" ".join(repr(float(x)) for x in pose.flatten()))
            (root / "cars" / oid / "intrinsics" / (f[:-4] + ".txt")).write_text("131.25 0 64 0 131.25 64 0 0 1")
    with open(root / "cars.pickle", "wb") as fh:
        pickle.dump(table, fh)
    return str(root / "cars"), str(root / "cars.pickle")


def _items(ds, seed, n):
    random.seed(seed)
    return [ds[i % len(ds)] for i in range(n)]


@pytest.mark.parametrize("imgsize", [128, 64])
def test_loader_contract_and_shards(srn_tree, imgsize):
    path, pk = srn_tree
    tr, va = srn_data.dataset("train", path, pk, imgsize), srn_data.dataset("val", path, pk, imgsize)
    assert len(tr) == 10 and len(va) == 2 and not set(tr.ids) & set(va.ids)
    imgs, R, T, K = tr[0]
    assert imgs.shape == (2, 3, imgsize, imgsize) and imgs.dtype == np.float32 and -1 <= imgs.min() and imgs.max() <= 1
    assert R.shape == (2, 3, 3) and T.shape == (2, 3) and K.shape == (3, 3) and R.dtype == np.float64
    assert K[0, 0] == 131.25                                   # not rescaled with the image size (SRNdataset.py:68-79)
    if os.path.exists(REF):                                    # bit-identical to the reference's own loader, same RNG consumption
        spec = importlib.util.spec_from_file_location("ref_srn", REF)
        ref = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(ref)
        rd = ref.dataset("train", path, pk, imgsize)
        assert rd.ids == tr.ids
        for a, b in zip(_items(tr, 3, 6), _items(rd, 3, 6)):
            assert all(np.array_equal(x, y) and x.dtype == y.dtype for x, y in zip(a, b))
        random.seed(4); tr[1]; s1 = random.random()
        random.seed(4); rd[1]; s2 = random.random()
        assert s1 == s2
    # the shard format serves the same items
    out = os.path.join(os.path.dirname(path), f"shards{imgsize}")
    assert srn_data.pack_shards(path, pk, out, imgsize=imgsize, objects_per_shard=5) == 3
    sd = srn_data.ShardDataset("train", out)
    assert sd.ids == tr.ids
    for a, b in zip(_items(tr, 5, 6), _items(sd, 5, 6)):
        assert all(np.array_equal(x, y) for x, y in zip(a, b))
    u8 = torch.randint(0, 256, (4, 3, 8, 8), dtype=torch.uint8)
    assert np.array_equal(srn_data.decode_pixels(u8).numpy(), (u8.numpy() / 255 * 2 - 1).astype(np.float32))


def test_multi_epochs_loader_and_batch_contract(srn_tree):
    path, pk = srn_tree
    ds = srn_data.dataset("train", path, pk, 64)
    dl = srn_data.MultiEpochsDataLoader(ds, batch_size=4, shuffle=True, drop_last=True, num_workers=0)
    assert len(dl) == 2
    for epoch in range(3):                                     # one iterator serves every epoch
        batches = list(dl)
        assert len(batches) == 2
        img, R, T, K = batches[0]
        assert img.shape == (4, 2, 3, 64, 64) and img.dtype == torch.float32 and R.shape == (4, 2, 3, 3) and R.dtype == torch.float64
        assert T.shape == (4, 2, 3) and K.shape == (4, 3, 3)


def test_object_views_and_png_dump(srn_tree, tmp_path):
    path, _ = srn_tree
    imgs, Rs, Ts, K = srn_data.load_object_views(os.path.join(path, "obj03"), imgsize=64)
    assert imgs.shape == (5, 3, 64, 64) and Rs.shape == (5, 3, 3) and Ts.shape == (5, 3) and K.dtype == torch.float64
    d = srn_data.save_view_pngs(str(tmp_path / "sampling"), 1, imgs[1], imgs[:3], w=torch.tensor([0, 1, 2]))
    assert sorted(os.listdir(d)) == ["0.png", "1.png", "2.png", "gt.png"]
    from PIL import Image
    back = np.array(Image.open(os.path.join(d, "gt.png")))
    assert np.array_equal(back, ((imgs[1].transpose(1, 2, 0) + 1) * 127.5).astype(np.uint8))
