"""SRN (ShapeNet cars/chairs) data path: the callers on either side of the hot path.

* `dataset(split, path, picklefile, imgsize)` -- the training-pair loader of the reference (SRNdataset.py:42-95): same
  constructor, same 90/10 split from `random.seed(0); random.shuffle`, same per-item `random.sample(frames, k=2)`, same
  return contract `(imgs f32 [2,3,S,S] in [-1,1], R f64 [2,3,3], T f64 [2,3], K f64 [3,3])` (K is NOT rescaled with the
  image size, SRNdataset.py:68-79).
* `MultiEpochsDataLoader` -- a DataLoader whose worker pool survives epoch boundaries (SRNdataset.py:12-40).
* `load_object_views` / `save_view_pngs` -- the sampler script's input and output (sampling.py:28-48,164-182).
* `pack_shards` / `ShardDataset` -- a pre-tensorised form of the same data (uint8 pixels + fp64 cameras per object in one
  file) so that a training job reads a few large files instead of ~2.5 M small ones; pixels are decoded with the same
  `u8 / 255 * 2 - 1` (float64, then float32) values, on whichever device the batch lives.

Nothing here is on the GPU hot path; the benchmarks use `synthetic.make_views`, which has this loader's output contract."""
import os
import pickle
import random

import numpy as np
import torch
from torch.utils.data import DataLoader, Dataset

# the 256 values `np.array(img) / 255 * 2 - 1` can take, computed like the reference does (float64) and rounded once
_PIXEL_LUT = (np.arange(256, dtype=np.float64) / 255 * 2 - 1).astype(np.float32)


def _read_matrix(filename, shape):
    with open(filename) as fh:
        return np.array(fh.read().strip().split()).astype(float).reshape(shape)


def _read_rgb_u8(filename, imgsize, native=128):
    """uint8 [3,S,S] (alpha dropped); resized with PIL's default filter when S differs from the stored size, like
    SRNdataset.py:75-77 / sampling.py:33-34"""
    from PIL import Image
    img = Image.open(filename)
    if native is None or imgsize != native:
        img = img.resize((imgsize, imgsize))
    a = np.array(img)
    assert a.dtype == np.uint8, "SRN renders are 8-bit PNGs"
    return np.ascontiguousarray(a.transpose(2, 0, 1)[:3])


def decode_pixels(u8):
    """uint8 -> float32 in [-1,1] with the reference's values; numpy array or torch tensor (any device)"""
    if torch.is_tensor(u8):
        return torch.from_numpy(_PIXEL_LUT).to(u8.device)[u8.long()]
    return _PIXEL_LUT[u8]


def _split_ids(all_ids, split):
    ids = sorted(all_ids)
    random.seed(0)                      # the reference reseeds the GLOBAL generator here (SRNdataset.py:51-52); so do we
    random.shuffle(ids)
    cut = int(len(ids) * 0.9)
    return ids[:cut] if split == 'train' else ids[cut:]


class dataset(Dataset):
    """SRNdataset.py:42-95."""

    def __init__(self, split, path='./data/SRN/cars_train', picklefile='./data/cars.pickle', imgsize=128):
        super().__init__()
        self.imgsize = imgsize
        self.path = path
        with open(picklefile, 'rb') as fh:
            self.picklefile = pickle.load(fh)
        self.ids = _split_ids(self.picklefile.keys(), split)

    def __len__(self):
        return len(self.ids)

    def __getitem__(self, idx):
        item = self.ids[idx]
        frames = self.picklefile[item]
        K = _read_matrix(os.path.join(self.path, item, 'intrinsics', frames[0][:-4] + ".txt"), (3, 3))
        pair = random.sample(frames, k=2)
        imgs = np.stack([decode_pixels(_read_rgb_u8(os.path.join(self.path, item, 'rgb', f), self.imgsize)) for f in pair], 0)
        poses = np.stack([_read_matrix(os.path.join(self.path, item, 'pose', f[:-4] + ".txt"), (4, 4)) for f in pair], 0)
        return imgs, poses[:, :3, :3], poses[:, :3, 3], K


class _Forever:
    def __init__(self, sampler):
        self.sampler = sampler

    def __iter__(self):
        while True:
            yield from iter(self.sampler)


class MultiEpochsDataLoader(DataLoader):
    """SRNdataset.py:12-40: one iterator (and one set of workers) for the whole run; `len()` batches per epoch."""

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        object.__setattr__(self, "batch_sampler", _Forever(self.batch_sampler))     # (DataLoader forbids re-assignment after init)
        self.iterator = super().__iter__()

    def __len__(self):
        return len(self.batch_sampler.sampler)

    def __iter__(self):
        for _ in range(len(self)):
            yield next(self.iterator)


def load_object_views(target_dir, imgsize=64):
    """sampling.py:28-48: every view of one object directory, sorted by file name.  Returns (imgs [V,3,S,S] f32, Rs [V,3,3],
    Ts [V,3], K [3,3] torch f64).  K is read from the LAST view's intrinsics file, as the reference does; the
    images are always resized (sampling.py:33)."""
    import glob
    names = sorted(glob.glob(os.path.join(target_dir, "rgb", "*.png")))
    assert names, f"no views under {target_dir}/rgb"
    imgs, Rs, Ts = [], [], []
    for fn in names:
        imgs.append(decode_pixels(_read_rgb_u8(fn, imgsize, native=None)))
        pose = _read_matrix(os.path.join(target_dir, 'pose', os.path.basename(fn)[:-4] + ".txt"), (4, 4))
        Rs.append(pose[:3, :3])
        Ts.append(pose[:3, 3])
    K = torch.tensor(_read_matrix(os.path.join(target_dir, 'intrinsics', os.path.basename(names[-1])[:-4] + ".txt"), (3, 3)))
    return np.stack(imgs, 0), np.stack(Rs, 0), np.stack(Ts, 0), K


def image_to_u8(img):
    """[3,H,W] in [-1,1] -> uint8 [H,W,3] exactly as sampling.py:166,180: ((x + 1) * 127.5).astype(uint8)"""
    return ((np.asarray(img).transpose(1, 2, 0) + 1) * 127.5).astype(np.uint8)


def save_view_pngs(out_root, step, gt, imgs, w=None):
    """sampling.py:164-182: `<out_root>/<step>/gt.png` and one `<i>.png` per guidance weight"""
    from PIL import Image
    d = os.path.join(out_root, str(step))
    os.makedirs(d, exist_ok=True)
    if gt is not None:
        Image.fromarray(image_to_u8(gt)).save(os.path.join(d, "gt.png"))
    for i in (range(len(imgs)) if w is None else [int(v) for v in w]):
        Image.fromarray(image_to_u8(imgs[i])).save(os.path.join(d, f"{i}.png"))
    return d


# ------------------------------------------------------------------ pre-tensorised shards
def pack_shards(path, picklefile, out_dir, imgsize=64, objects_per_shard=64):
    """Reads the SRN tree once and writes `<out_dir>/shard_%05d.pt`, each {ids, frames, rgb uint8 [n,V,3,S,S], pose f64
    [n,V,4,4], K f64 [n,3,3]} for `objects_per_shard` objects with the same number of views, plus `index.pt`."""
    with open(picklefile, 'rb') as fh:
        table = pickle.load(fh)
    os.makedirs(out_dir, exist_ok=True)
    by_views = {}
    for item in sorted(table.keys()):
        by_views.setdefault(len(table[item]), []).append(item)
    index, n_shard = [], 0
    for V, items in sorted(by_views.items()):
        for s in range(0, len(items), objects_per_shard):
            chunk = items[s:s + objects_per_shard]
            rgb = np.stack([np.stack([_read_rgb_u8(os.path.join(path, it, 'rgb', f), imgsize) for f in table[it]], 0) for it in chunk], 0)
            pose = np.stack([np.stack([_read_matrix(os.path.join(path, it, 'pose', f[:-4] + ".txt"), (4, 4)) for f in table[it]], 0) for it in chunk], 0)
            K = np.stack([_read_matrix(os.path.join(path, it, 'intrinsics', table[it][0][:-4] + ".txt"), (3, 3)) for it in chunk], 0)
            name = f"shard_{n_shard:05d}.pt"
            torch.save({"ids": chunk, "frames": [table[it] for it in chunk], "rgb": torch.from_numpy(rgb), "pose": torch.from_numpy(pose),
                        "K": torch.from_numpy(K), "imgsize": imgsize}, os.path.join(out_dir, name))
            index += [(it, name, j) for j, it in enumerate(chunk)]
            n_shard += 1
    torch.save({"index": index, "imgsize": imgsize}, os.path.join(out_dir, "index.pt"))
    return n_shard


class ShardDataset(Dataset):
    """The same items as `dataset` (same split, same `random.sample` draw per item) served from `pack_shards` files."""

    def __init__(self, split, shard_dir, cache_shards=4):
        super().__init__()
        meta = torch.load(os.path.join(shard_dir, "index.pt"), weights_only=False)
        self.dir, self.imgsize = shard_dir, meta["imgsize"]
        self.where = {it: (name, j) for it, name, j in meta["index"]}
        self.ids = _split_ids(self.where.keys(), split)
        self._cache, self._cache_cap = {}, cache_shards

    def _shard(self, name):
        if name not in self._cache:
            if len(self._cache) >= self._cache_cap:
                self._cache.pop(next(iter(self._cache)))
            self._cache[name] = torch.load(os.path.join(self.dir, name), weights_only=False)
        return self._cache[name]

    def __len__(self):
        return len(self.ids)

    def __getitem__(self, idx):
        item = self.ids[idx]
        name, j = self.where[item]
        sh = self._shard(name)
        frames = sh["frames"][j]
        pair = [frames.index(f) for f in random.sample(frames, k=2)]
        poses = sh["pose"][j][pair].numpy()
        return decode_pixels(sh["rgb"][j][pair].numpy()), poses[:, :3, :3], poses[:, :3, 3], sh["K"][j].numpy()
