"""ImageNet-in-one-file dataset.  The reference reads ``imagenet-shuffled.hdf5`` through a ``DatasetHDF5`` class
whose import is commented out and whose source is not in its tree (dl_trainer.py:29, 305-336); the interface
is rebuilt from that use and from the writer (scripts/create_hdf5.py:66-72): datasets ``train_img`` /
``val_img`` uint8 [N,224,224,3] and ``train_labels`` / ``val_labels`` int16.  Needs h5py (optional dependency)."""
from torch.utils.data import Dataset


class DatasetHDF5(Dataset):
    def __init__(self, hdf5fn, split, transform=None, target_transform=None):
        if split not in ('train', 'val'):
            raise ValueError("split must be 'train' or 'val'")
        self.hdf5fn, self.split = hdf5fn, split
        self.transform, self.target_transform = transform, target_transform
        self._file = None
        import h5py
        with h5py.File(hdf5fn, 'r') as f:
            self.n = f[split + '_img'].shape[0]

    def _open(self):
        if self._file is None:          # one handle per worker process
            import h5py
            self._file = h5py.File(self.hdf5fn, 'r', libver='latest', swmr=True)
        return self._file

    def __getitem__(self, index):
        f = self._open()
        img = f[self.split + '_img'][index]
        target = int(f[self.split + '_labels'][index])
        if self.transform is not None:
            img = self.transform(img)
        if self.target_transform is not None:
            target = self.target_transform(target)
        return img, target

    def __len__(self):
        return self.n
