"""Dataset descriptors (role of ``inception/dataset.py:44-105`` and ``inception/imagenet_data.py:24-57``)."""
from __future__ import annotations

import glob
import os
from typing import List


class Dataset:
    """A named dataset subset living as ``<data_dir>/<subset>-*`` TFRecord shards."""

    def __init__(self, name: str, subset: str, data_dir: str = ""):
        assert subset in self.available_subsets(), self.available_subsets()
        self.name, self.subset, self.data_dir = name, subset, data_dir

    def num_classes(self) -> int:
        raise NotImplementedError

    def num_examples_per_epoch(self) -> int:
        raise NotImplementedError

    def download_message(self) -> str:
        raise NotImplementedError

    def available_subsets(self) -> List[str]:
        return ["train", "validation"]

    def data_files(self) -> List[str]:
        """Shard paths; an empty list means the dataset is absent (the reference prints the download
        message and exits, dataset.py:86-95)."""
        return sorted(glob.glob(os.path.join(self.data_dir, "%s-*" % self.subset)))

    def available(self) -> bool:
        return bool(self.data_dir) and len(self.data_files()) > 0


class ImagenetData(Dataset):
    """ILSVRC-2012: 1000 classes (+ background label 0 => 1001 logits), 1,281,167 train / 50,000 validation."""

    def __init__(self, subset: str, data_dir: str = ""):
        super().__init__("ImageNet", subset, data_dir)

    def num_classes(self) -> int:
        return 1000

    def num_examples_per_epoch(self) -> int:
        return 1281167 if self.subset == "train" else 50000

    def download_message(self) -> str:
        return ("Failed to find any ImageNet %s files under %s.  Build them with "
                "`python -m distributed_tensorflow_models_b200.data.build_imagenet_data` from a raw ImageNet tree "
                "(there is no network on this box; benchmarks use --synthetic)." % (self.subset, self.data_dir))


class FlowersData(Dataset):
    """The 5-class flowers example the reference's ``build_image_data`` tooling targets."""

    def __init__(self, subset: str, data_dir: str = ""):
        super().__init__("Flowers", subset, data_dir)

    def num_classes(self) -> int:
        return 5

    def num_examples_per_epoch(self) -> int:
        return 3170 if self.subset == "train" else 500
