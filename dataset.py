"""Upstream module name kept: ``from dataset import TrainDataset, NewsDataset, ...``."""
from fedrec_b200.data.dataset import (  # noqa: F401
    NewsDataset, NewsPartDataset, NewsUpdatorDataset, TrainDataset, UserDataset, newsample, npratio, max_his_len)
