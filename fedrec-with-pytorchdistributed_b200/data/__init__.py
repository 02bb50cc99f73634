from .dataset import (  # noqa: F401
    NewsDataset, NewsPartDataset, NewsUpdatorDataset, TrainDataset, UserDataset,
    ImpressionTable, BatchSampler, DevicePrefetcher, newsample, npratio, max_his_len, load_user_data,
)
from .synthetic import SyntheticMind, make_synthetic_mind  # noqa: F401
