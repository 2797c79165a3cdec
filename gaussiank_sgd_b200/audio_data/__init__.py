from .data_loader import (AudioDataLoader, BucketingSampler, DistributedBucketingSampler,  # noqa: F401
                          SpectrogramDataset, SpectrogramParser, SyntheticSpeechDataset)
